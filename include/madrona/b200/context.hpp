// Context / CustomContext of the reference (include/madrona/context.hpp:17-159, custom_context.hpp:15-41)
// as a device-side view of one world: the object a system function receives. Structural calls have the
// reference's call-and-return semantics (makeEntityNow returns the Entity; any number per invocation):
// see mdr::ecs::Context::waitTurn in ecs_device.cuh for how row order is kept on 32 lanes.
#pragma once
#include "ecs.hpp"

namespace madrona {

// ResultRef<T> (include/madrona/utils.hpp): valid() + value()
template <typename T> class ResultRef {
public:
    __device__ ResultRef(T *p) : ptr_(p) {}
    __device__ bool valid() const { return ptr_ != nullptr; }
    __device__ T &value() { return *ptr_; }
private:
    T *ptr_;
};

class Context {
public:
    __device__ explicit Context(mdr::ecs::Context &low) : low_(low) {}
    Context(const Context &) = delete;

    __device__ Loc getLoc(Entity e) const { return low_.getLoc(e); }

    // makeEntityNow<ArchetypeT>(components in archetype order...) or with no arguments
    // (components zero-initialised); include/madrona/state.inl:398-449
    template <typename ArchetypeT, typename... Args> __device__ Entity makeEntityNow(Args &&...args)
    {
        constexpr int words = (0 + ... + (int)(sizeof(std::remove_reference_t<Args>) / 4));
        uint32_t payload[words > 0 ? words : 1];
        int at = 0;
        auto put = [&](const auto &c) {
            static_assert(b200::ComponentCheck<std::remove_cv_t<std::remove_reference_t<decltype(c)>>>::ok);
            const uint32_t *src = reinterpret_cast<const uint32_t *>(&c);
            for (int i = 0; i < (int)(sizeof(c) / 4); i++) payload[at + i] = src[i];
            at += (int)(sizeof(c) / 4);
        };
        (put(args), ...);
        return low_.makeEntityNowImmediate(b200::g_type_id<ArchetypeT>, payload, words);
    }
    __device__ void destroyEntityNow(Entity e) { low_.destroyEntityNowImmediate(e); }

    template <typename ArchetypeT> __device__ Loc makeTemporary() { return low_.makeTemporary(b200::g_type_id<ArchetypeT>); }

    template <typename ComponentT> __device__ ResultRef<ComponentT> get(Entity e)
    {
        return ResultRef<ComponentT>(low_.get<ComponentT>(low_.getLoc(e), (uint32_t)b200::deviceTypeId<ComponentT>()));
    }
    template <typename ComponentT> __device__ ComponentT &getUnsafe(Entity e) { return getUnsafe<ComponentT>(low_.getLoc(e)); }
    template <typename ComponentT> __device__ ComponentT &getUnsafe(Loc l)
    {
        return *low_.get<ComponentT>(l, (uint32_t)b200::deviceTypeId<ComponentT>());
    }
    template <typename ComponentT> __device__ ComponentT &getDirect(int32_t column_idx, Loc loc)
    {
        return low_.getDirect<ComponentT>(column_idx, loc);
    }
    template <typename SingletonT> __device__ SingletonT &getSingleton()
    {
        return low_.getSingleton<SingletonT>(b200::g_singleton_arch<SingletonT>);
    }
    template <typename ArchetypeT> __device__ CountT numRows() const { return low_.numRows(b200::g_type_id<ArchetypeT>); }

    __device__ void *tmpAlloc(uint64_t num_bytes) { return low_.tmpAlloc(num_bytes); }
    __device__ WorldID worldID() const { return WorldID { (int32_t)low_.world_id }; }
    __device__ WorldBase &data() { return *(WorldBase *)low_.worldData(); }
    __device__ mdr::ecs::Context &low() { return low_; }

protected:
    mdr::ecs::Context &low_;
};

// class Engine : public madrona::CustomContext<Engine, MyWorld> { using CustomContext::CustomContext; };
template <typename ContextT, typename DataT> class CustomContext : public Context {
public:
    using Context::Context;
    __device__ DataT &data() { return *(DataT *)low_.worldData(); }
};

}  // namespace madrona
