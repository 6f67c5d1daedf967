// ECSRegistry of the reference (include/madrona/state.hpp:109-134, state.inl:17-52) over the C ABI
// (mdr_ecs_register_*): same member templates, so a world's registerTypes() compiles unchanged.
// Tables here have a fixed per-world row capacity (SoA tiles in HBM): registerArchetype takes it as
// an optional argument, default ECSRegistry::defaultRowsPerWorld.
#pragma once
#include "ecs.hpp"

namespace madrona {

class ECSRegistry {
public:
    ECSRegistry(mdr_ecs *ecs, int32_t default_rows_per_world) : ecs_(ecs), defaultRowsPerWorld(default_rows_per_world) {}

    template <typename ComponentT> void registerComponent()
    {
        b200::checkComponent<ComponentT>();
        b200::publishTypeId<ComponentT>(
            b200::req(mdr_ecs_register_component(ecs_, (int32_t)sizeof(ComponentT), (int32_t)alignof(ComponentT)),
                      "registerComponent"));
    }
    template <typename ArchetypeT> void registerArchetype(CountT max_rows_per_world = 0)
    {
        registerArchetypeImpl<ArchetypeT>((b200::ArchetypeBaseOf<ArchetypeT> *)nullptr,
                                          max_rows_per_world > 0 ? max_rows_per_world : defaultRowsPerWorld, 0);
    }
    template <typename ArchetypeT> void registerFixedSizeArchetype(CountT max_num_per_world)
    {
        registerArchetypeImpl<ArchetypeT>((b200::ArchetypeBaseOf<ArchetypeT> *)nullptr, max_num_per_world, MDR_ARCH_FIXED);
    }
    template <typename SingletonT> void registerSingleton()
    {
        b200::checkComponent<SingletonT>();
        const int32_t arch = b200::req(mdr_ecs_register_singleton(ecs_, (int32_t)sizeof(SingletonT), (int32_t)alignof(SingletonT)),
                                       "registerSingleton");
        // the singleton's component id is the one registered just now: the highest so far
        b200::publishTypeId<SingletonT>(mdr_ecs_num_components(ecs_) - 1);
        singletonArch<SingletonT>() = arch;
        b200::reqCuda(cudaMemcpyToSymbol(b200::g_singleton_arch<SingletonT>, &arch, sizeof(arch)), "singleton archetype id");
    }
    template <typename ArchetypeT, typename ComponentT> void exportColumn(int32_t slot)
    {
        b200::req(mdr_ecs_export_column(ecs_, b200::TypeInfo<ArchetypeT>::id, b200::TypeInfo<ComponentT>::id, slot), "exportColumn");
    }
    template <typename SingletonT> void exportSingleton(int32_t slot)
    {
        b200::req(mdr_ecs_export_column(ecs_, singletonArch<SingletonT>(), b200::TypeInfo<SingletonT>::id, slot), "exportSingleton");
    }

    template <typename SingletonT> static int32_t &singletonArch() { static int32_t arch = -1; return arch; }
    mdr_ecs *handle() const { return ecs_; }
    int32_t defaultRowsPerWorld;

private:
    template <typename ArchetypeT, typename... Cs>
    void registerArchetypeImpl(Archetype<Cs...> *, CountT rows, int32_t flags)
    {
        const int32_t comps[] = { b200::TypeInfo<Cs>::id... };
        for (int32_t c : comps) if (c <= 0) b200::fatal("registerArchetype: component not registered", c);
        b200::publishTypeId<ArchetypeT>(
            b200::req(mdr_ecs_register_archetype(ecs_, comps, (int32_t)sizeof...(Cs), rows, flags), "registerArchetype"));
    }
    mdr_ecs *ecs_;
};

}  // namespace madrona
