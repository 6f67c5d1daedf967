// Source-compatible ECS vocabulary for worlds written against the reference (include/madrona/ecs.hpp,
// components.hpp): Entity, Loc, Archetype<Cs...>, WorldBase, WorldID -- over the B200 device tables
// (ecs_device.cuh). Type ids are assigned at registration, in registration order, like the reference's
// TypeTracker (src/common/type_tracker.cpp:87-168) -- kept per type on the host and, for device code,
// in a __device__ variable of the translation unit that registered the type.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <type_traits>
#include <utility>

#include "../../madrona_b200.h"
#include "ecs_device.cuh"

namespace madrona {

using Entity = mdr::ecs::Entity;
using Loc = mdr::ecs::Loc;
using CountT = int32_t;

struct WorldID { int32_t idx; };
struct WorldBase {};

// struct Sphere : madrona::Archetype<Position, Rotation> {};
template <typename... ComponentTs> struct Archetype {};

namespace b200 {

[[noreturn]] inline void fatal(const char *what, int32_t rc)
{
    // the reference's FATAL(): print and abort (include/madrona/crash.hpp)
    fprintf(stderr, "madrona_b200: %s failed (%d)\n", what, rc);
    abort();
}
inline int32_t req(int32_t rc, const char *what) { if (rc < 0) fatal(what, rc); return rc; }
inline void reqCuda(cudaError_t e, const char *what)
{
    if (e != cudaSuccess) { fprintf(stderr, "madrona_b200: %s: %s\n", what, cudaGetErrorString(e)); abort(); }
}

// component / archetype ids: host side, and the device copy written at registration
template <typename T> struct TypeInfo { static inline int32_t id = -1; };
template <> struct TypeInfo<Entity> { static inline int32_t id = MDR_COMP_ENTITY; };
template <typename T> __device__ int32_t g_type_id = -1;            // component id / archetype id of T
template <typename T> __device__ int32_t g_singleton_arch = -1;     // archetype holding singleton T

template <typename T> __device__ __forceinline__ int32_t deviceTypeId()
{
    if constexpr (std::is_same_v<T, Entity>) return MDR_COMP_ENTITY;
    else return g_type_id<T>;
}

template <typename T> inline void publishTypeId(int32_t id)
{
    TypeInfo<T>::id = id;
    reqCuda(cudaMemcpyToSymbol(g_type_id<T>, &id, sizeof(id)), "cudaMemcpyToSymbol(type id)");
}

// Archetype<Cs...> base of a user archetype struct
template <typename... Cs> Archetype<Cs...> archetypeBase(const Archetype<Cs...> *);
template <typename A> using ArchetypeBaseOf = decltype(archetypeBase((const A *)nullptr));

template <typename T> struct ComponentCheck {
    static_assert(sizeof(T) % 4 == 0, "component sizes are multiples of 4 bytes");
    static_assert(alignof(T) <= 16, "components are at most 16-byte aligned");
    static_assert(std::is_trivially_copyable_v<T>, "components are trivially copyable");
    static constexpr bool ok = true;
};
template <typename T> constexpr void checkComponent() { static_assert(ComponentCheck<T>::ok); }

}  // namespace b200
}  // namespace madrona
