// Archetype column map: component id -> column index through a perfect hash.
//
// Restates StaticIntegerMap<128> / StaticMapHelper (include/madrona/hashmap.hpp:26-58, hashmap.inl:14-74,
// src/common/hashmap.cpp:17-116), which the reference keeps per archetype as `ColumnMap columnLookup`
// (include/madrona/state.hpp:245-293) and consults in makeQuery (src/core/state.cpp:315,351), exportColumn
// (:451) and every `get<T>(Loc)` (state.inl:246,284). Table layout and search are the reference's, so the
// tables come out bit-identical (pinned in tests/test_static_map.py against the compiled reference):
//   idx(key) = ((key ^ key >> shift) * multiplier) >> shift,   shift = 32 - log2(capacity)
//   slots 14 and 15 of BOTH arrays hold shift and multiplier (one cache line serves a lookup);
//   capacity = max(16, nextPow2(n + 2)), doubled while no multiplier among 32 tries
//   (0x7feb352d, then hash-prospector rehash | 1 after every collision) is collision-free.
// The lookup is host + device code: the megakernel resolves query columns on the host at mdr_ecs_build, and
// reaches a component of an entity whose archetype is only known at run time through ColumnMapDev::lookup.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define MDR_SM_HD __host__ __device__ __forceinline__
#else
#define MDR_SM_HD inline
#endif

namespace mdr {

constexpr uint32_t kMapSlots = 128;          // StaticIntegerMap<128>
constexpr uint32_t kMapShiftIdx = 14, kMapConstIdx = 15;
constexpr uint32_t kMapInvalidKey = ~0u;
constexpr uint32_t kMapMaxKeys = kMapSlots - 2;

MDR_SM_HD uint32_t staticMapIndex(uint32_t key, uint32_t shift, uint32_t constant)
{
    uint32_t hash = key;
    hash ^= hash >> shift;
    return (hash * constant) >> shift;
}

struct ColumnMapDev {
    uint32_t keys[kMapSlots];
    uint32_t values[kMapSlots];

    MDR_SM_HD bool exists(uint32_t key) const
    {
        return keys[staticMapIndex(key, keys[kMapShiftIdx], keys[kMapConstIdx])] == key;
    }
    // operator[] of the reference: no check
    MDR_SM_HD uint32_t at(uint32_t key) const
    {
        return values[staticMapIndex(key, values[kMapShiftIdx], values[kMapConstIdx])];
    }
    // Optional<uint32_t> lookup(): -1 when absent
    MDR_SM_HD int32_t lookup(uint32_t key) const
    {
        const uint32_t idx = staticMapIndex(key, keys[kMapShiftIdx], keys[kMapConstIdx]);
        return keys[idx] == key ? (int32_t)values[idx] : -1;
    }
};

inline uint32_t mapRehash(uint32_t x)       // utils::int32Hash (hash-prospector), include/madrona/utils.hpp:108-116
{
    x ^= x >> 16u; x *= 0x7feb352du; x ^= x >> 15u; x *= 0x846ca68bu; x ^= x >> 16u;
    return x;
}

// false: no perfect hash within kMapSlots (the reference FATALs), or a key equal to the invalid marker,
// or more keys than fit. `values` untouched slots stay as the caller left them (the reference leaves them
// uninitialised; callers zero the struct first so tables compare).
inline bool buildColumnMap(ColumnMapDev &m, const uint32_t *keys, const uint32_t *values, uint32_t n)
{
    if (n > kMapMaxKeys) return false;
    uint32_t capacity = n + 2 <= 1 ? 1 : 1u << (32 - __builtin_clz(n + 2 - 1));
    if (capacity < 16u) capacity = 16u;
    uint32_t shift = 32u - (31u - (uint32_t)__builtin_clz(capacity));
    uint32_t mult = 0x7feb352du;
    // Only the slots of the FIRST capacity are marked invalid; slots gained by doubling keep the zero the
    // storage starts with (hashmap.cpp:78-80 runs before the search loop).
    for (uint32_t i = 0; i < capacity; i++) m.keys[i] = kMapInvalidKey;
    bool used[kMapSlots];
    while (true) {
        bool found = false;
        // a failed search leaves the multiplier as it was: every capacity starts from the same constant
        uint32_t cur = mult;
        for (int attempt = 0; attempt < 32 && !found; attempt++) {
            for (uint32_t i = 0; i < capacity; i++) used[i] = false;
            used[kMapShiftIdx] = used[kMapConstIdx] = true;
            found = true;
            for (uint32_t i = 0; i < n; i++) {
                if (keys[i] == kMapInvalidKey) return false;
                const uint32_t idx = staticMapIndex(keys[i], shift, cur);
                if (used[idx]) { cur = mapRehash(cur) | 1u; found = false; break; }
                used[idx] = true;
            }
        }
        if (found) { mult = cur; break; }
        capacity *= 2;
        shift -= 1;
        if (capacity > kMapSlots) return false;
    }
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t idx = staticMapIndex(keys[i], shift, mult);
        m.keys[idx] = keys[i];
        m.values[idx] = values[i];
    }
    m.keys[kMapShiftIdx] = m.values[kMapShiftIdx] = shift;
    m.keys[kMapConstIdx] = m.values[kMapConstIdx] = mult;
    return true;
}

}  // namespace mdr
