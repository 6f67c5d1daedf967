// PUBLIC DEVICE HEADER (include/madrona/b200/): user systems are compiled ahead of time against it
// (nvcc, sm_100a) and the task-graph megakernel is instantiated in the user's translation unit, so
// that systems inline into it -- the reference reaches the same end with NVRTC + LTO at run time
// (src/mw/cuda_exec.cpp:553-818). The library (libmadrona_b200.so) owns the tables and the host side.
//
// Device side of the ECS state layer: SoA archetype tables in HBM, the per-world
// entity id map, and the warp-level structural operations (ordered create /
// destroy, clear, stable sort) that the task-graph megakernel and the script
// kernel share. One WARP owns one world for the duration of a step; everything
// here is called with all 32 lanes converged unless stated otherwise.
//
// Reference semantics restated (paths under /root/reference):
//   id map            include/madrona/impl/id_map_impl.inl:68-331
//   makeEntityNow     include/madrona/state.inl:398-449
//   destroyEntityNow  src/core/state.cpp:181-202 (+ Table::removeRow, src/common/table.cpp:59-71)
//   makeTemporary     include/madrona/state.inl:451-464
//   clear             src/core/state.cpp:565-581
//   sort node (spec)  src/mw/device/sort_archetype.cpp:960-1392 (orphan file)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <utility>

#include "static_map.hpp"

namespace mdr {
namespace ecs {

// Capacities of the device descriptor (the reference: 128 columns per table, include/madrona/table.hpp:24-52;
// the physics graph at 4 substeps is 40 nodes, src/physics/physics.cpp:1142-1205).
constexpr int kMaxArch = 32;
constexpr int kMaxCols = 32;
constexpr int kMaxNodes = 96;
constexpr int kMaxMatch = 8;
constexpr int kMaxQueryComps = 12;
constexpr int kMaxExports = 16;
constexpr int kMaxPayloadWords = 24;     // component bytes of one created row, in 4-byte words
constexpr int32_t kSentinel = (int32_t)0xFFFFFFFF;
constexpr int32_t kIdsPerCache = 64;
constexpr uint32_t kNoArch = 0xFFFFFFFFu;
constexpr unsigned kFull = 0xFFFFFFFFu;

enum { kErrRowOverflow = 1, kErrIdOverflow = 2, kErrHandleOverflow = 4, kErrTmpAllocOverflow = 8,
       kErrPendingOverflow = 16 };
enum { kNodeParallelFor = 1, kNodeClearTmp = 2, kNodeResetTmpAlloc = 3, kNodeSort = 4, kNodePhysics = 5 };

struct Entity { uint32_t gen; int32_t id; };
struct Loc { uint32_t archetype; int32_t row; };

// IDMap::Node (id_map.hpp:36-47): Loc and the free links share the first 8 bytes.
struct IdNode { int32_t a, b; uint32_t gen; };
struct IdCache { int32_t free_head, num_free, overflow_head, num_overflow; };
struct IdState { IdCache init_cache, cache; int32_t global_head, num_ids; };

// One archetype: every column is [num_worlds][stride][bytes], world w's rows are
// the contiguous tile starting at row w * stride (stride is a multiple of 16 so
// every tile is 16-byte aligned for any component size that is a multiple of 4).
struct ArchDev {
    int32_t num_cols, max_rows, stride, flags;
    int32_t *row_count;             // [num_worlds]
    char *col[kMaxCols];
    int32_t col_bytes[kMaxCols];
};
struct MatchDev { int32_t arch; int32_t col[kMaxQueryComps]; };
struct Context;
struct NodeDev;
// A node compiled in the user's translation unit: the whole row loop of one node for one world,
// called by the megakernel with all 32 lanes converged (UserFuncID of the reference,
// src/mw/device/include/madrona/taskgraph.inl:9-28).
typedef void (*NodeFn)(Context &, const NodeDev &);
struct NodeDev {
    int32_t kind, func, num_match, arch, key_col, remap;
    NodeFn fn;                      // kNodeParallelFor with func == 0: device function pointer
    MatchDev match[kMaxMatch];
};
struct SortScratch { uint32_t *key[2]; int32_t *idx[2]; char *staging; int32_t staging_bytes; };
struct ExportDev { int32_t arch, col, bytes, fixed; char *buffer; int64_t *offsets; };

struct EcsDev {
    int32_t num_worlds, first_world, num_arch, num_nodes, id_cap, app, num_exports, max_handles;
    ArchDev arch[kMaxArch];
    NodeDev node[kMaxNodes];
    SortScratch sort[kMaxArch];
    ExportDev exports[kMaxExports];
    IdNode *id_nodes;               // [num_worlds][id_cap]
    IdState *id_state;              // [num_worlds]
    Entity *handles;                // script app: [num_worlds][max_handles]
    int32_t *num_handles;           // [num_worlds]
    int32_t *error;                 // one word of kErr* flags
    ColumnMapDev *col_maps;         // [num_arch]: component id -> column (ColumnMap columnLookup, state.hpp:245-293)
    // Context::tmpAlloc (include/madrona/context.hpp, src/core/state.cpp:83-127): one bump arena per world,
    // reset by ResetTmpAllocNode; tmp_used[w] = bytes handed out since the last reset
    char *tmp_arena;                // [num_worlds][tmp_bytes]
    int32_t *tmp_used;              // [num_worlds]
    int32_t tmp_bytes;
    int32_t world_data_bytes;       // user WorldT objects, Context::data()
    char *world_data;               // [num_worlds][world_data_bytes]
};

// Per-warp shared memory: the world's row counts and id caches live here for the
// whole step (loaded once, written back once), plus the radix-sort histogram.
constexpr int kTileArch = 4, kTileCols = 8;      // column tile addresses kept per warp for the first archetypes / columns
struct WarpShared {
    char *tile[kTileArch * kTileCols];              // this world's tile of (archetype a < 4, column c < 8), filled by loadWorld
    int32_t count[kMaxArch];
    IdState ids;
    int32_t tmp_used;               // bump pointer of the world's tmpAlloc arena
    uint32_t done;                  // lanes of the current slice whose system call has returned (immediate mode)
    int32_t hist[256];
};

__device__ __forceinline__ int laneId() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lowMask(int k) { return k >= 32 ? kFull : ((1u << k) - 1u); }

// ------------------------------------------------------------------ id map --
// Executed by ONE lane; restates id_map_impl.inl line by line.
__host__ __device__ inline Entity idAssign(IdNode *nodes, int32_t *head)   // :70-97
{
    const int32_t id = *head;
    const IdNode n = nodes[id];
    if (n.b == 1) {
        *head = n.a;
    } else {
        nodes[id + 1] = IdNode { n.a, n.b - 1, 0u };
        *head = id + 1;
    }
    return Entity { n.gen, id };
}

__host__ __device__ inline Entity idAcquire(IdNode *nodes, IdState &st, IdCache &c, int32_t id_cap, int32_t *error)   // :68-186
{
    if (c.num_overflow > 0) { c.num_overflow--; return idAssign(nodes, &c.overflow_head); }
    if (c.num_free > 0) { c.num_free--; return idAssign(nodes, &c.free_head); }
    if (st.global_head != kSentinel) {
        const int32_t h = st.global_head;
        st.global_head = nodes[h].b;
        nodes[h].b = 1;
        c.free_head = h;
        c.num_free = kIdsPerCache - 1;
        return idAssign(nodes, &c.free_head);
    }
    const int32_t first = st.num_ids;
    if (first + kIdsPerCache > id_cap) {
#ifdef __CUDA_ARCH__
        atomicOr(error, kErrIdOverflow);
#else
        *error |= kErrIdOverflow;
#endif
        return Entity { 0u, -1 };
    }
    st.num_ids = first + kIdsPerCache;
    nodes[first].gen = 0u;
    nodes[first + 1] = IdNode { kSentinel, kIdsPerCache - 1, 0u };
    c.free_head = first + 1;
    c.num_free = kIdsPerCache - 1;
    return Entity { 0u, first };
}

__device__ inline void idRelease(IdNode *nodes, IdState &st, IdCache &c, int32_t id, uint32_t gen)   // :188-224
{
    if (c.num_free < kIdsPerCache) {
        nodes[id] = IdNode { c.free_head, 1, gen + 1u };
        c.free_head = id;
        c.num_free++;
        return;
    }
    if (c.num_overflow < kIdsPerCache) {
        nodes[id] = IdNode { c.overflow_head, 1, gen + 1u };
        c.overflow_head = id;
        c.num_overflow++;
    } else {
        nodes[id].gen = gen + 1u;
        nodes[id].b = 1;
    }
    if (c.num_overflow == kIdsPerCache) {
        nodes[c.overflow_head].b = st.global_head;
        st.global_head = c.overflow_head;
        c.overflow_head = kSentinel;
        c.num_overflow = 0;
    }
}

// bulkRelease (:226-331) including the double generation bump of the last node of
// every full 64-block. One lane; `keys` is the Entity column of the cleared table.
__device__ inline void idBulkRelease(IdNode *nodes, IdState &st, IdCache &c, const Entity *keys, int32_t n)
{
    if (n <= 0) return;
    int32_t base, rem = 0;
    int32_t tail = kSentinel;
    for (base = 0; base < n; base += kIdsPerCache) {
        rem = n - base;
        if (rem < kIdsPerCache) break;
        const int32_t head_id = keys[base].id;
        for (int s = 0; s < kIdsPerCache - 1; s++) {
            IdNode &nd = nodes[keys[base + s].id];
            nd.gen += 1u; nd.a = keys[base + s + 1].id; nd.b = 1;
        }
        IdNode &last = nodes[keys[base + kIdsPerCache - 1].id];
        last.gen += 2u; last.a = kSentinel; last.b = 1;
        if (tail != kSentinel) nodes[tail].b = head_id;
        tail = head_id;
    }
    if (rem != kIdsPerCache) {
        const int32_t start = keys[base].id;
        for (int32_t i = base; i < n - 1; i++) {
            IdNode &nd = nodes[keys[i].id];
            nd.gen += 1u; nd.a = keys[i + 1].id; nd.b = 1;
        }
        IdNode &t = nodes[keys[n - 1].id];
        t.gen += 1u; t.b = 1; t.a = c.overflow_head;
        const int32_t from_overflow = kIdsPerCache - rem;
        if (c.num_overflow < from_overflow) {
            c.overflow_head = start;
            c.num_overflow += rem;
        } else {
            int32_t next = c.overflow_head, cut = kSentinel;
            for (int32_t i = 0; i < from_overflow; i++) { cut = next; next = nodes[cut].a; }
            nodes[cut].a = kSentinel;
            c.overflow_head = next;
            c.num_overflow -= from_overflow;
            if (tail != kSentinel) nodes[tail].b = start;
            tail = start;
        }
    }
    if (tail == kSentinel) return;
    nodes[tail].b = st.global_head;
    st.global_head = keys[0].id;
}

// ------------------------------------------------------------------ context --
// What a system function sees: the reference's Context (include/madrona/
// context.hpp:17-159) reduced to the calls the task-graph path uses. One object
// per lane. Structural changes requested by a system (makeEntityNow /
// destroyEntityNow) are applied at the convergence point that follows every
// 32-row slice of a ParallelFor, in row order, so that ids, generations and row
// order come out as the reference's sequential loop produces them (a node must
// not structurally change an archetype it iterates -- the reference skips rows
// in that case as well).
struct Context {
    const EcsDev *E;
    const ArchDev *archs = nullptr;   // E->arch, or the block's shared-memory copy of it (BlockDescriptors)
    WarpShared *ws;
    int32_t world;          // index on this GPU
    uint32_t world_id;      // global index (seeds)
    uint32_t pend;          // bit 0: destroy pending, bit 1: create pending, bit 2: nested job pending
    uint32_t nested[4];     // arguments of the pending nested parallel-for
    Entity pend_e;
    int32_t create_arch;
    Entity created;         // result of the applied create (script kernel)
    int32_t created_row;
    uint32_t payload[kMaxPayloadWords];

    __device__ __forceinline__ char *colPtr(int arch, int col) const
    {
        // the address of this world's tile: from the per-warp table (one shared-memory load) for the first
        // archetypes / columns, computed from the descriptor otherwise
        if (arch < kTileArch && col < kTileCols) return ws->tile[arch * kTileCols + col];
        const ArchDev &A = archs[arch];
        return A.col[col] + (size_t)world * A.stride * A.col_bytes[col];
    }
    template <typename T> __device__ __forceinline__ T *column(int arch, int col) const { return (T *)colPtr(arch, col); }
    __device__ __forceinline__ int32_t numRows(int arch) const { return ((volatile int32_t *)ws->count)[arch]; }
    __device__ __forceinline__ IdNode *idNodes() const { return E->id_nodes + (size_t)world * E->id_cap; }

    // Context::getLoc -> EntityStore::getLoc -> IDMap::lookup (id_map.hpp:59-68)
    __device__ __forceinline__ Loc getLoc(Entity e) const
    {
        if (e.id < 0 || e.id >= ((volatile IdState *)&ws->ids)->num_ids) return Loc { kNoArch, 0 };
        const IdNode n = idNodes()[e.id];
        if (n.gen != e.gen) return Loc { kNoArch, 0 };
        return Loc { (uint32_t)n.a, n.b };
    }
    // Context::getDirect (include/madrona/state.inl:292-303)
    template <typename T> __device__ __forceinline__ T &getDirect(int col, Loc l) const { return column<T>((int)l.archetype, col)[l.row]; }
    // Context::get<T>(Loc) (state.inl:240-262): the component's column in the entity's archetype through the
    // archetype's perfect-hash column map; nullptr when the archetype has no such component (ResultRef).
    __device__ __forceinline__ int32_t columnOf(int arch, uint32_t comp_id) const { return E->col_maps[arch].lookup(comp_id); }
    template <typename T> __device__ __forceinline__ T *get(Loc l, uint32_t comp_id) const
    {
        if (l.archetype == kNoArch) return nullptr;
        const int32_t col = columnOf((int)l.archetype, comp_id);
        return col < 0 ? nullptr : &column<T>((int)l.archetype, col)[l.row];
    }
    // Context::getSingleton (state.inl:205-215): column 1, row 0 of the singleton archetype
    template <typename T> __device__ __forceinline__ T &getSingleton(int arch) const { return column<T>(arch, 1)[0]; }

    // Context::makeTemporary: a row without an id. Rows appended by different lanes
    // of one slice land in unspecified order (a sort node restores a defined one).
    __device__ __forceinline__ Loc makeTemporary(int arch) const
    {
        const int32_t row = atomicAdd(&ws->count[arch], 1);
        if (row >= archs[arch].max_rows) {
            atomicSub(&ws->count[arch], 1);
            atomicOr(E->error, kErrRowOverflow);
            return Loc { kNoArch, 0 };
        }
        return Loc { (uint32_t)arch, row };
    }
    template <typename ArchT> __device__ __forceinline__ Loc makeTemporary() const { return makeTemporary(ArchT::kId); }
    template <typename ArchT> __device__ __forceinline__ int32_t numRows() const { return numRows(ArchT::kId); }
    template <typename T> __device__ __forceinline__ T &getSingleton() const { return getSingleton<T>(T::kSingletonArch); }
    // ---- immediate mode: the reference's call-and-return semantics -------------------------
    // makeEntityNow returns the Entity, destroyEntityNow acts before it returns, any number per
    // invocation. Rows of a slice run on 32 lanes at once; a lane's structural call waits until
    // every LOWER lane of the slice has returned from its system call (ws->done), so structural
    // changes happen in row order, one lane at a time: ids, generations and row order come out as
    // the reference's sequential row loop produces them (include/madrona/state.inl:398-449,
    // src/core/state.cpp:181-202). parallelForImmediate() runs a node in this mode; world
    // constructors run on lane 0 alone (serial != 0: no waiting).
    uint32_t serial;
    __device__ __forceinline__ void waitTurn() const
    {
        if (serial) return;
        const unsigned below = lowMask((int)(threadIdx.x & 31));
        while ((((volatile WarpShared *)ws)->done & below) != below) { }
        __threadfence_block();
    }
    __device__ inline Entity makeEntityNowImmediate(int arch, const uint32_t *words, int num_words)
    {
        waitTurn();
        const ArchDev &A = archs[arch];
        volatile int32_t *cnt = ws->count;
        const int32_t r = cnt[arch];
        if (r >= A.max_rows) { atomicOr(E->error, kErrRowOverflow); return Entity { 0u, -1 }; }
        const Entity e = idAcquire(idNodes(), ws->ids, ws->ids.cache, E->id_cap, E->error);
        if (e.id < 0) return e;
        cnt[arch] = r + 1;
        column<Entity>(arch, 0)[r] = e;
        int w = 0;
        for (int c = 1; c < A.num_cols; c++) {
            const int n = A.col_bytes[c] >> 2;
            uint32_t *dst = (uint32_t *)colPtr(arch, c) + (size_t)r * n;
            for (int k = 0; k < n; k++) dst[k] = w + k < num_words ? words[w + k] : 0u;
            w += n;
        }
        IdNode *nodes = idNodes();
        nodes[e.id].a = arch;       // EntityStore::setLoc (state.inl:443-446)
        nodes[e.id].b = r;
        __threadfence_block();
        return e;
    }
    __device__ inline void destroyEntityNowImmediate(Entity e)
    {
        waitTurn();
        const Loc l = getLoc(e);
        if (l.archetype == kNoArch) return;         // stale handle: the reference's lookup fails the same way
        const int arch = (int)l.archetype;
        const ArchDev &A = archs[arch];
        volatile int32_t *cnt = ws->count;
        const int32_t from = cnt[arch] - 1, to = l.row;
        IdNode *nodes = idNodes();
        if (from != to) {                           // Table::removeRow: the last row fills the hole (table.cpp:59-71)
            for (int c = 0; c < A.num_cols; c++) {
                const int n = A.col_bytes[c] >> 2;
                uint32_t *base = (uint32_t *)colPtr(arch, c);
                for (int k = 0; k < n; k++) base[(size_t)to * n + k] = base[(size_t)from * n + k];
            }
            nodes[column<Entity>(arch, 0)[to].id].b = to;   // EntityStore::setRow (state.cpp:196-200)
        }
        cnt[arch] = from;
        idRelease(nodes, ws->ids, ws->ids.cache, e.id, e.gen);
        __threadfence_block();
    }
    // Context::tmpAlloc: bytes from the world's arena in 256-byte grains (TmpAllocator::alloc rounds the
    // same way, src/core/state.cpp:95-115), valid until the next ResetTmpAllocNode; nullptr (and
    // kErrTmpAllocOverflow) when the arena is full.
    __device__ __forceinline__ void *tmpAlloc(uint64_t num_bytes) const
    {
        const int32_t need = (int32_t)((num_bytes + 255u) & ~(uint64_t)255u);
        const int32_t at = atomicAdd(&ws->tmp_used, need);
        if (at + need > E->tmp_bytes) {
            atomicSub(&ws->tmp_used, need);
            atomicOr(E->error, kErrTmpAllocOverflow);
            return nullptr;
        }
        return E->tmp_arena + (size_t)world * E->tmp_bytes + at;
    }
    __device__ __forceinline__ void *worldData() const { return E->world_data + (size_t)world * E->world_data_bytes; }

    // ---- deferred mode (the built-in worlds): destroyEntityNow / makeEntityNow<Archetype>(components...),
    // at most one of each per invocation, applied in row order after the slice (destroy first);
    // a second request in one invocation raises kErrPendingOverflow instead of being dropped.
    __device__ __forceinline__ void destroyEntityNow(Entity e)
    {
        if (pend & 1u) atomicOr(E->error, kErrPendingOverflow);
        pend |= 1u; pend_e = e;
    }
    // Context::parallelFor issued from inside a system (examples/fantasy_vs/fvs.cpp:180-190):
    // the nested loop is handed to the whole warp after the slice (see parallelForNested);
    // at most one per invocation, four words of arguments.
    __device__ __forceinline__ void submitNested(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3)
    {
        pend |= 4u;
        nested[0] = a0; nested[1] = a1; nested[2] = a2; nested[3] = a3;
    }
    template <typename ArchT, typename... Cs> __device__ __forceinline__ void makeEntityNow(const Cs &...cs)
    {
        makeEntityNowIn(ArchT::kId, cs...);
    }
    template <typename... Cs> __device__ __forceinline__ void makeEntityNowIn(int arch, const Cs &...cs)
    {
        if (pend & 2u) atomicOr(E->error, kErrPendingOverflow);
        pend |= 2u;
        create_arch = arch;
        int w = 0;
        auto put = [&](const auto &c) {
            static_assert(sizeof(c) % 4 == 0, "component sizes are multiples of 4 bytes");
            const uint32_t *src = reinterpret_cast<const uint32_t *>(&c);
            for (int i = 0; i < (int)(sizeof(c) / 4); i++) payload[w + i] = src[i];
            w += (int)(sizeof(c) / 4);
        };
        (put(cs), ...);
    }
};

// ------------------------------------------------------- ordered destroys --
// The lanes of `run` each destroy ctx.pend_e; the result equals destroying them
// one after the other in lane order. The sequential part (which row fills which
// hole, id releases) runs on indices only: each lane that caused a move holds one
// (archetype, destination, ORIGINAL source row) entry, composed so that the row
// data itself moves once, after the loop, all movers in parallel.
__device__ inline void destroyBatch(Context &ctx, unsigned run)
{
    const int lane = laneId();
    const EcsDev *E = ctx.E;
    WarpShared *ws = ctx.ws;
    IdNode *nodes = ctx.idNodes();
    const bool mine = (run >> lane) & 1u;

    bool valid = false;
    uint32_t gen = 0;
    int32_t arch = -1, row = 0;
    const Entity e = ctx.pend_e;
    if (mine && e.id >= 0 && e.id < ws->ids.num_ids) {
        const IdNode n = nodes[e.id];
        valid = n.gen == e.gen;
        gen = n.gen; arch = n.a; row = n.b;
    }
    // the same live entity named twice in one run: only the first destroy acts
    if (mine) {
        const unsigned peers = __match_any_sync(run, e.id);
        if ((peers & lowMask(lane)) & __ballot_sync(run, valid)) valid = false;
    }
    unsigned todo = __ballot_sync(kFull, valid);
    int32_t ent_arch = -1, ent_to = 0, ent_orig = 0;
    while (todo) {
        const int i = __ffs(todo) - 1;
        todo &= todo - 1;
        const int32_t a_i = __shfl_sync(kFull, arch, i);
        const int32_t r_i = __shfl_sync(kFull, row, i);
        const int32_t id_i = __shfl_sync(kFull, e.id, i);
        const uint32_t g_i = __shfl_sync(kFull, gen, i);
        // where the entity's row content lives now: the latest move of its original row
        const unsigned hit = __ballot_sync(kFull, ent_arch == a_i && ent_orig == r_i);
        const int32_t to = hit ? __shfl_sync(kFull, ent_to, 31 - __clz(hit)) : r_i;
        const int32_t from = ws->count[a_i] - 1;
        __syncwarp();
        if (lane == 0) {
            ws->count[a_i] = from;
            idRelease(nodes, ws->ids, ws->ids.cache, id_i, g_i);
        }
        if (from != to) {
            const unsigned hit2 = __ballot_sync(kFull, ent_arch == a_i && ent_to == from);
            const int32_t orig = hit2 ? __shfl_sync(kFull, ent_orig, 31 - __clz(hit2)) : from;
            if (lane == i) { ent_arch = a_i; ent_to = to; ent_orig = orig; }
        }
        __syncwarp();
    }
    // Movers: the latest entry per (archetype, destination) whose destination survived.
    const unsigned emask = __ballot_sync(kFull, ent_arch >= 0);
    bool mover = false;
    if (ent_arch >= 0) {
        const unsigned peers = __match_any_sync(emask, (ent_arch << 24) ^ ent_to);
        mover = (peers >> lane) == 1u && ent_to < ws->count[ent_arch];
    }
    if (mover) {
        const ArchDev &A = ctx.archs[ent_arch];
        for (int c = 0; c < A.num_cols; c++) {
            const int words = A.col_bytes[c] >> 2;
            uint32_t *base = (uint32_t *)ctx.colPtr(ent_arch, c);
            const uint32_t *src = base + (size_t)ent_orig * words;
            uint32_t *dst = base + (size_t)ent_to * words;
            for (int k = 0; k < words; k++) dst[k] = src[k];
        }
        // EntityStore::setRow of the moved entity (state.cpp:196-200)
        const Entity moved = ctx.column<Entity>(ent_arch, 0)[ent_orig];
        nodes[moved.id].b = ent_to;
    }
    __syncwarp();
}

// ------------------------------------------------------- ordered creates --
__device__ inline void createBatch(Context &ctx, unsigned run)
{
    const int lane = laneId();
    const EcsDev *E = ctx.E;
    WarpShared *ws = ctx.ws;
    IdNode *nodes = ctx.idNodes();
    Entity my_e { 0u, -1 };
    int32_t my_row = -1;
    unsigned todo = run;
    while (todo) {
        const int i = __ffs(todo) - 1;
        todo &= todo - 1;
        const int32_t a_i = __shfl_sync(kFull, ctx.create_arch, i);
        Entity e { 0u, -1 };
        int32_t r = -1;
        if (lane == 0) {
            r = ws->count[a_i];
            if (r >= ctx.archs[a_i].max_rows) {
                atomicOr(E->error, kErrRowOverflow);
                r = -1;
            } else {
                e = idAcquire(nodes, ws->ids, ws->ids.cache, E->id_cap, E->error);
                if (e.id < 0) r = -1;
                else ws->count[a_i] = r + 1;
            }
        }
        e.gen = __shfl_sync(kFull, e.gen, 0);
        e.id = __shfl_sync(kFull, e.id, 0);
        r = __shfl_sync(kFull, r, 0);
        if (lane == i) { my_e = e; my_row = r; }
    }
    if (((run >> lane) & 1u) && my_row >= 0) {
        const int a = ctx.create_arch;
        const ArchDev &A = ctx.archs[a];
        ctx.column<Entity>(a, 0)[my_row] = my_e;
        int w = 0;
        for (int c = 1; c < A.num_cols; c++) {
            const int words = A.col_bytes[c] >> 2;
            uint32_t *dst = (uint32_t *)ctx.colPtr(a, c) + (size_t)my_row * words;
            for (int k = 0; k < words; k++) dst[k] = w + k < kMaxPayloadWords ? ctx.payload[w + k] : 0u;
            w += words;
        }
        nodes[my_e.id].a = a;      // EntityStore::setLoc (state.inl:443-446)
        nodes[my_e.id].b = my_row;
    }
    ctx.created = my_e;
    ctx.created_row = my_row;
    __syncwarp();
}

// Applies the structural requests of one slice in slot order
// (lane 0 destroy, lane 0 create, lane 1 destroy, ...), batching maximal runs of
// one kind.
__device__ inline void applyPending(Context &ctx)
{
    unsigned rd = __ballot_sync(kFull, (ctx.pend & 1u) != 0u);
    unsigned rc = __ballot_sync(kFull, (ctx.pend & 2u) != 0u);
    ctx.created = Entity { 0u, -1 };
    ctx.created_row = -1;
    while (rd | rc) {
        const int ld = rd ? __ffs(rd) - 1 : 32;
        const int lc = rc ? __ffs(rc) - 1 : 32;
        if (ld <= lc) {
            const unsigned run = rd & lowMask(lc + 1);
            destroyBatch(ctx, run);
            rd &= ~run;
        } else {
            const unsigned run = rc & lowMask(ld);
            const Entity keep = ctx.created;
            const int32_t keep_row = ctx.created_row;
            createBatch(ctx, run);
            if (!((run >> laneId()) & 1u)) { ctx.created = keep; ctx.created_row = keep_row; }
            rc &= ~run;
        }
    }
    ctx.pend &= 4u;
}

// ------------------------------------------------------------------- clear --
// StateManager::clear (src/core/state.cpp:565-581).
__device__ inline void clearArchetype(Context &ctx, int arch, bool is_temporary)
{
    WarpShared *ws = ctx.ws;
    __syncwarp();
    if (laneId() == 0) {
        const int32_t n = ws->count[arch];
        if (!is_temporary && n > 0)
            idBulkRelease(ctx.idNodes(), ws->ids, ws->ids.cache, ctx.column<Entity>(arch, 0), n);
        ws->count[arch] = 0;
    }
    __syncwarp();
}

// -------------------------------------------------------------------- sort --
// Stable sort of one world's rows of `arch` by the u32 column `key_col`:
// LSD radix, 8-bit digits, passes = ceil(bits(max key) / 8)
// (sort_archetype.cpp:960-1073), ranks by warp match instead of CUB's block
// ranker; then every column is gathered through a staging tile and the moved
// entities' rows are patched (:1215-1279).
__device__ inline void sortArchetype(Context &ctx, int arch, int key_col, bool remap)
{
    const int lane = laneId();
    const EcsDev *E = ctx.E;
    WarpShared *ws = ctx.ws;
    const ArchDev &A = ctx.archs[arch];
    const SortScratch &S = E->sort[arch];
    __syncwarp();
    const int32_t n = ws->count[arch];
    if (n <= 1) return;
    const size_t tile = (size_t)ctx.world * A.stride;
    uint32_t *key_a = S.key[0] + tile, *key_b = S.key[1] + tile;
    int32_t *idx_a = S.idx[0] + tile, *idx_b = S.idx[1] + tile;
    const uint32_t *keys = ctx.column<uint32_t>(arch, key_col);
    uint32_t mx = 0;
    bool sorted = true;
    for (int32_t i = lane; i < n; i += 32) {
        const uint32_t k = keys[i];
        key_a[i] = k; idx_a[i] = i;
        mx = max(mx, k);
        if (i + 1 < n && keys[i + 1] < k) sorted = false;
    }
    if (__all_sync(kFull, sorted)) return;
    for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(kFull, mx, o));
    const int bits = mx ? 32 - __clz(mx) : 1;
    const int passes = (bits + 7) / 8;
    __syncwarp();
    for (int p = 0; p < passes; p++) {
        const int shift = 8 * p;
        for (int b = lane; b < 256; b += 32) ws->hist[b] = 0;
        __syncwarp();
        for (int32_t i = lane; i < n; i += 32) atomicAdd(&ws->hist[(key_a[i] >> shift) & 255u], 1);
        __syncwarp();
        // exclusive scan of 256 bins: lane owns bins [8*lane, 8*lane+8)
        int32_t loc[8], sum = 0;
        for (int k = 0; k < 8; k++) { loc[k] = sum; sum += ws->hist[8 * lane + k]; }
        int32_t incl = sum;
        for (int o = 1; o < 32; o <<= 1) {
            const int32_t v = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += v;
        }
        const int32_t excl = incl - sum;
        __syncwarp();
        for (int k = 0; k < 8; k++) ws->hist[8 * lane + k] = excl + loc[k];
        __syncwarp();
        for (int32_t base = 0; base < n; base += 32) {
            const int32_t i = base + lane;
            const bool act = i < n;
            const unsigned amask = __ballot_sync(kFull, act);
            uint32_t k = 0; int32_t src = 0, dst = 0; unsigned peers = 0; uint32_t d = 0;
            if (act) {
                k = key_a[i]; src = idx_a[i];
                d = (k >> shift) & 255u;
                peers = __match_any_sync(amask, d);
                dst = ws->hist[d] + __popc(peers & lowMask(lane));
            }
            __syncwarp();
            if (act && (peers >> lane) == 1u) ws->hist[d] += __popc(peers);
            __syncwarp();
            if (act) { key_b[dst] = k; idx_b[dst] = src; }
        }
        uint32_t *tk = key_a; key_a = key_b; key_b = tk;
        int32_t *ti = idx_a; idx_a = idx_b; idx_b = ti;
        __syncwarp();
    }
    // gather every column through the staging tile
    char *staging = S.staging + (size_t)ctx.world * A.stride * S.staging_bytes;
    for (int c = 0; c < A.num_cols; c++) {
        const int bytes = A.col_bytes[c];
        char *col = ctx.colPtr(arch, c);
        if ((bytes & 15) == 0) {
            const int q = bytes >> 4;
            for (int32_t i = lane; i < n * q; i += 32) {
                const int32_t r = i / q, k = i - r * q;
                ((uint4 *)staging)[i] = ((const uint4 *)col)[(size_t)idx_a[r] * q + k];
            }
            __syncwarp();
            for (int32_t i = lane; i < n * q; i += 32) ((uint4 *)col)[i] = ((const uint4 *)staging)[i];
        } else {
            const int q = bytes >> 2;
            for (int32_t i = lane; i < n * q; i += 32) {
                const int32_t r = i / q, k = i - r * q;
                ((uint32_t *)staging)[i] = ((const uint32_t *)col)[(size_t)idx_a[r] * q + k];
            }
            __syncwarp();
            for (int32_t i = lane; i < n * q; i += 32) ((uint32_t *)col)[i] = ((const uint32_t *)staging)[i];
        }
        __syncwarp();
    }
    if (remap) {
        IdNode *nodes = ctx.idNodes();
        const Entity *ents = ctx.column<Entity>(arch, 0);
        for (int32_t i = lane; i < n; i += 32) nodes[ents[i].id].b = i;
    }
    __syncwarp();
}

// ------------------------------------------------------ world load / store --
// What one walk of the graph starts from besides the world's tables: no pending structural batch, lane turns at
// zero. A warp that walks SEVERAL steps of its world between one loadWorld and one storeWorld (the batched step
// kernels of csrc/ecs.cu) calls this before every walk after the first.
__device__ __forceinline__ void beginStep(Context &ctx)
{
    ctx.pend = 0u;
    ctx.serial = 0u;
    ctx.created = Entity { 0u, -1 };
    ctx.created_row = -1;
    __syncwarp();
    if (laneId() == 0) ctx.ws->done = 0u;
}

__device__ inline void loadWorld(Context &ctx, const EcsDev *E, WarpShared *ws, int world)
{
    const int lane = laneId();
    ctx.E = E; ctx.ws = ws; ctx.world = world; ctx.world_id = (uint32_t)(E->first_world + world);
    if (!ctx.archs) ctx.archs = E->arch;
    beginStep(ctx);
    if (lane == 0) ws->tmp_used = E->tmp_used ? E->tmp_used[world] : 0;
    {
        const int a = lane / kTileCols, c = lane % kTileCols;      // 32 lanes = kTileArch x kTileCols entries
        char *ptr = nullptr;
        if (a < E->num_arch && c < ctx.archs[a].num_cols) {
            const ArchDev &A = ctx.archs[a];
            ptr = A.col[c] + (size_t)world * A.stride * A.col_bytes[c];
        }
        ws->tile[lane] = ptr;
    }
    if (lane < E->num_arch) ws->count[lane] = ctx.archs[lane].row_count[world];
    if (lane < (int)(sizeof(IdState) / 4)) ((int32_t *)&ws->ids)[lane] = ((const int32_t *)(E->id_state + world))[lane];
    __syncwarp();
}

__device__ inline void storeWorld(Context &ctx)
{
    const int lane = laneId();
    const EcsDev *E = ctx.E;
    WarpShared *ws = ctx.ws;
    __syncwarp();
    if (lane == 0 && E->tmp_used) E->tmp_used[ctx.world] = ws->tmp_used;
    if (lane < E->num_arch) ctx.archs[lane].row_count[ctx.world] = ws->count[lane];
    if (lane < (int)(sizeof(IdState) / 4)) ((int32_t *)(E->id_state + ctx.world))[lane] = ((const int32_t *)&ws->ids)[lane];
    __syncwarp();
}

// ParallelForNode<ContextT, Fn, Components...>::run (include/madrona/taskgraph.inl:
// 63-71) -> Context::forEach -> StateManager::iterateEntities (state.inl:358-396):
// matching archetypes in id order, rows in order; lanes take consecutive rows.
// The 32-row slices of one table; the column tile addresses arrive as arguments (computed once per table).
// Asks for the rows of the NEXT slice while this one is worked on (non-binding L1 prefetch): a slice otherwise
// starts with a load nobody has asked for yet, and the warp has nothing else in flight.
template <typename T> __device__ __forceinline__ void prefetchRow(const T *p)
{
#ifndef MDR_ECS_NO_PREFETCH
    asm volatile("prefetch.global.L1 [%0];" ::"l"(p));
#endif
}

template <auto Fn, typename... Cs>
__device__ __forceinline__ void sliceLoop(Context &ctx, int32_t n, int lane, Cs *...cols)
{
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t row = base + lane;
        if (row + 32 < n) (prefetchRow(cols + row + 32), ...);
        if (row < n) Fn(ctx, cols[row]...);
        __syncwarp();
        if (__any_sync(kFull, ctx.pend != 0u)) applyPending(ctx);
    }
}

template <auto Fn, typename... Cs, int... Is>
__device__ __forceinline__ void parallelForImpl(Context &ctx, const NodeDev &nd, std::integer_sequence<int, Is...>)
{
    const int lane = laneId();
    for (int m = 0; m < nd.num_match; m++) {
        const MatchDev &mt = nd.match[m];
        const int32_t n = ctx.numRows(mt.arch);
        // the tile addresses of this world's columns, once per table (not once per row: the descriptors sit in
        // shared memory and the compiler must assume every store of a system may change them)
        sliceLoop<Fn, Cs...>(ctx, n, lane, ctx.column<Cs>(mt.arch, mt.col[Is])...);
        __syncwarp();
    }
}
template <auto Fn, typename... Cs>
__device__ __forceinline__ void parallelFor(Context &ctx, const NodeDev &nd)
{
    parallelForImpl<Fn, Cs...>(ctx, nd, std::make_integer_sequence<int, (int)sizeof...(Cs)>());
}

// A system that may submit one nested parallel-for per invocation (Context::
// submitNested). The nested jobs of a slice run right after it, one at a time in row
// order, each executed by ALL 32 lanes: Nested(ctx, args) is called converged and
// strides its own loop by lane. Valid when the nested job's effects are not read by
// later rows of the same node (the reference runs it inline).
template <auto Fn, auto Nested, typename... Cs>
__device__ __forceinline__ void sliceLoopNested(Context &ctx, int32_t n, int lane, Cs *...cols)
{
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t row = base + lane;
        if (row + 32 < n) (prefetchRow(cols + row + 32), ...);
        if (row < n) Fn(ctx, cols[row]...);
        __syncwarp();
        unsigned jobs = __ballot_sync(kFull, (ctx.pend & 4u) != 0u);
        while (jobs) {
            const int src = __ffs(jobs) - 1;
            jobs &= jobs - 1;
            Nested(ctx, __shfl_sync(kFull, ctx.nested[0], src), __shfl_sync(kFull, ctx.nested[1], src),
                   __shfl_sync(kFull, ctx.nested[2], src), __shfl_sync(kFull, ctx.nested[3], src));
        }
        ctx.pend &= ~4u;
        __syncwarp();
        if (__any_sync(kFull, ctx.pend != 0u)) applyPending(ctx);
    }
}
template <auto Fn, auto Nested, typename... Cs, int... Is>
__device__ __forceinline__ void parallelForNestedImpl(Context &ctx, const NodeDev &nd, std::integer_sequence<int, Is...>)
{
    const int lane = laneId();
    for (int m = 0; m < nd.num_match; m++) {
        const MatchDev &mt = nd.match[m];
        const int32_t n = ctx.numRows(mt.arch);
        sliceLoopNested<Fn, Nested, Cs...>(ctx, n, lane, ctx.column<Cs>(mt.arch, mt.col[Is])...);
        __syncwarp();
    }
}
template <auto Fn, auto Nested, typename... Cs>
__device__ __forceinline__ void parallelForNested(Context &ctx, const NodeDev &nd)
{
    parallelForNestedImpl<Fn, Nested, Cs...>(ctx, nd, std::make_integer_sequence<int, (int)sizeof...(Cs)>());
}

// ParallelForNode in immediate mode: systems may call makeEntityNowImmediate / destroyEntityNowImmediate
// (any number, results returned at once); see Context::waitTurn. The row count of a table is read once
// per archetype, as the reference's loop bound is (state.inl:358-396).
template <auto Fn, typename... Cs>
__device__ __forceinline__ void sliceLoopImmediate(Context &ctx, int32_t n, int lane, Cs *...cols)
{
    volatile WarpShared *vws = ctx.ws;
    for (int32_t base = 0; base < n; base += 32) {
        const int32_t row = base + lane;
        const unsigned active = __ballot_sync(kFull, row < n);
        if (lane == 0) vws->done = ~active;
        __syncwarp();
        if (row < n) {
            Fn(ctx, cols[row]...);
            __threadfence_block();
            atomicOr(&ctx.ws->done, 1u << lane);
        }
        __syncwarp();
    }
}
template <auto Fn, typename... Cs, int... Is>
__device__ __forceinline__ void parallelForImmediateImpl(Context &ctx, const NodeDev &nd, std::integer_sequence<int, Is...>)
{
    const int lane = laneId();
    for (int m = 0; m < nd.num_match; m++) {
        const MatchDev &mt = nd.match[m];
        const int32_t n = ctx.numRows(mt.arch);
        sliceLoopImmediate<Fn, Cs...>(ctx, n, lane, ctx.column<Cs>(mt.arch, mt.col[Is])...);
        __syncwarp();
    }
}
template <auto Fn, typename... Cs>
__device__ __forceinline__ void parallelForImmediate(Context &ctx, const NodeDev &nd)
{
    parallelForImmediateImpl<Fn, Cs...>(ctx, nd, std::make_integer_sequence<int, (int)sizeof...(Cs)>());
}

// ---------------------------------------------------------------- megakernel --
// TaskGraph::run (src/core/taskgraph.cpp:117-122) for nodes [first, last) of one world, one warp.
// `builtin(ctx, nd)` dispatches a library-compiled func id (the generated switch of the reference,
// src/mw/cuda_exec.cpp:553-564); nodes compiled in the user's translation unit arrive as pointers.
// The descriptors every warp of a block reads for every world -- archetype tables (column addresses, sizes)
// and the nodes of the launch -- copied to shared memory once per block: the walk of one world is a chain of
// dependent loads descriptor -> column -> row, and its first link then costs a shared-memory access
// instead of a trip to L2 (profiles/r01e_summary.md: long scoreboard on exactly these loads).
constexpr int kSharedArch = 8, kSharedNodes = 16;
struct BlockDescriptors {
    ArchDev arch[kSharedArch];
    NodeDev node[kSharedNodes];
};
// All threads of the block. Returns the node array to walk (indexed from `first`).
__device__ __forceinline__ const NodeDev *loadDescriptors(BlockDescriptors *bd, Context &ctx, const EcsDev *E, int first, int last)
{
    const bool arch_fits = E->num_arch <= kSharedArch, nodes_fit = last - first <= kSharedNodes;
    if (arch_fits) {
        const int words = E->num_arch * (int)(sizeof(ArchDev) / 4);
        for (int i = threadIdx.x; i < words; i += blockDim.x) ((uint32_t *)bd->arch)[i] = ((const uint32_t *)E->arch)[i];
    }
    if (nodes_fit) {
        const int words = (last - first) * (int)(sizeof(NodeDev) / 4);
        for (int i = threadIdx.x; i < words; i += blockDim.x) ((uint32_t *)bd->node)[i] = ((const uint32_t *)(E->node + first))[i];
    }
    __syncthreads();
    ctx.archs = arch_fits ? bd->arch : E->arch;
    return nodes_fit ? bd->node : E->node + first;
}

#ifdef MDR_ECS_NODE_CLOCKS
__device__ unsigned long long g_node_clocks[kMaxNodes + 2];     // profiling builds: clock64 per node, summed over warps
#endif

// block_aligned: every warp of the block walks a world (none idles) and the caller wants the block's warps to enter
// each node together: the 8 warps then fetch the same instructions at the same time instead of 8 different nodes'
// worth (with 64 warps per SM at 64 different places the step stalls on instruction fetch: ncu `no_instruction`).
template <typename Builtin>
__device__ __forceinline__ void runNodes(Context &ctx, const NodeDev *nodes, int count, Builtin builtin,
                                         bool block_aligned = false)
{
#ifdef MDR_ECS_NODE_CLOCKS
    long long t_last = clock64();
#endif
    for (int ni = 0; ni < count; ni++) {
        if (block_aligned) __syncthreads();
#ifdef MDR_ECS_NODE_CLOCKS
        if (ni > 0 && laneId() == 0) { const long long now = clock64(); atomicAdd(&g_node_clocks[ni - 1], (unsigned long long)(now - t_last)); t_last = now; }
#endif
        const NodeDev &nd = nodes[ni];
        switch (nd.kind) {
        case kNodeParallelFor:
            if (nd.func == 0) nd.fn(ctx, nd);
            else builtin(ctx, nd);
            break;
        case kNodeClearTmp: clearArchetype(ctx, nd.arch, true); break;       // taskgraph.inl:88-92
        case kNodeSort: sortArchetype(ctx, nd.arch, nd.key_col, nd.remap != 0); break;
        case kNodeResetTmpAlloc:                                              // taskgraph.inl:94-101
            __syncwarp();
            if (laneId() == 0) ctx.ws->tmp_used = 0;
            break;
        default: break;     // kNodePhysics: a launch boundary of the host-side run graph, not a megakernel node
        }
        __syncwarp();
    }
#ifdef MDR_ECS_NODE_CLOCKS
    if (count > 0 && laneId() == 0) atomicAdd(&g_node_clocks[count - 1], (unsigned long long)(clock64() - t_last));
#endif
}

constexpr int kWarpsPerBlock = 8;

struct NoBuiltin { __device__ __forceinline__ void operator()(Context &, const NodeDev &) const { } };

// The persistent task-graph megakernel: one warp walks nodes [first, last) of one world, worlds
// strided over the grid. Instantiated by the library for its built-in worlds and, through
// madrona/b200/mw_gpu.hpp, in the translation unit that holds a user's systems.
template <int kMinBlocks, typename Builtin>
__global__ void __launch_bounds__(32 * kWarpsPerBlock, kMinBlocks)
ecsGraphKernel(const EcsDev *__restrict__ E, int first, int last)
{
    __shared__ WarpShared shared[kWarpsPerBlock];
    __shared__ BlockDescriptors descriptors;
    WarpShared *ws = &shared[threadIdx.x >> 5];
    const int warps = gridDim.x * kWarpsPerBlock;
    Context ctx;
    const NodeDev *nodes = loadDescriptors(&descriptors, ctx, E, first, last);
    for (int w = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < E->num_worlds; w += warps) {
        loadWorld(ctx, E, ws, w);
        runNodes(ctx, nodes, last - first, Builtin());
        storeWorld(ctx);
    }
}

}  // namespace ecs
}  // namespace mdr
