// The whole rigid-body step of one world (node order of the reference graph:
// src/physics/physics.cpp:1142-1205) and the rare BVH rebuild.
#pragma once
#include "phys_world.cuh"
#include "phys_solver.cuh"
#include "phys_narrow.cuh"
#include "phys_sat_warp.cuh"

namespace mdr {

// ------------------------------------------------- dense solver lanes -------
// Contacts of one dependency level commute exactly, whichever thread solves
// them. Per substep the block gathers the contacts of ALL its world slots into
// one list bucketed by level, and every level is then solved by as few warps as
// it has contacts (instead of every world slot walking its own pair list with
// mostly idle lanes, level after level). Levels nobody has cost nothing, not
// even a barrier. Falls back to the per-slot walk when a level is deeper than
// kDenseLevels or the block holds more than kDenseItems contacts.
constexpr int kDenseLevels = 30;
constexpr int kDenseItems = 1024;
constexpr int kDenseSlots = 16;
struct BlockSolve {
    int32_t cnt[32];                  // per level (index 31: deeper than kDenseLevels)
    int32_t off[33];
    int32_t dense;
    uint32_t next_task;               // the block's next task (phys_kernel.cuh)
    uint32_t level_any[2];            // per substep parity: bit lv set when a contact of level lv was emitted
    uint16_t items[kDenseItems];      // world slot << 10 | pair slot
    WorldView views[kDenseSlots];
};

#if defined(__CUDACC__)
extern __shared__ __align__(16) char mdr_smem[];
#endif

// Layout: the shared-memory layout of a world slot (row stride, pair capacity, contact slots) and the threads per world
// either come with the launch (RuntimeLayout) or are compile-time constants (FixedLayout). With constants every array of
// the world view is "slot base + immediate" and the loops stride by an immediate: the view costs one register instead of
// twenty-odd, the 128-register kernel stops spilling, and the step is 8-10 % faster (profiles/r02_summary.md section 7).
// A fixed variant is compiled per registered shape (the table below holds collisions-64's); physics.cu selects it when
// the configuration matches and falls back to the runtime layout otherwise.
struct RuntimeLayout { static constexpr int kNp = 0, kPairs = 0, kContacts = 0, kTpw = 0; };
template <int N, int P, int C, int T> struct FixedLayout { static constexpr int kNp = N, kPairs = P, kContacts = C, kTpw = T; };
using Collisions64Layout = FixedLayout<68, 640, 71, 64>;     // 65 bodies, 640 candidate pairs, 71 SMEM contacts (two
                                                            // blocks then fit the 196 KB carve-out: physics.cu), 64 threads

#if defined(__CUDA_ARCH__)
// All threads of the block, once per step, after every world slot has its pair list and levels (block
// barrier passed): the PAIRS of all slots bucketed by level. Which of them carry a contact changes from
// substep to substep; the solver skips empty slots and, through `level_any`, whole levels nobody has.
// Returns the block-uniform verdict.
__device__ __forceinline__ bool buildLevelLists(const BlockCtx &c, const WorldView &v, BlockSolve *bs, int slot,
                                                int max_level)
{
    const int t = threadIdx.x;
    if (t < 32) bs->cnt[t] = 0;
    if (t == 32) { bs->level_any[0] = 0u; bs->level_any[1] = 0u; }
    __syncthreads();
    const int np = v.misc[0];
    for (int i = c.tid(); i < np; i += c.nthreads()) {
        const int lv = pairLevel(v.pairs[i]);
        atomicAdd(&bs->cnt[lv <= kDenseLevels ? lv : 31], 1);
    }
    __syncthreads();
    if (t < 32) {
        const int x = bs->cnt[t];
        int incl = x;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, incl, d);
            if (t >= d) incl += y;
        }
        bs->off[t] = incl - x;
        bs->cnt[t] = 0;
        if (t == 31) {
            bs->off[32] = incl;
            bs->dense = x == 0 && incl <= kDenseItems && max_level <= kDenseLevels;
        }
    }
    __syncthreads();
    if (!bs->dense) return false;
    for (int i = c.tid(); i < np; i += c.nthreads()) {
        const int lv = pairLevel(v.pairs[i]);
        const int at = bs->off[lv] + atomicAdd(&bs->cnt[lv], 1);
        bs->items[at] = (uint16_t)((slot << 10) | i);
    }
    __syncthreads();
    return true;
}

// `any`: the levels that hold a contact in this substep (bit lv), read by every thread after a block barrier.
// With a compile-time layout the view of the contact's world slot is arithmetic on the slot index; with the runtime
// layout it is read from the block's table of views in shared memory (a dependent load at the head of every contact).
template <bool kVelocity, typename Layout>
__device__ __forceinline__ void solveDense(BlockSolve *bs, const StepParams &p, int max_level, uint32_t any)
{
    for (int lv = 1; lv <= max_level; lv++) {
        if (!((any >> lv) & 1u)) continue;             // block-uniform
        const int begin = bs->off[lv], end = bs->off[lv + 1];
        for (int k = begin + (int)threadIdx.x; k < end; k += (int)blockDim.x) {
            const uint32_t code = bs->items[k];
            const int slot = (int)(code >> 10), pi = (int)(code & 1023u);
            if constexpr (Layout::kNp > 0) {
                WorldView vv;
                const size_t slot_bytes = carveWorld(vv, nullptr, Layout::kNp, Layout::kPairs, Layout::kContacts);
                carveWorld(vv, mdr_smem + (size_t)slot * slot_bytes, Layout::kNp, Layout::kPairs, Layout::kContacts);
                SolverContact *ct = pi < Layout::kContacts ? vv.contacts + pi : bs->views[slot].overflow + (pi - Layout::kContacts);
                if (ct->hdr == 0) continue;
                if (kVelocity) solveContactVelocities(vv, p, ct);
                else solveContactPositions(vv, p, ct);
            } else {
                WorldView &vv = bs->views[slot];
                SolverContact *ct = contactSlot(vv, pi);
                if (ct->hdr == 0) continue;
                if (kVelocity) solveContactVelocities(vv, p, ct);
                else solveContactPositions(vv, p, ct);
            }
        }
        __syncthreads();
    }
}
#endif

// Steps world `w` in world slot `v`. `group_level` is one int shared by the
// whole block: the level loops must run the same number of barriers in every
// world slot, so they run to the block-wide maximum. A slot without a world
// this round (`!v.active`) walks the same barriers with zero rows.
template <bool kJoints = true, int kHullSlots = 1, typename Layout = RuntimeLayout>
MDR_HD void stepWorld(const BlockCtx &c, WorldView &v, const StepParams &p, int w, int *group_level,
                      BlockSolve *bs = nullptr, int slot = 0)
{
    v.n = v.active ? p.t.row_count[w] : 0;
    if (v.n > v.np) v.n = v.np;
    c.syncBlock();                           // world slots re-align once per step
    MDR_PHASE_BEGIN
    if (c.tid() == 0) *group_level = 0;     // every slot writes the same value
    loadWorld(c, v, p, w);
    c.sync();
    MDR_PHASE(0);
    updateLeavesAndRefit(c, v, p);
    c.sync();
    MDR_PHASE(1);
    findPairs(c, v, p, w);
    if (c.lockstep() && c.tid() == 0 && v.misc[1] > 0) {
#if defined(__CUDA_ARCH__)
        atomicMax(group_level, v.misc[1]);
#else
        *group_level = v.misc[1];
#endif
    }
    c.sync();
    MDR_PHASE(2);
    const int max_level = c.lockstep() ? *group_level : v.misc[1];
    const bool probing = v.active && p.probe.world == w;
#if defined(__CUDA_ARCH__)
    const bool dense = bs && buildLevelLists(c, v, bs, slot, max_level);
    MDR_PHASE(6);
#endif
    for (int sub = 0; sub < p.substeps; sub++) {
        substepBodies(c, v, p, w, sub);
        worldBoxes(c, v, p, sub);
        c.sync();
        MDR_PHASE(3);
        const int np = v.misc[0];
#if defined(__CUDA_ARCH__)
        // nobody reads the other parity's flags any more (its solves ended before this substep's first barrier)
        if (bs && threadIdx.x == 0) bs->level_any[(sub + 1) & 1] = 0u;
#endif
        for (int i = c.tid(); i < np; i += c.nthreads()) narrowphasePair(v, p, i, probing, sub);
        c.sync();
        MDR_PHASE(4);
#if defined(__CUDA_ARCH__)
        satQueueWarp<kHullSlots>(c, v, p, probing, sub);
        c.sync();
        MDR_PHASE(5);
#endif
#if defined(__CUDA_ARCH__)
        if (dense) {
            const uint32_t any = bs->level_any[sub & 1];      // final: a block barrier has passed since the last emit
            solveDense<false, Layout>(bs, p, max_level, any);
            MDR_PHASE(7);
            if (kJoints) solveJoints(c, v, p, w);
            setVelocities(c, v, p, sub);
            c.sync();
            MDR_PHASE(8);
            solveDense<true, Layout>(bs, p, max_level, any);
            MDR_PHASE(9);
            continue;
        }
#endif
        solveLevels<false>(c, v, p, max_level);
        if (kJoints) solveJoints(c, v, p, w);
        setVelocities(c, v, p, sub);
        c.sync();
        solveLevels<true>(c, v, p, max_level);
    }
#if defined(__CUDA_ARCH__)
    if (c.tid() < 32 && v.active) {
        int nc = 0;
        for (int i = c.tid(); i < v.misc[0]; i += 32) nc += contactSlot(v, i)->hdr != 0;
        nc = __reduce_add_sync(0xffffffffu, nc);
        if (c.tid() == 0) p.world_contacts[w] = nc;
    }
#else
    if (c.tid() == 0 && v.active) {
        int nc = 0;
        for (int i = 0; i < v.misc[0]; i++) nc += contactSlot(v, i)->hdr != 0;
        p.world_contacts[w] = nc;
    }
#endif
    MDR_PHASE(10);
    updateLeavesAndRefitGlobal(c, v, p, w);
    c.sync();
    MDR_PHASE(11);
    storeWorld(c, v, p, w);
    c.sync();
    MDR_PHASE(12);
}

// ------------------------------------------------------------ BVH rebuild ---
// BVH::rebuild (src/physics/broadphase.cpp:42-281) reduced to what the step
// needs from it: the order in which the reference's LIFO traversal
// (include/madrona/physics.inl:61-100) meets the leaves, and slot boxes reset
// to the current leaf boxes. The tree is the reference's: every node of more
// than four leaves is cut by two levels of midpoint splits on the widest axis
// of the leaf-centre bounds (fallback: half), children in range order; the
// traversal pops the LAST pushed child first, so sub-ranges are ranked fourth,
// third, second, first, recursively, and the leaves of a <=4 range in order.
// Runs only when the reference would rebuild (init / reset), one thread per world.
MDR_HD V3 leafCentre(const float *leaf, int l)
{
    Box b = ldBox(leaf, l);
    return (b.lo + b.hi) / 2.f;
}

MDR_HD int midpointSplit(const float *leaf, uint16_t *sorted, int base, int num)
{
    V3 cmin = mk3(FLT_MAX, FLT_MAX, FLT_MAX), cmax = mk3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    for (int i = 0; i < num; i++) {
        V3 cc = leafCentre(leaf, sorted[base + i]);
        cmin = vmin(cmin, cc);
        cmax = vmax(cmax, cc);
    }
    V3 d = cmax - cmin;
    int axis = 2;
    if (d.x > d.y && d.x > d.z) axis = 0;
    else if (d.y > d.x && d.y > d.z) axis = 1;
    float split_val = 0.5f * (comp(cmin, axis) + comp(cmax, axis));
    int start = 0, end = num;
    while (start < end) {
        while (start < end && comp(leafCentre(leaf, sorted[base + start]), axis) < split_val) ++start;
        while (start < end && comp(leafCentre(leaf, sorted[base + end - 1]), axis) >= split_val) --end;
        if (start < end) {
            uint16_t t = sorted[base + start];
            sorted[base + start] = sorted[base + end - 1];
            sorted[base + end - 1] = t;
            ++start;
            --end;
        }
    }
    if (start > 0 && start < num) return start;
    return num / 2;
}

// leaf: this world's current leaf boxes [n][6]; slot/order/sorted: outputs.
// The traversal order alone; `sorted` holds 0..n-1 on entry.
MDR_HD void bvhRankOrder(const float *leaf, uint16_t *order, uint16_t *sorted, int n);

MDR_HD void bvhRebuildWorld(const float *leaf, float *slot, uint16_t *order, uint16_t *sorted, int n)
{
    for (int i = 0; i < n; i++) {
        sorted[i] = (uint16_t)i;
        for (int k = 0; k < 6; k++) slot[6 * i + k] = leaf[6 * i + k];
    }
    bvhRankOrder(leaf, order, sorted, n);
}

MDR_HD void bvhRankOrder(const float *leaf, uint16_t *order, uint16_t *sorted, int n)
{
    struct Range { int32_t off, num; };
    Range stack[128];
    int sz = 0, rank = 0;
    stack[sz].off = 0; stack[sz].num = n; sz++;
    while (sz > 0) {
        Range e = stack[--sz];
        if (e.num <= 4) {
            for (int i = 0; i < e.num; i++) order[rank++] = sorted[e.off + i];
            continue;
        }
        int second = midpointSplit(leaf, sorted, e.off, e.num);
        int h1 = second, h2 = e.num - second;
        int first = midpointSplit(leaf, sorted, e.off, h1);
        int third = midpointSplit(leaf, sorted, e.off + second, h2);
        if (sz + 4 > 128) break;   // deeper than the reference's own stack
        stack[sz].off = e.off; stack[sz].num = first; sz++;
        stack[sz].off = e.off + first; stack[sz].num = h1 - first; sz++;
        stack[sz].off = e.off + h1; stack[sz].num = third; sz++;
        stack[sz].off = e.off + h1 + third; stack[sz].num = h2 - third; sz++;
    }
}

}  // namespace mdr
