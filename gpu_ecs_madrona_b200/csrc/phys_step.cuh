// The whole rigid-body step of one world (node order of the reference graph:
// src/physics/physics.cpp:1142-1205) and the rare BVH rebuild.
#pragma once
#include "phys_world.cuh"
#include "phys_solver.cuh"
#include "phys_narrow.cuh"
#include "phys_sat_warp.cuh"

namespace mdr {

// Steps world `w` in world slot `v`. `group_level` is one int shared by the
// whole block: the level loops must run the same number of barriers in every
// world slot, so they run to the block-wide maximum. A slot without a world
// this round (`!v.active`) walks the same barriers with zero rows.
MDR_HD void stepWorld(const BlockCtx &c, WorldView &v, const StepParams &p, int w, int *group_level)
{
    v.n = v.active ? p.t.row_count[w] : 0;
    if (v.n > v.np) v.n = v.np;
    c.syncBlock();                           // world slots re-align once per step
    if (c.tid() == 0) *group_level = 0;     // every slot writes the same value
    loadWorld(c, v, p, w);
    c.sync();
    updateLeavesAndRefit(c, v, p);
    c.sync();
    findPairs(c, v, p, w);
    if (c.lockstep() && c.tid() == 0 && v.misc[1] > 0) {
#if defined(__CUDA_ARCH__)
        atomicMax(group_level, v.misc[1]);
#else
        *group_level = v.misc[1];
#endif
    }
    c.sync();
    const int max_level = c.lockstep() ? *group_level : v.misc[1];
    const bool probing = v.active && p.probe.world == w;
    for (int sub = 0; sub < p.substeps; sub++) {
        substepBodies(c, v, p, w);
        c.sync();
        worldBoxes(c, v, p);
        c.sync();
        const int np = v.misc[0];
        for (int i = c.tid(); i < np; i += c.nthreads()) narrowphasePair(v, p, i, probing, sub);
        c.sync();
#if defined(__CUDA_ARCH__)
        satQueueWarp(c, v, p, probing, sub);
        c.sync();
#endif
        solveLevels<false>(c, v, p, max_level);
        setVelocities(c, v, p);
        c.sync();
        solveLevels<true>(c, v, p, max_level);
    }
    if (c.tid() == 0 && v.active) {
        int nc = 0;
        for (int i = 0; i < v.misc[0]; i++) nc += contactSlot(v, i)->hdr != 0;
        p.world_contacts[w] = nc;
    }
    updateLeavesAndRefitGlobal(c, v, p, w);
    c.sync();
    storeWorld(c, v, p.t, w);
    c.sync();
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
MDR_HD void bvhRebuildWorld(const float *leaf, float *slot, uint16_t *order, uint16_t *sorted, int n)
{
    for (int i = 0; i < n; i++) {
        sorted[i] = (uint16_t)i;
        for (int k = 0; k < 6; k++) slot[6 * i + k] = leaf[6 * i + k];
    }
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
