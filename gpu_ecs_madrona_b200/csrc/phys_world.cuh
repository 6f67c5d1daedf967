// One world's rigid-body step, executed by a group of threads out of shared
// memory: the world's rows are staged HBM -> SMEM once, every node of the
// reference's physics task graph (src/physics/physics.cpp:1142-1205,
// broadphase.cpp:934-991, narrowphase.cpp:1768-1782) runs out of SMEM with a
// block barrier where the graph has an edge, and the rows go back once.
// A thread block steps several worlds side by side (one "world slot" each) in
// barrier lock-step, so that its warps share instruction fetches: the step is a
// long instruction stream and the SM's instruction cache is small.
//
// What is restructured relative to the reference (results stay bit-identical):
//  * Broadphase. The reference walks a 4-wide BVH whose boxes only ever grow
//    between rebuilds (broadphase.cpp:545-642). Ancestor boxes are supersets of
//    leaf slot boxes, so the pair SET is {(a,b): leaf_aabb[a] overlaps
//    slot_box[b]} and the pair ORDER is (row of a, rank of b in the tree's LIFO
//    traversal). We keep only the per-leaf grow-only slot box and the rank
//    permutation `order` (rebuilt by bvhRebuildWorld when the reference would
//    rebuild), and test all leaves in rank order out of SMEM.
//  * Solver. The reference's Gauss-Seidel is sequential over contacts
//    (physics.cpp:650-671,995-1008). Contacts that share no mutable body commute
//    exactly, so pairs get dependency levels from the reference order once per
//    step and each level is solved in parallel, one thread per contact.
//  * Stable static bodies (zero inverse mass and inertia, pose a fixed point of
//    the solver's write-back) contribute exact zeros; their arithmetic is skipped
//    where the result is provably the same bits (see phys_solver.cuh).
#pragma once
#include "phys_types.cuh"

namespace mdr {

// ---------------------------------------------------------------- context ---
// The threads that step one world: `nth` threads with index `tid`; sync() is
// a barrier of the whole block (every world slot of the block takes part).
#if defined(__CUDACC__)
struct BlockCtx {
    int tid_, nth_;
    int bar_;     // 0: barriers are block-wide (world slots in lock-step); else this slot's named barrier
    __host__ __device__ __forceinline__ int tid() const { return tid_; }
    __host__ __device__ __forceinline__ int nthreads() const { return nth_; }
    __host__ __device__ __forceinline__ bool lockstep() const { return bar_ == 0; }
    __host__ __device__ __forceinline__ void sync() const
    {
#if defined(__CUDA_ARCH__)
        if (bar_ == 0) __syncthreads();
        else asm volatile("bar.sync %0, %1;" ::"r"(bar_), "r"(nth_) : "memory");
#endif
    }
    // barrier of the whole block regardless of the mode
    __host__ __device__ __forceinline__ void syncBlock() const
    {
#if defined(__CUDA_ARCH__)
        __syncthreads();
#endif
    }
};
#else
// Host emulation of a 1-thread group (tests/emu only).
struct BlockCtx {
    int tid() const { return 0; }
    int nthreads() const { return 1; }
    bool lockstep() const { return true; }
    void sync() const {}
    void syncBlock() const {}
};
#endif
#if defined(__CUDA_ARCH__)
#define MDR_LDG(p) __ldg(p)
#else
#define MDR_LDG(p) (*(p))
#endif

// Per-body flags packed with the object id.
enum { META_STATIC = 1u << 16, META_DEP = 1u << 17, META_HULL = 1u << 18 };

// SMEM view of one world.
struct WorldView {
    int n;                 // live rows
    int np;                // row capacity of the arrays below (>= n)
    bool active;           // this slot holds a world in this round
    float *pos, *rot, *vel, *prev, *prev_vel, *scale;
    uint32_t *meta;        // object id | META_*
    float *box;            // [np][6] leaf AABBs (broadphase) / world AABBs (narrowphase)
    uint32_t *pairs;       // [max_pairs] row a | row b << 10 | level << 20 (11 bits) | PAIR_SAT
    uint32_t *level_any;   // block-shared [2] (substep parity): levels that hold a contact; null without dense lists
    uint8_t *hint;         // [max_pairs] per pair slot: 1 + face + 32 * (face of b), the face that separated the
                           // pair's hulls when a warp last ran the SAT on this slot (see satHintSeparates)
    int32_t *misc;         // [0] num pairs  [1] max level  [2] queued hull-hull pairs
    // --- broadphase scratch, aliased with the contact slots ---
    float *slot;           // [np][6] grow-only slot boxes
    int32_t *ent_id;       // [np]
    int32_t *cnt;          // [np+1]
    uint32_t *mask;        // [np][mask_words] hit bits in rank order
    uint16_t *order;       // [np] leaf at traversal rank r
    uint16_t *last_level;  // [np]
    uint16_t *row_perm;    // [np] rows in the order the per-row substep loops take them: stable statics last
    uint16_t *rank_of;     // [np] traversal rank of leaf i (inverse of order)
    float *rbox;           // [np][8] per traversal RANK, two 16-byte loads: slot lo.xyz, entity id bits,
                           //         slot hi.xyz, (leaf | static << 15) bits
    int mask_words;
    // --- substep phase ---
    SolverContact *contacts;   // [smem_contacts]
    SolverContact *overflow;   // this slot's global overflow contacts
    int smem_contacts;
};

// Pair word: the dependency level, and the per-substep mark "this hull-hull pair
// passed the box test and waits for a whole warp" (phys_sat_warp.cuh).
constexpr uint32_t PAIR_SAT = 1u << 31;
MDR_HD int pairLevel(uint32_t word) { return (int)((word >> 20) & 0x7ffu); }

MDR_HD int popcount32(uint32_t m)
{
#if defined(__CUDA_ARCH__)
    return __popc(m);
#else
    return __builtin_popcount(m);
#endif
}
// Position of the j-th (0-based) set bit of m; m has more than j set bits.
MDR_HD int nthSetBit(uint32_t m, int j)
{
#if defined(__CUDA_ARCH__)
    return (int)__fns(m, 0u, j + 1);
#else
    for (; j > 0; j--) m &= m - 1;
    return __builtin_ctz(m);
#endif
}

MDR_HD size_t alignUp(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Carves the view out of one SMEM (or host) buffer; returns bytes used. With
// base == nullptr only sizes the buffer.
MDR_HD size_t carveWorld(WorldView &v, char *base, int np, int max_pairs, int smem_contacts)
{
    size_t off = 0;
#define MDR_CARVE(field, type, count) \
    off = alignUp(off, 16); v.field = (type *)(base + off); off += sizeof(type) * (size_t)(count);
    v.np = np;
    v.level_any = nullptr;
    v.mask_words = (np + 31) / 32;
    v.smem_contacts = smem_contacts;
    MDR_CARVE(rot, float, 4 * np)
    MDR_CARVE(pos, float, 3 * np)
    MDR_CARVE(vel, float, 6 * np)
    MDR_CARVE(prev, float, 7 * np)
    MDR_CARVE(prev_vel, float, 6 * np)
    MDR_CARVE(scale, float, 3 * np)
    MDR_CARVE(meta, uint32_t, np)
    MDR_CARVE(box, float, 6 * np)
    MDR_CARVE(pairs, uint32_t, max_pairs)
    MDR_CARVE(hint, uint8_t, (max_pairs + 3) / 4 * 4)
    MDR_CARVE(misc, int32_t, 4)
    MDR_CARVE(row_perm, uint16_t, np)
    off = alignUp(off, 16);
    const size_t region = off;
    MDR_CARVE(contacts, SolverContact, smem_contacts)
    const size_t end_contacts = off;
    off = region;
    MDR_CARVE(slot, float, 6 * np)
    MDR_CARVE(ent_id, int32_t, np)
    MDR_CARVE(cnt, int32_t, np + 1)
    MDR_CARVE(mask, uint32_t, np * v.mask_words)
    MDR_CARVE(order, uint16_t, np)
    MDR_CARVE(last_level, uint16_t, np)
    MDR_CARVE(rank_of, uint16_t, np)
    MDR_CARVE(rbox, float, 8 * np)
#undef MDR_CARVE
    if (end_contacts > off) off = end_contacts;
    return alignUp(off, 16);
}

MDR_HD V3 ld3(const float *p, int i) { return mk3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
MDR_HD void st3(float *p, int i, V3 v) { p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z; }
MDR_HD Q4 ld4(const float *p, int i) { return mk4(p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]); }
MDR_HD void st4(float *p, int i, Q4 q) { p[4 * i] = q.w; p[4 * i + 1] = q.x; p[4 * i + 2] = q.y; p[4 * i + 3] = q.z; }
MDR_HD Box ldBox(const float *p, int i)
{
    Box b;
    b.lo = mk3(p[6 * i], p[6 * i + 1], p[6 * i + 2]);
    b.hi = mk3(p[6 * i + 3], p[6 * i + 4], p[6 * i + 5]);
    return b;
}
MDR_HD void stBox(float *p, int i, const Box &b)
{
    p[6 * i] = b.lo.x; p[6 * i + 1] = b.lo.y; p[6 * i + 2] = b.lo.z;
    p[6 * i + 3] = b.hi.x; p[6 * i + 4] = b.hi.y; p[6 * i + 5] = b.hi.z;
}
// SubstepPrevState / PreSolvePositional: {Vector3, Quat} packed in 28 B.
MDR_HD V3 ldPoseX(const float *p, int i) { return mk3(p[7 * i], p[7 * i + 1], p[7 * i + 2]); }
MDR_HD Q4 ldPoseQ(const float *p, int i) { return mk4(p[7 * i + 3], p[7 * i + 4], p[7 * i + 5], p[7 * i + 6]); }
MDR_HD void stPose(float *p, size_t i, V3 x, Q4 q)
{
    p[7 * i] = x.x; p[7 * i + 1] = x.y; p[7 * i + 2] = x.z;
    p[7 * i + 3] = q.w; p[7 * i + 4] = q.x; p[7 * i + 5] = q.y; p[7 * i + 6] = q.z;
}
MDR_HD V3 ldVelLin(const float *p, int i) { return mk3(p[6 * i], p[6 * i + 1], p[6 * i + 2]); }
MDR_HD V3 ldVelAng(const float *p, int i) { return mk3(p[6 * i + 3], p[6 * i + 4], p[6 * i + 5]); }
MDR_HD void stVel(float *p, int i, V3 l, V3 a)
{
    p[6 * i] = l.x; p[6 * i + 1] = l.y; p[6 * i + 2] = l.z;
    p[6 * i + 3] = a.x; p[6 * i + 4] = a.y; p[6 * i + 5] = a.z;
}

// Copies `words` 32-bit words between a world tile and SMEM, all threads of the world.
template <typename T>
MDR_HD void copyWords(const BlockCtx &c, T *dst, const T *src, int words)
{
    for (int i = c.tid(); i < words; i += c.nthreads()) dst[i] = src[i];
}

// HBM -> SMEM without a register round trip: every copy of the tile is in flight
// before the first one is waited for (world tiles are only 4-byte aligned, so
// the copies are word-sized). Completed by fetchWait() + a barrier.
MDR_HD void fetchWords(const BlockCtx &c, float *dst, const float *src, int words)
{
#if defined(__CUDA_ARCH__)
    for (int i = c.tid(); i < words; i += c.nthreads()) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + i)),
                     "l"(src + i) : "memory");
    }
#else
    copyWords(c, dst, src, words);
#endif
}
// The same for the columns the step itself writes (poses, velocities, slot boxes, hints): read at L2 (.cg). Inside a
// batch of steps in one launch (phys_kernel.cuh) the previous step of this world may have run on another SM, and an
// L1 line this SM kept from an even earlier step would be stale; L2 is the point of coherence. 16-byte copies when
// the tile starts on a 16-byte boundary (a row stride that is a multiple of 4 rows: every tile of the library's own
// tables), a quarter of the copy instructions; the last few words, and unaligned tiles, through registers.
MDR_HD void fetchWordsL2(const BlockCtx &c, float *dst, const float *src, int words)
{
#if defined(__CUDA_ARCH__)
    int done = 0;
    if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        const int chunks = words >> 2;
        for (int i = c.tid(); i < chunks; i += c.nthreads()) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + 4 * i)),
                         "l"(src + 4 * i) : "memory");
        }
        done = chunks << 2;
    }
    for (int i = done + c.tid(); i < words; i += c.nthreads()) dst[i] = __ldcg(src + i);
#else
    copyWords(c, dst, src, words);
#endif
}
MDR_HD void fetchWait()
{
#if defined(__CUDA_ARCH__)
    asm volatile("cp.async.wait_all;" ::: "memory");
#endif
}

// A static body whose pose the positional solver's write-back leaves bit-identical
// (physics.cpp:185-211 with zero inverse mass and inertia): no component is -0
// and q is a fixed point of Quat::normalize. Such a body does not order contacts.
MDR_HD bool staticStable(V3 x, Q4 q)
{
    const float f[7] = { x.x, x.y, x.z, q.w, q.x, q.y, q.z };
    bool neg_zero = false;
#pragma unroll
    for (int i = 0; i < 7; i++) neg_zero = neg_zero || (f[i] == 0.f && signbit(f[i]));
    return !neg_zero && qBitsEqual(qnormalize(q), q);
}

// Profiling builds (profiles/exp, -DMDR_PHASE_CLOCKS): clock64 between block barriers, per phase.
#if defined(MDR_PHASE_CLOCKS) && defined(__CUDA_ARCH__)
#define MDR_PHASE(k)                                                                     \
    do {                                                                                 \
        if (threadIdx.x == 0) {                                                          \
            const long long now_ = clock64();                                            \
            atomicAdd(&p.phase_clocks[k], (unsigned long long)(now_ - phase_last_));     \
            phase_last_ = now_;                                                          \
        }                                                                                \
    } while (0)
#define MDR_PHASE_BEGIN long long phase_last_ = clock64();
#else
#define MDR_PHASE(k) do { } while (0)
#define MDR_PHASE_BEGIN
#endif

MDR_HD void raiseError(const StepParams &p, int bit)
{
#if defined(__CUDA_ARCH__)
    atomicOr(p.error_flags, bit);
#else
    *p.error_flags |= bit;
#endif
}

// ------------------------------------------------------------- load/store ---
MDR_HD void loadWorld(const BlockCtx &c, WorldView &v, const StepParams &p, int w)
{
    const PhysTables &t = p.t;
    const size_t r0 = (size_t)w * t.stride;
    const int n = v.n;
    fetchWords(c, v.scale, t.scale + 3 * r0, 3 * n);
    fetchWordsL2(c, v.pos, t.pos + 3 * r0, 3 * n);
    fetchWordsL2(c, v.rot, t.rot + 4 * r0, 4 * n);
    fetchWordsL2(c, v.vel, t.vel + 6 * r0, 6 * n);
    fetchWordsL2(c, v.slot, t.slot_box + 6 * r0, 6 * n);
    if (p.sat_hint) {
        const int hw = (p.max_pairs + 3) / 4;
        fetchWordsL2(c, reinterpret_cast<float *>(v.hint), reinterpret_cast<const float *>(p.sat_hint + (size_t)w * hw * 4), hw);
    }
    for (int i = c.tid(); i < n; i += c.nthreads()) {
        // four independent loads, then the object table (cached)
        const int32_t ent = t.entity[2 * (r0 + i) + 1];
        uint32_t obj = (uint32_t)t.obj[r0 + i];
        // the ObjectID column is the caller's (raw writes, tensors, an ECS system): an id outside the object table is
        // flagged and treated as object 0 instead of being dereferenced (mdr_phys_sync answers MDR_EINVAL)
        if (obj >= (uint32_t)p.a.num_objs) { raiseError(p, MDR_ERR_BAD_OBJECT); obj = 0u; }
        const uint32_t resp = t.resp[r0 + i];
        const uint32_t b = t.order[r0 + i];
        v.ent_id[i] = ent;
        uint32_t m = obj & 0xffffu;
        if (resp == 2u) m |= META_STATIC;
        if (MDR_LDG(&p.a.objs[obj].type) == 2) m |= META_HULL;
        v.meta[i] = m;
        // thread i also owns traversal rank i: the leaf met i-th by the reference's walk
        v.order[i] = (uint16_t)b;
        v.rank_of[b] = (uint16_t)i;
    }
    fetchWait();
}

MDR_HD void storeWorld(const BlockCtx &c, const WorldView &v, const StepParams &p, int w)
{
    const PhysTables &t = p.t;
    const size_t r0 = (size_t)w * t.stride;
    const int n = v.n;
    if (p.sat_hint && v.active) {     // only the slots of this step's pairs can have changed
        const int hw = (v.misc[0] + 3) / 4;
        copyWords(c, reinterpret_cast<uint32_t *>(p.sat_hint + (size_t)w * ((p.max_pairs + 3) / 4) * 4),
                  reinterpret_cast<const uint32_t *>(v.hint), hw);
    }
    copyWords(c, t.pos + 3 * r0, v.pos, 3 * n);
    copyWords(c, t.rot + 4 * r0, v.rot, 4 * n);
    copyWords(c, t.vel + 6 * r0, v.vel, 6 * n);
    copyWords(c, t.prev + 7 * r0, v.prev, 7 * n);
    copyWords(c, t.prev_vel + 6 * r0, v.prev_vel, 6 * n);
    copyWords(c, t.leaf_aabb + 6 * r0, v.box, 6 * n);
}

// ------------------------------------------------------------- broadphase ---
MDR_HD Box objAABB(const ObjDev *o)
{
    Box b;
    b.lo = mk3(MDR_LDG(&o->aabb[0]), MDR_LDG(&o->aabb[1]), MDR_LDG(&o->aabb[2]));
    b.hi = mk3(MDR_LDG(&o->aabb[3]), MDR_LDG(&o->aabb[4]), MDR_LDG(&o->aabb[5]));
    return b;
}

MDR_HD Box growBox(Box s, const Box &leaf)
{
    if (leaf.lo.x < s.lo.x) s.lo.x = leaf.lo.x;
    if (leaf.lo.y < s.lo.y) s.lo.y = leaf.lo.y;
    if (leaf.lo.z < s.lo.z) s.lo.z = leaf.lo.z;
    if (leaf.hi.x > s.hi.x) s.hi.x = leaf.hi.x;
    if (leaf.hi.y > s.hi.y) s.hi.y = leaf.hi.y;
    if (leaf.hi.z > s.hi.z) s.hi.z = leaf.hi.z;
    return s;
}

MDR_HD Box leafBox(const WorldView &v, const StepParams &p, int i)
{
    const ObjDev *o = p.a.objs + (v.meta[i] & 0xffffu);
    Box wb = applyTRS(objAABB(o), ld3(v.pos, i), ld4(v.rot, i), ld3(v.scale, i));
    return expandWithMotion(wb, ldVelLin(v.vel, i), p.vel_expand, p.accel_expand);
}

// updateLeafPositionsEntry -> BVH::updateLeafPosition (broadphase.cpp:858-873,
// 461-480) followed by refitEntry -> BVH::refitLeaf (:891-895,545-642) reduced
// to the leaf's own slot box (see the header comment). Start-of-step form: the
// slot boxes are in SMEM for the pair search.
MDR_HD void updateLeavesAndRefit(const BlockCtx &c, WorldView &v, const StepParams &p)
{
    for (int i = c.tid(); i < v.n; i += c.nthreads()) {
        Box leaf = leafBox(v, p, i);
        stBox(v.box, i, leaf);
        const Box grown = growBox(ldBox(v.slot, i), leaf);
        stBox(v.slot, i, grown);
        float *rb = v.rbox + 8 * v.rank_of[i];
        rb[0] = grown.lo.x; rb[1] = grown.lo.y; rb[2] = grown.lo.z;
        reinterpret_cast<int32_t *>(rb)[3] = v.ent_id[i];
        rb[4] = grown.hi.x; rb[5] = grown.hi.y; rb[6] = grown.hi.z;
        reinterpret_cast<uint32_t *>(rb)[7] = (uint32_t)i | ((v.meta[i] & META_STATIC) ? 1u << 15 : 0u);
    }
}

// End-of-step form: the SMEM slot boxes have been recycled for contacts, the
// ones written to HBM after the first refit are grown in place.
MDR_HD void updateLeavesAndRefitGlobal(const BlockCtx &c, WorldView &v, const StepParams &p, int w)
{
    float *gslot = p.t.slot_box + 6 * (size_t)w * p.t.stride;
    for (int i = c.tid(); i < v.n; i += c.nthreads()) {
        Box leaf = leafBox(v, p, i);
        stBox(v.box, i, leaf);
        stBox(gslot, i, growBox(ldBox(gslot, i), leaf));
    }
}

// Exclusive prefix sum of cnt[0..n) in place; cnt[n] = total. Run by the first
// warp of the world's thread group.
MDR_HD void groupExclusiveScan(const BlockCtx &c, int32_t *cnt, int n)
{
#if defined(__CUDA_ARCH__)
    if (c.tid() < 32) {
        const int lane = c.tid();
        int carry = 0;
        for (int base = 0; base < n; base += 32) {
            int i = base + lane;
            int x = i < n ? cnt[i] : 0;
            int incl = x;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += y;
            }
            if (i < n) cnt[i] = carry + incl - x;
            carry += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) cnt[n] = carry;
    }
#else
    int carry = 0;
    for (int i = 0; i < n; i++) { int x = cnt[i]; cnt[i] = carry; carry += x; }
    cnt[n] = carry;
#endif
}


// One rank-ordered broadphase entry against row a's current box.
struct alignas(16) RankEntry { float lx, ly, lz; int32_t ent; float hx, hy, hz; uint32_t info; };
MDR_HD bool pairTest(const float *rbox, int r, const Box &q, int a_id, bool a_static)
{
    const RankEntry e = *reinterpret_cast<const RankEntry *>(rbox + 8 * r);
    return a_id < e.ent && !(a_static && (e.info >> 15)) &&
           q.lo.x < e.hx && e.lx < q.hi.x && q.lo.y < e.hy && e.ly < q.hi.y && q.lo.z < e.hz && e.lz < q.hi.z;
}

// findOverlappingEntry + BVH::findOverlaps (broadphase.cpp:897-932,
// include/madrona/physics.inl:61-100): pairs (a, b) with a's entity id below
// b's, not both static, a's CURRENT leaf box strictly overlapping b's slot box,
// ordered by (row a, traversal rank of b). Then dependency levels. The grown
// slot boxes go back to HBM here (their SMEM is recycled for contacts).
MDR_HD void findPairs(const BlockCtx &c, WorldView &v, const StepParams &p, int w)
{
    const int n = v.n;
    const size_t r0 = (size_t)w * p.t.stride;
    MDR_PHASE_BEGIN
    // Rows are taken one per thread; the rows left over after the last full pass
    // (row 64 of 65 with 64 threads) are swept by the whole group instead, one rank
    // per thread, so that no thread walks all ranks alone while the others wait.
#if defined(__CUDA_ARCH__)
    const int full = n / c.nthreads() * c.nthreads();
#else
    const int full = n;
#endif
    for (int a = c.tid(); a < full; a += c.nthreads()) {
        const Box q = ldBox(v.box, a);
        const int a_id = v.ent_id[a];
        const bool a_static = (v.meta[a] & META_STATIC) != 0;
        // one mask word at a time (a branch-free, eight-in-flight form of the test measured
        // slower: it costs registers the rest of the kernel needs)
        int count = 0;
        for (int wd = 0; wd * 32 < n; wd++) {
            uint32_t word = 0;
#pragma unroll 8
            for (int b = 0; b < 32; b++) {
                const int r = wd * 32 + b;
                if (r < n && pairTest(v.rbox, r, q, a_id, a_static)) word |= 1u << b;
            }
            v.mask[a * v.mask_words + wd] = word;
            count += popcount32(word);
        }
        v.cnt[a] = count;
        v.last_level[a] = 0;
        if (!(a_static && staticStable(ld3(v.pos, a), ld4(v.rot, a)))) v.meta[a] |= META_DEP;
    }
#if defined(__CUDA_ARCH__)
    for (int a = full + c.tid(); a < n; a += c.nthreads()) {
        v.cnt[a] = 0;
        v.last_level[a] = 0;
        if (!((v.meta[a] & META_STATIC) && staticStable(ld3(v.pos, a), ld4(v.rot, a)))) v.meta[a] |= META_DEP;
    }
    c.sync();       // every world slot of the block takes this barrier, tail rows or not
    MDR_PHASE(16);
    if (full < n) {
        const int padded = (n + 31) & ~31;
        for (int a = full; a < n; a++) {
            const Box q = ldBox(v.box, a);
            const int a_id = v.ent_id[a];
            const bool a_static = (v.meta[a] & META_STATIC) != 0;
            for (int r = c.tid(); r < padded; r += c.nthreads()) {      // whole warps iterate together
                const bool hit = r < n && pairTest(v.rbox, r, q, a_id, a_static);
                const uint32_t word = __ballot_sync(0xffffffffu, hit);
                if ((c.tid() & 31) == 0) {
                    v.mask[a * v.mask_words + (r >> 5)] = word;
                    if (word) atomicAdd(&v.cnt[a], __popc(word));
                }
            }
        }
    }
#endif
    if (v.active) copyWords(c, p.t.slot_box + 6 * r0, v.slot, 6 * n);
    c.sync();
    MDR_PHASE(17);
    groupExclusiveScan(c, v.cnt, n);
    c.sync();
    MDR_PHASE(18);
    const int total = v.cnt[n];
    // One pair slot per thread: its row by bisection of the prefix sums, its partner the
    // j-th set bit of the row's hit mask (a row with many pairs, the ground plane's 64,
    // is shared by as many threads).
    const int fill = total < p.max_pairs ? total : p.max_pairs;
    for (int k = c.tid(); k < fill; k += c.nthreads()) {
        int lo = 0, hi = n - 1;              // largest row a with cnt[a] <= k
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (v.cnt[mid] <= k) lo = mid; else hi = mid - 1;
        }
        const int a = lo;
        int j = k - v.cnt[a], wd = 0;
        uint32_t m = v.mask[a * v.mask_words];
        for (int cbits = popcount32(m); j >= cbits; cbits = popcount32(m)) {
            j -= cbits;
            m = v.mask[a * v.mask_words + ++wd];
        }
        v.pairs[k] = (uint32_t)a | ((uint32_t)v.order[wd * 32 + nthSetBit(m, j)] << 10);
    }
    c.sync();
    MDR_PHASE(19);
    // A stable static body (see staticStable) comes out of every substep exactly as it
    // went in, so the per-row substep loops only visit it in the first substep: such
    // rows go to the end of the row order (misc[3] = rows before them). Built by the
    // second warp's first lane while thread 0 assigns the levels.
#if defined(__CUDA_ARCH__)
    if ((c.tid() >> 5) == (c.nthreads() > 32 ? 1 : 0)) {
        const int lane = c.tid() & 31;
        int deps = 0;
        for (int base = 0; base < n; base += 32)
            deps += __popc(__ballot_sync(0xffffffffu, base + lane < n && (v.meta[base + lane] & META_DEP)));
        int at_dep = 0, at_rest = deps;
        for (int base = 0; base < n; base += 32) {
            const int i = base + lane;
            const bool live = i < n, dep = live && (v.meta[i] & META_DEP);
            const uint32_t md = __ballot_sync(0xffffffffu, dep), mr = __ballot_sync(0xffffffffu, live && !dep);
            const uint32_t below = (1u << lane) - 1u;
            if (dep) v.row_perm[at_dep + __popc(md & below)] = (uint16_t)i;
            else if (live) v.row_perm[at_rest + __popc(mr & below)] = (uint16_t)i;
            at_dep += __popc(md);
            at_rest += __popc(mr);
        }
        if (lane == 0) v.misc[3] = deps;
    }
#else
    {
        int k = 0;
        for (int i = 0; i < n; i++) if (v.meta[i] & META_DEP) v.row_perm[k++] = (uint16_t)i;
        v.misc[3] = k;
        for (int i = 0; i < n; i++) if (!(v.meta[i] & META_DEP)) v.row_perm[k++] = (uint16_t)i;
    }
#endif
    // Level of a pair = 1 + the latest earlier pair sharing an ordering body: a walk over
    // the pairs in order. The pairs of one row are contiguous and their second bodies
    // all differ, so when the row itself does not order contacts (a stable static: the
    // ground plane and its 64 pairs) the whole run is independent and the first warp
    // takes it at once; the other pairs are walked by its first lane.
#if defined(__CUDA_ARCH__)
    if (c.tid() < 32) {
        const int lane = c.tid();
        int np = total;
        if (np > p.max_pairs) { np = p.max_pairs; if (lane == 0) raiseError(p, MDR_ERR_PAIR_OVERFLOW); }
        int max_level = 0;
        bool too_deep = false;
        int i = 0;
        while (i < np) {
            const uint32_t pr = v.pairs[i];
            const int a = pr & 1023;
            if (!(v.meta[a] & META_DEP)) {
                int end = v.cnt[a + 1];
                if (end > np) end = np;
                for (int k = i + lane; k < end; k += 32) {
                    const uint32_t q = v.pairs[k];
                    const int b = (q >> 10) & 1023;
                    int lv = 1;
                    if (v.meta[b] & META_DEP) {
                        lv = v.last_level[b] + 1;
                        if (lv > MDR_MAX_LEVEL) { lv = MDR_MAX_LEVEL; too_deep = true; }
                        v.last_level[b] = (uint16_t)lv;
                    }
                    v.pairs[k] = q | ((uint32_t)lv << 20);
                    if (lv > max_level) max_level = lv;
                }
                i = end;
            } else {
                // a stretch of pairs whose first body orders contacts: the first lane walks it alone, without a
                // warp barrier per pair, up to the next run the whole warp can take
                int j = i;
                if (lane == 0) {
                    while (j < np) {
                        const uint32_t q = v.pairs[j];
                        const int qa = q & 1023;
                        if (!(v.meta[qa] & META_DEP)) break;
                        const int b = (q >> 10) & 1023;
                        const bool db = (v.meta[b] & META_DEP) != 0;
                        const int la = v.last_level[qa];
                        const int lb = db ? v.last_level[b] : 0;
                        int lv = (la > lb ? la : lb) + 1;
                        if (lv > MDR_MAX_LEVEL) { lv = MDR_MAX_LEVEL; too_deep = true; }
                        v.last_level[qa] = (uint16_t)lv;
                        if (db) v.last_level[b] = (uint16_t)lv;
                        v.pairs[j] = q | ((uint32_t)lv << 20);
                        if (lv > max_level) max_level = lv;
                        j++;
                    }
                }
                i = __shfl_sync(0xffffffffu, j, 0);
            }
            __syncwarp();
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            const int o = __shfl_xor_sync(0xffffffffu, max_level, d);
            if (o > max_level) max_level = o;
        }
        if (__any_sync(0xffffffffu, too_deep) && lane == 0) raiseError(p, MDR_ERR_PAIR_OVERFLOW);
    if (lane == 0) {
#else
    if (c.tid() == 0) {
        int np = total;
        if (np > p.max_pairs) { np = p.max_pairs; raiseError(p, MDR_ERR_PAIR_OVERFLOW); }
        int max_level = 0;
        for (int i = 0; i < np; i++) {
            const uint32_t pr = v.pairs[i];
            const int a = pr & 1023, b = (pr >> 10) & 1023;
            const bool da = (v.meta[a] & META_DEP) != 0, db = (v.meta[b] & META_DEP) != 0;
            int la = da ? v.last_level[a] : 0;
            int lb = db ? v.last_level[b] : 0;
            int lv = (la > lb ? la : lb) + 1;
            if (lv > MDR_MAX_LEVEL) { lv = MDR_MAX_LEVEL; raiseError(p, MDR_ERR_PAIR_OVERFLOW); }
            if (da) v.last_level[a] = (uint16_t)lv;
            if (db) v.last_level[b] = (uint16_t)lv;
            v.pairs[i] = pr | ((uint32_t)lv << 20);
            if (lv > max_level) max_level = lv;
        }
    {
#endif
        v.misc[0] = np;
        v.misc[1] = max_level;
        if (v.active) {
            p.world_pairs[w] = total;
            if (p.probe.world == w) {
                *p.probe.num_pairs = total;
                for (int i = 0; i < np && i < p.probe.max_pairs; i++) {
                    p.probe.pairs[2 * i] = v.pairs[i] & 1023;
                    p.probe.pairs[2 * i + 1] = (v.pairs[i] >> 10) & 1023;
                }
            }
        }
    }
    }
    c.sync();
    MDR_PHASE(20);
}

}  // namespace mdr
