// Warp-cooperative hull-hull SAT (device only). The thread-per-pair narrowphase
// marks the hull-hull pairs that pass the AABB reject; here one warp takes one
// marked pair: lane k holds world vertex k and world face plane k of both hulls
// (32 per slot, up to MDR_MAX_HULL_VERTS / FACES = 64 in the two-slot variant), face queries run one face per lane, the
// edge-edge query one edge pair per lane, and warp arg-max reductions pick the
// same winner as the reference's sequential scans (first strict maximum =
// largest value, lowest index; src/physics/narrowphase.cpp:352-378,474-576).
// The reference's early-outs only ever trigger on a positive separation, whose
// outcome (no contact) does not depend on which axis found it. Contact
// generation for the surviving pair is done by lane 0 with the serial code.
#pragma once
#include "phys_narrow.cuh"

#if defined(__CUDACC__)
namespace mdr {

#define MDR_FULL 0xffffffffu

__device__ __forceinline__ V3 shfl3(V3 v, int src)
{
    return mk3(__shfl_sync(MDR_FULL, v.x, src), __shfl_sync(MDR_FULL, v.y, src),
               __shfl_sync(MDR_FULL, v.z, src));
}

// Hulls of up to 32 * kSlots vertices / faces: lane l holds vertex (and face plane) l of every slot, element k
// lives in slot k >> 5 of lane k & 31. kSlots = 1 is the common case (cubes, the 4-32 vertex hulls of the asset
// tests); kSlots = 2 (64 vertices / faces) is its own kernel variant, chosen at creation when an object needs it,
// so that the common case does not carry the second set of registers.
template <int kSlots> struct LaneV3 { V3 v[kSlots]; };
template <int kSlots> struct LanePlane { Plane p[kSlots]; };

// Element k; every lane of the warp takes part, k may differ from lane to lane.
template <int kSlots> __device__ __forceinline__ V3 laneGet(const LaneV3<kSlots> &a, int k)
{
    V3 r = shfl3(a.v[0], k & 31);
#pragma unroll
    for (int s = 1; s < kSlots; s++) {
        const V3 o = shfl3(a.v[s], k & 31);
        if ((k >> 5) == s) r = o;
    }
    return r;
}
template <int kSlots> __device__ __forceinline__ V3 laneGetNormal(const LanePlane<kSlots> &a, int k)
{
    V3 r = shfl3(a.p[0].n, k & 31);
#pragma unroll
    for (int s = 1; s < kSlots; s++) {
        const V3 o = shfl3(a.p[s].n, k & 31);
        if ((k >> 5) == s) r = o;
    }
    return r;
}
template <int kSlots> __device__ __forceinline__ float laneGetD(const LanePlane<kSlots> &a, int k)
{
    float r = __shfl_sync(MDR_FULL, a.p[0].d, k & 31);
#pragma unroll
    for (int s = 1; s < kSlots; s++) {
        const float o = __shfl_sync(MDR_FULL, a.p[s].d, k & 31);
        if ((k >> 5) == s) r = o;
    }
    return r;
}

// Vertex k of a hull whose world vertices sit in the lanes. Every lane of the
// warp must call at() together (contact generation below runs on all lanes
// with identical inputs, so its control flow is uniform).
template <int kSlots> struct LaneVerts {
    LaneV3<kSlots> mine;
    __device__ __forceinline__ V3 at(int k) const { return laneGet<kSlots>(mine, k); }
};

// emitContact with one contact point per lane (lanes 0..3); every lane holds
// the same manifold.
__device__ __forceinline__ void emitContactLanes(WorldView &v, const StepParams &p, SolverContact *sc,
                                                 const Manifold &m, int ref, int alt, int lane)
{
    const bool s1 = !(v.meta[ref] & META_DEP), s2 = !(v.meta[alt] & META_DEP);
    if (lane == 0) {
        sc->hdr = (uint32_t)ref | ((uint32_t)alt << 10) | ((uint32_t)m.n << 20) |
                  ((uint32_t)s1 << 24) | ((uint32_t)s2 << 25);
        sc->normal[0] = m.normal.x; sc->normal[1] = m.normal.y; sc->normal[2] = m.normal.z;
    }
    if (lane >= 4) return;
    const int i = lane;
    sc->lambda[i] = 0.f;
    if (i >= m.n) return;
    const V3 c1 = i == 0 ? m.pts[0] : i == 1 ? m.pts[1] : i == 2 ? m.pts[2] : m.pts[3];
    const float depth = i == 0 ? m.depths[0] : i == 1 ? m.depths[1] : i == 2 ? m.depths[2] : m.depths[3];
    const V3 x1 = ld3(v.pos, ref), x2 = ld3(v.pos, alt);
    const Q4 q1 = ld4(v.rot, ref), q2 = ld4(v.rot, alt);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    const V3 c2 = c1 - m.normal * depth;
    const V3 r1 = rotate(qinv(q1), c1 - x1);
    const V3 r2 = rotate(qinv(q2), c2 - x2);
    const V3 r1w = rotate(q1, r1), r2w = rotate(q2, r2);
    const V3 va = s1 ? zero : ldVelLin(v.prev_vel, ref) + cross(ldVelAng(v.prev_vel, ref), r1w);
    const V3 vb = s2 ? zero : ldVelLin(v.prev_vel, alt) + cross(ldVelAng(v.prev_vel, alt), r2w);
    sc->vn_bar[i] = dot(m.normal, va - vb);
    const V3 o1 = s1 ? r1w + x1 : r1;
    const V3 o2 = s2 ? r2w + x2 : r2;
    sc->r1[i][0] = o1.x; sc->r1[i][1] = o1.y; sc->r1[i][2] = o1.z;
    sc->r2[i][0] = o2.x; sc->r2[i][1] = o2.y; sc->r2[i][2] = o2.z;
}

// Arg-max over the warp: largest value, lowest index on ties. All lanes get it.
__device__ __forceinline__ void warpArgMax(float &val, int &idx)
{
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        float ov = __shfl_xor_sync(MDR_FULL, val, d);
        int oi = __shfl_xor_sync(MDR_FULL, idx, d);
        if (ov > val || (ov == val && oi < idx)) { val = ov; idx = oi; }
    }
}

// The faces of `a` held by this lane (one per slot): distance of the other hull (vertices in `vb`, nvb of them)
// from each, reduced to this lane's first strict maximum (value, face index).
template <int kSlots>
__device__ __forceinline__ void faceSeparationLane(const LanePlane<kSlots> &pa, int nfa, const LaneV3<kSlots> &vb, int nvb,
                                                   int lane, float &best, int &best_face)
{
    float m[kSlots];
#pragma unroll
    for (int s = 0; s < kSlots; s++) m[s] = FLT_MAX;
    for (int k = 0; k < nvb; k++) {                      // k is warp-uniform
        const V3 v = laneGet<kSlots>(vb, k);
#pragma unroll
        for (int s = 0; s < kSlots; s++) {
            const float d = dot(pa.p[s].n, v);
            if (d < m[s]) m[s] = d;
        }
    }
    best = -FLT_MAX;
    best_face = lane;
#pragma unroll
    for (int s = 0; s < kSlots; s++) {
        const int f = lane + 32 * s;
        const float sep = f < nfa ? m[s] - pa.p[s].d : -FLT_MAX;
        if (s == 0 || sep > best) { best = sep; best_face = f; }
    }
}

// Edge-edge query, one (edge of a, edge of b) pair per lane and round
// (queryEdgeDirections CPU branch, narrowphase.cpp:474-576; buildsMinkowskiFace
// :418-426; edgeDistance :433-472). Returns the warp-wide winner.
struct EdgeWin { float sep; V3 normal; int ea, eb; };

template <int kSlots>
__device__ __forceinline__ EdgeWin edgeQueryWarp(int lane, const LaneV3<kSlots> &va, const LanePlane<kSlots> &pa,
                                                 const HEdgeDev *hea_tab, const uint32_t *edges_a, int nea, V3 centre_a,
                                                 const LaneV3<kSlots> &vb, const LanePlane<kSlots> &pb,
                                                 const HEdgeDev *heb_tab, const uint32_t *edges_b, int neb)
{
    const int total = nea * neb;
    float best = -FLT_MAX;
    int best_e = 0x7fffffff;
    V3 best_n = mk3(0.f, 0.f, 0.f);
    for (int base = 0; base < total; base += 32) {
        const int e = base + lane;
        const bool valid = e < total;
        const int ia = valid ? e / neb : 0, ib = valid ? e % neb : 0;
        const uint32_t ha = MDR_LDG(edges_a + ia), hb = MDR_LDG(edges_b + ib);
        const HEdgeDev ea = ldHE(hea_tab, ha), eb = ldHE(heb_tab, hb);
        const int ea_twin_poly = (int)MDR_LDG(&hea_tab[ea.twin].poly);
        const int eb_twin_poly = (int)MDR_LDG(&heb_tab[eb.twin].poly);
        const int ea_next_root = (int)MDR_LDG(&hea_tab[ea.next].root);
        const int eb_next_root = (int)MDR_LDG(&heb_tab[eb.next].root);
        // every lane takes part in every shuffle
        const V3 an1 = laneGetNormal<kSlots>(pa, (int)ea.poly), an2 = laneGetNormal<kSlots>(pa, ea_twin_poly);
        const V3 bn1 = laneGetNormal<kSlots>(pb, (int)eb.poly), bn2 = laneGetNormal<kSlots>(pb, eb_twin_poly);
        const V3 a1 = laneGet<kSlots>(va, (int)ea.root), a2 = laneGet<kSlots>(va, ea_next_root);
        const V3 b1 = laneGet<kSlots>(vb, (int)eb.root), b2 = laneGet<kSlots>(vb, eb_next_root);
        float sep = -FLT_MAX;
        V3 normal = mk3(0.f, 0.f, 0.f);
        if (valid && minkowskiFace(an1, an2, -bn1, -bn2)) {
            V3 da = a2 - a1, db = b2 - b1;
            V3 cr = cross(da, db);
            float l2 = len2(cr);
            if (l2 != 0) {
                float inv = 1.f / sqrtf(l2);
                normal = cr * inv;
                if (dot(normal, a1 - centre_a) < 0.0f) normal = -normal;
                sep = dot(normal, b1 - a1);
            }
        }
        if (valid && sep > best) { best = sep; best_e = e; best_n = normal; }
    }
    float wv = best;
    int we = best_e;
    warpArgMax(wv, we);
    EdgeWin r;
    r.sep = wv;
    if (we == 0x7fffffff || wv == -FLT_MAX) {
        // nothing beat the reference's initial (-FLT_MAX, zero normal, 0, 0)
        r.normal = mk3(0.f, 0.f, 0.f); r.ea = 0; r.eb = 0;
        return r;
    }
    const int owner = we & 31;       // lane that evaluated pair `we`
    r.normal = shfl3(best_n, owner);
    r.ea = (int)MDR_LDG(edges_a + we / neb);
    r.eb = (int)MDR_LDG(edges_b + we % neb);
    return r;
}

// One queued hull-hull pair, one warp. doSAT + generateContacts for two hulls
// (narrowphase.cpp:678-758,1366-1513).
template <int kSlots>
__device__ __forceinline__ void hullHullWarp(WorldView &v, const StepParams &p, int pi, int lane,
                                             bool probing, int sub)
{
    const uint32_t pr = v.pairs[pi];
    const int a = pr & 1023, b = (pr >> 10) & 1023;     // both hulls: no reordering by type
    const BodyXf xa = bodyXf(v, p, a), xb = bodyXf(v, p, b);
    const ObjDev *oa = xa.o, *ob = xb.o;
    const float *src_va = p.a.verts + 3 * MDR_LDG(&oa->vert_off), *src_vb = p.a.verts + 3 * MDR_LDG(&ob->vert_off);
    const float *src_pa = p.a.planes + 4 * MDR_LDG(&oa->face_off), *src_pb = p.a.planes + 4 * MDR_LDG(&ob->face_off);
    const int nva = MDR_LDG(&oa->num_verts), nvb = MDR_LDG(&ob->num_verts);
    const int nfa = MDR_LDG(&oa->num_faces), nfb = MDR_LDG(&ob->num_faces);
    const int nea = MDR_LDG(&oa->num_edges), neb = MDR_LDG(&ob->num_edges);
    HullTopo ta, tb;
    ta.he = p.a.hedges + MDR_LDG(&oa->hedge_off); ta.polys = p.a.polys + MDR_LDG(&oa->face_off);
    tb.he = p.a.hedges + MDR_LDG(&ob->hedge_off); tb.polys = p.a.polys + MDR_LDG(&ob->face_off);
    const uint32_t *edges_a = p.a.edges + MDR_LDG(&oa->edge_off), *edges_b = p.a.edges + MDR_LDG(&ob->edge_off);

    const V3 zero = mk3(0.f, 0.f, 0.f);
    Plane none;
    none.n = zero; none.d = 0.f;
    // The first face query needs A's planes and B's vertices only; most pairs that reach a warp in a resting
    // scene (neighbours whose boxes overlap after a few degrees of yaw) end there, separated: the other
    // half of the world-space hulls is built only for those that go on.
    LaneV3<kSlots> va, vb;
    LanePlane<kSlots> pa, pb;
#pragma unroll
    for (int s = 0; s < kSlots; s++) {
        const int k = lane + 32 * s;
        vb.v[s] = k < nvb ? worldVert(xb, src_vb, k) : zero;
        pa.p[s] = k < nfa ? worldPlane(xa, src_pa, k) : none;
    }

    Manifold m;
    int ref = -1, alt = -1;
    bool hit = false;
    uint32_t hint = 0u;         // what the pair's thread tries first next time (satHintSeparates)
    do {
        float fa_sep;
        int fa_face;
        faceSeparationLane<kSlots>(pa, nfa, vb, nvb, lane, fa_sep, fa_face);
        warpArgMax(fa_sep, fa_face);
        if (fa_sep > 0.0f) { hint = 1u + (uint32_t)fa_face; break; }
#pragma unroll
        for (int s = 0; s < kSlots; s++) {
            const int k = lane + 32 * s;
            va.v[s] = k < nva ? worldVert(xa, src_va, k) : zero;
            pb.p[s] = k < nfb ? worldPlane(xb, src_pb, k) : none;
        }
        float fb_sep;
        int fb_face;
        faceSeparationLane<kSlots>(pb, nfb, va, nva, lane, fb_sep, fb_face);
        warpArgMax(fb_sep, fb_face);
        if (fb_sep > 0.0f) { hint = 1u + (uint32_t)kHintFaceB + (uint32_t)fb_face; break; }
        const EdgeWin eq = edgeQueryWarp<kSlots>(lane, va, pa, ta.he, edges_a, nea, xa.t, vb, pb, tb.he, edges_b, neb);
        if (eq.sep > 0.0f) break;
        const bool face_a = fa_sep > eq.sep, face_b = fb_sep > eq.sep;
        LaneVerts<kSlots> xva, xvb;
        xva.mine = va; xvb.mine = vb;
        if (face_a || face_b) {
            const bool a_ref = fa_sep >= fb_sep;
            const int rf = a_ref ? fa_face : fb_face;
            Plane rp;
            rp.n = a_ref ? laneGetNormal<kSlots>(pa, rf) : laneGetNormal<kSlots>(pb, rf);
            rp.d = a_ref ? laneGetD<kSlots>(pa, rf) : laneGetD<kSlots>(pb, rf);
            // incident face: most anti-parallel normal of the other hull, first
            // strict minimum (narrowphase.cpp:578-624) = arg-max of the negation
            const int nfo = a_ref ? nfb : nfa;
            float neg = -FLT_MAX;
            int inc = lane;
#pragma unroll
            for (int s = 0; s < kSlots; s++) {
                const int f = lane + 32 * s;
                const V3 on = a_ref ? pb.p[s].n : pa.p[s].n;
                const float val = f < nfo ? -dot(on, rp.n) : -FLT_MAX;
                if (s == 0 || val > neg) { neg = val; inc = f; }
            }
            warpArgMax(neg, inc);
            // all lanes, same inputs: vertices come by shuffle instead of being recomputed
            if (a_ref) faceContact(m, rp, rf, inc, ta, xva, tb, xvb);
            else faceContact(m, rp, rf, inc, tb, xvb, ta, xva);
            ref = a_ref ? a : b;
            alt = a_ref ? b : a;
        } else {
            edgeContact(m, eq.normal, eq.sep, eq.ea, eq.eb, ta, xva, tb, xvb);
            ref = a; alt = b;
        }
        hit = true;
    } while (false);
    if (lane == 0) v.hint[pi] = (uint8_t)hint;
    const bool emit = hit && m.n > 0;
    if (emit) {
        emitContactLanes(v, p, contactSlot(v, pi), m, ref, alt, lane);
        if (lane == 0 && v.level_any) {
            const int lv = pairLevel(pr);
            atomicOr(&v.level_any[sub & 1], 1u << (lv < 31 ? lv : 31));
        }
    }
    if (lane == 0 && probing) probeContact(p, sub, pi, emit ? &m : nullptr, ref, alt);
}

// The marked pairs of the world, in turn over the warps of its thread group.
// Every warp walks the same marks (nobody writes the pair words in this phase)
// and takes every nwarps-th one.
template <int kSlots>
__device__ __forceinline__ void satQueueWarp(const BlockCtx &c, WorldView &v, const StepParams &p,
                                             bool probing, int sub)
{
    const int np = v.misc[0];
    const int warp = c.tid() >> 5, nwarps = c.nthreads() >> 5, lane = c.tid() & 31;
    int turn = 0;
    for (int base = 0; base < np; base += 32) {
        const int i = base + lane;
        uint32_t marks = __ballot_sync(MDR_FULL, i < np && (v.pairs[i] & PAIR_SAT));
        while (marks) {
            const int bit = __ffs(marks) - 1;
            marks &= marks - 1;
            if (turn == warp) hullHullWarp<kSlots>(v, p, base + bit, lane, probing, sub);
            if (++turn == nwarps) turn = 0;
        }
    }
}

}  // namespace mdr
#endif
