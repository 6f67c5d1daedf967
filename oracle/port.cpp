// TEST INFRASTRUCTURE ONLY — never linked into or imported by the product path
// (gpu_ecs_madrona_b200/). Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may load oracle/libmadrona_port.so.
//
// Scalar CPU restatement ("port") of the reference's rigid-body step for one
// archetype of N bodies per world: broadphase (4-wide BVH, grow-only refit,
// LIFO traversal), narrowphase (hull-hull / hull-plane SAT, clipping, manifold
// reduction) and the XPBD solver, in the exact node order of the reference's
// task graph. Every function cites the reference file:line it restates.
// Parity is PINNED: tests/test_oracle_pin.py checks this library bit-for-bit
// against oracle/_ref/libmadrona_ref.so (the unmodified reference compiled by
// oracle/Makefile) and against tests/golden/*.npz generated from it.
//
// Floating point: plain IEEE fp32, no contraction (-ffp-contract=off), the
// reference's CPU expressions in the reference's association order.
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <map>
#include <thread>
#include <utility>
#include <vector>

#include "oracle_api.h"

namespace {

// ---------------------------------------------------------------- math ----
// include/madrona/math.hpp:195-476 (Vector3), :478-730 (Quat), :807-905
// (Mat3x3), :985-1128 (AABB). Association order is part of the contract.
struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };
struct Box { V3 lo, hi; };
struct M3 { V3 c[3]; };

inline V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V3 operator-(V3 a, V3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 operator-(V3 a) { return { -a.x, -a.y, -a.z }; }
inline V3 operator*(V3 a, float s) { return { a.x * s, a.y * s, a.z * s }; }
inline V3 operator*(float s, V3 a) { return { a.x * s, a.y * s, a.z * s }; }
// math.hpp:345-350: division multiplies by the reciprocal.
inline V3 operator/(V3 a, float s) { float inv = 1.f / s; return a * inv; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b)
{
    return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x };
}
inline float len2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
inline float len(V3 a) { return sqrtf(len2(a)); }
// math.hpp:239-247, CPU branch: 1.f / length().
inline V3 normalize(V3 a) { return a * (1.f / len(a)); }
inline V3 vmin(V3 a, V3 b) { return { fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z) }; }
inline V3 vmax(V3 a, V3 b) { return { fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z) }; }
inline V3 mulDiag(V3 d, V3 v) { return { d.x * v.x, d.y * v.y, d.z * v.z }; }
inline float &comp(V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
inline float comp(const V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

inline Q4 operator+(Q4 a, Q4 b) { return { a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z }; }
inline Q4 operator-(Q4 a, Q4 b) { return { a.w - b.w, a.x - b.x, a.y - b.y, a.z - b.z }; }
// math.hpp:707-715
inline Q4 operator*(Q4 a, Q4 b)
{
    return {
        (a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z),
        (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y),
        (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x),
        (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w),
    };
}
inline Q4 qinv(Q4 q) { return { q.w, -q.x, -q.y, -q.z }; }
// math.hpp:500-520
inline Q4 qnormalize(Q4 q)
{
    float inv = 1.f / sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    return { q.w * inv, q.x * inv, q.y * inv, q.z * inv };
}
// math.hpp:539-548
inline V3 rotate(Q4 q, V3 v)
{
    V3 pure { q.x, q.y, q.z };
    V3 pxv = cross(pure, v);
    V3 ppxv = cross(pure, pxv);
    return v + 2.f * ((pxv * q.w) + ppxv);
}
inline Q4 fromAngular(V3 v) { return { 0, v.x, v.y, v.z }; }

// math.hpp:810-834
inline M3 matFromQuat(Q4 r)
{
    float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    return {{
        { 1.f - 2.f * (y2 + z2), 2.f * (xy + wz), 2.f * (xz - wy) },
        { 2.f * (xy - wz), 1.f - 2.f * (x2 + z2), 2.f * (yz + wx) },
        { 2.f * (xz + wy), 2.f * (yz - wx), 1.f - 2.f * (x2 + y2) },
    }};
}
// math.hpp:835-866
inline M3 matFromRS(Q4 r, V3 s)
{
    float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    V3 ds { s.x * 2.f, s.y * 2.f, s.z * 2.f };
    return {{
        { s.x - ds.x * (y2 + z2), ds.x * (xy + wz), ds.x * (xz - wy) },
        { ds.y * (xy - wz), s.y - ds.y * (x2 + z2), ds.y * (yz + wx) },
        { ds.z * (xz + wy), ds.z * (yz - wx), s.z - ds.z * (x2 + y2) },
    }};
}
inline V3 matMul(const M3 &m, V3 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z; }
inline M3 matMulDiag(const M3 &m, V3 d) { return {{ m.c[0] * d.x, m.c[1] * d.y, m.c[2] * d.z }}; }

// math.hpp:999-1007 (strict)
inline bool overlaps(const Box &a, const Box &b)
{
    return a.lo.x < b.hi.x && b.lo.x < a.hi.x &&
           a.lo.y < b.hi.y && b.lo.y < a.hi.y &&
           a.lo.z < b.hi.z && b.lo.z < a.hi.z;
}
// math.hpp:1071-1103 (RTCD p.86)
inline Box applyTRS(const Box &b, V3 t, Q4 r, V3 s)
{
    M3 m = matFromRS(r, s);
    Box o;
    for (int i = 0; i < 3; i++) {
        comp(o.lo, i) = comp(o.hi, i) = comp(t, i);
        for (int j = 0; j < 3; j++) {
            float e = comp(m.c[j], i) * comp(b.lo, j);
            float f = comp(m.c[j], i) * comp(b.hi, j);
            if (e < f) { comp(o.lo, i) += e; comp(o.hi, i) += f; }
            else { comp(o.lo, i) += f; comp(o.hi, i) += e; }
        }
    }
    return o;
}

// ------------------------------------------------------------ geometry ----
struct HEdge { uint32_t next, twin, root, poly; };
struct Plane { V3 n; float d; };

// Half-edge mesh tables; construction restates
// src/physics/geometry.cpp:52-194 (HalfEdgeMesh::construct), including its
// quirks: a polygon's entry half-edge is the LAST one created for it, an
// "edge" is recorded (as the earlier twin's index) when the second twin appears.
struct Hull {
    std::vector<V3> verts;
    std::vector<Plane> planes;
    std::vector<HEdge> hedges;
    std::vector<uint32_t> edges;
    std::vector<uint32_t> polys;
};

void buildHull(Hull &h, const float *verts, int num_verts,
               const uint32_t *words, int num_words)
{
    h.verts.resize(num_verts);
    memcpy(h.verts.data(), verts, sizeof(V3) * num_verts);
    std::map<std::pair<uint32_t, uint32_t>, uint32_t> pair_to_hedge;
    int k = 0;
    uint32_t poly_idx = 0;
    while (k < num_words) {
        uint32_t n = words[k];
        const uint32_t *poly = words + k + 1;
        uint32_t first = (uint32_t)h.hedges.size();
        uint32_t entry = 0;
        int64_t prev = -1;
        for (uint32_t v = 0; v < n; v++) {
            uint32_t a = poly[v], b = poly[(v + 1) % n];
            if (pair_to_hedge.count({ a, b })) {
                fprintf(stderr, "port: invalid mesh for half-edge construction\n");
                abort();
            }
            uint32_t idx = (uint32_t)h.hedges.size();
            h.hedges.push_back({ 0, 0, a, poly_idx });
            auto twin = pair_to_hedge.find({ b, a });
            if (twin != pair_to_hedge.end()) {
                h.hedges[idx].twin = twin->second;
                h.hedges[twin->second].twin = idx;
                h.edges.push_back(twin->second);
            }
            if (prev >= 0) h.hedges[prev].next = idx;
            prev = idx;
            pair_to_hedge[{ a, b }] = idx;
            entry = idx;
        }
        h.hedges[prev].next = first;
        h.polys.push_back(entry);

        V3 pts[3];
        uint32_t he = entry;
        for (int i = 0; i < 3; i++) {
            pts[i] = h.verts[h.hedges[he].root];
            he = h.hedges[he].next;
        }
        V3 nrm = normalize(cross(pts[1] - pts[0], pts[2] - pts[0]));
        h.planes.push_back({ nrm, dot(nrm, pts[0]) });

        k += n + 1;
        poly_idx++;
    }
}

struct Object {
    int32_t type;       // 2 hull, 4 plane (physics.hpp:251-256)
    V3 inv_inertia;
    float inv_mass, mu_s, mu_d;
    Box aabb;
    Hull hull;
};

// ------------------------------------------------------------ world -------
struct Contact {
    int32_t ref, alt;
    float pts[4][4];
    int32_t num_points;
    V3 normal;
    float lambda_n[4];
};

// broadphase::BVH (include/madrona/physics.hpp:302-396)
struct BVHNode {
    float minX[4], minY[4], minZ[4], maxX[4], maxY[4], maxZ[4];
    int32_t children[4];
    int32_t parent;
};
constexpr int32_t kSentinel = (int32_t)0xFFFFFFFF;

struct World {
    int N;
    std::vector<V3> pos, scale, ext_force, ext_torque;
    std::vector<Q4> rot;
    std::vector<V3> vel_lin, vel_ang;
    std::vector<int32_t> obj;
    std::vector<uint32_t> response;
    std::vector<V3> prev_pos, pre_x, pre_v, pre_omega;
    std::vector<Q4> prev_rot, pre_q;

    std::vector<BVHNode> nodes;
    std::vector<Box> leaf_aabbs;
    std::vector<uint32_t> leaf_parents;
    std::vector<int32_t> sorted_leaves;
    bool force_rebuild = false;

    std::vector<std::pair<int32_t, int32_t>> candidates;
    std::vector<Contact> contacts;
    std::vector<orc_joint> joints;     // ConstraintData rows, creation order
    int64_t indeterminate = 0;
};

struct Sim {
    orc_config cfg;
    std::vector<Object> objs;
    std::vector<World> worlds;
    float h, vel_expand, accel_expand, restitution_threshold;
    V3 g;
};

// ---------------------------------------------------------- broadphase ----
// src/physics/broadphase.cpp:435-459
inline Box expandWithMotion(Box b, V3 v, float vf, float af)
{
    for (int i = 0; i < 3; i++) {
        float pos_delta = vf * comp(v, i);
        float min_delta = pos_delta - af;
        float max_delta = pos_delta + af;
        if (min_delta < 0.f) comp(b.lo, i) += min_delta;
        if (max_delta > 0.f) comp(b.hi, i) += max_delta;
    }
    return b;
}

// updateLeafPositionsEntry -> BVH::updateLeafPosition, broadphase.cpp:858-873,461-480
void updateLeafPositions(const Sim &s, World &w)
{
    for (int i = 0; i < w.N; i++) {
        Box world_box = applyTRS(s.objs[w.obj[i]].aabb, w.pos[i], w.rot[i], w.scale[i]);
        w.leaf_aabbs[i] = expandWithMotion(world_box, w.vel_lin[i],
                                           s.vel_expand, s.accel_expand);
        w.sorted_leaves[i] = i;
    }
}

inline int numBVHNodes(int num_leaves)
{
    // broadphase.cpp:33-40
    int a = (num_leaves - 1 + 2) / 3;
    return (a > 1 ? a : 1) + num_leaves;
}

// BVH::rebuild, broadphase.cpp:42-281: top-down, two levels of midpoint split
// per 4-wide node, explicit stack, DFS node numbering.
int midpointSplit(World &w, int base, int num)
{
    auto center = [&](int off) {
        const Box &b = w.leaf_aabbs[w.sorted_leaves[base + off]];
        return (b.lo + b.hi) / 2.f;
    };
    V3 cmin { FLT_MAX, FLT_MAX, FLT_MAX }, cmax { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (int i = 0; i < num; i++) {
        V3 c = center(i);
        cmin = vmin(cmin, c);
        cmax = vmax(cmax, c);
    }
    auto split = [&](int axis) {
        float split_val = 0.5f * (comp(cmin, axis) + comp(cmax, axis));
        int start = 0, end = num;
        while (start < end) {
            while (start < end && comp(center(start), axis) < split_val) ++start;
            while (start < end && comp(center(end - 1), axis) >= split_val) --end;
            if (start < end) {
                std::swap(w.sorted_leaves[base + start], w.sorted_leaves[base + end - 1]);
                ++start;
                --end;
            }
        }
        if (start > 0 && start < num) return start;
        return num / 2;
    };
    V3 d = cmax - cmin;
    if (d.x > d.y && d.x > d.z) return split(0);
    if (d.y > d.x && d.y > d.z) return split(1);
    return split(2);
}

void rebuildBVH(World &w)
{
    struct Entry { int32_t node, parent, offset, num; };
    Entry stack[128];
    stack[0] = { kSentinel, kSentinel, 0, w.N };
    int cur = 0, sz = 1;
    while (sz > 0) {
        Entry &e = stack[sz - 1];
        int32_t node_id;
        if (e.num <= 4) {
            node_id = cur++;
            BVHNode &n = w.nodes[node_id];
            n.parent = e.parent;
            for (int i = 0; i < 4; i++) {
                if (i < e.num) {
                    int32_t leaf = w.sorted_leaves[e.offset + i];
                    const Box &b = w.leaf_aabbs[leaf];
                    w.leaf_parents[leaf] = ((uint32_t)node_id << 2) | (uint32_t)i;
                    n.children[i] = (int32_t)(0x80000000u | (uint32_t)leaf);
                    n.minX[i] = b.lo.x; n.minY[i] = b.lo.y; n.minZ[i] = b.lo.z;
                    n.maxX[i] = b.hi.x; n.maxY[i] = b.hi.y; n.maxZ[i] = b.hi.z;
                } else {
                    n.children[i] = kSentinel;
                    n.minX[i] = n.minY[i] = n.minZ[i] = FLT_MAX;
                    n.maxX[i] = n.maxY[i] = n.maxZ[i] = -FLT_MAX;
                }
            }
        } else if (e.node == kSentinel) {
            node_id = cur++;
            e.node = node_id;
            BVHNode &n = w.nodes[node_id];
            for (int i = 0; i < 4; i++) n.children[i] = kSentinel;
            n.parent = e.parent;

            int second = midpointSplit(w, e.offset, e.num);
            int h1 = second, h2 = e.num - second;
            int first = midpointSplit(w, e.offset, h1);
            int third = midpointSplit(w, e.offset + second, h2);
            // copy: pushing may not alias e (the array does not move, but e is
            // read after the first push)
            Entry cur_e = e;
            stack[sz++] = { -1, cur_e.node, cur_e.offset + h1 + third, h2 - third };
            stack[sz++] = { -1, cur_e.node, cur_e.offset + h1, third };
            stack[sz++] = { -1, cur_e.node, cur_e.offset + first, h1 - first };
            stack[sz++] = { -1, cur_e.node, cur_e.offset, first };
            continue;
        } else {
            node_id = e.node;
        }
        sz -= 1;

        BVHNode &n = w.nodes[node_id];
        if (n.parent == -1) continue;

        V3 lo { FLT_MAX, FLT_MAX, FLT_MAX }, hi { -FLT_MAX, -FLT_MAX, -FLT_MAX };
        for (int i = 0; i < 4; i++) {
            if (n.children[i] == kSentinel) break;
            lo = vmin(lo, V3 { n.minX[i], n.minY[i], n.minZ[i] });
            hi = vmax(hi, V3 { n.maxX[i], n.maxY[i], n.maxZ[i] });
        }
        BVHNode &p = w.nodes[n.parent];
        int slot = 0;
        while (p.children[slot] != kSentinel) slot++;
        p.children[slot] = node_id;
        p.minX[slot] = lo.x; p.minY[slot] = lo.y; p.minZ[slot] = lo.z;
        p.maxX[slot] = hi.x; p.maxY[slot] = hi.y; p.maxZ[slot] = hi.z;
    }
}

// BVH::refitLeaf, broadphase.cpp:545-642 (grow-only; single-threaded here so
// the atomic min/max reduce to plain ones).
inline bool growSlot(BVHNode &n, int s, const Box &b)
{
    bool expanded = b.lo.x < n.minX[s] || b.lo.y < n.minY[s] || b.lo.z < n.minZ[s] ||
                    b.hi.x > n.maxX[s] || b.hi.y > n.maxY[s] || b.hi.z > n.maxZ[s];
    if (b.lo.x < n.minX[s]) n.minX[s] = b.lo.x;
    if (b.lo.y < n.minY[s]) n.minY[s] = b.lo.y;
    if (b.lo.z < n.minZ[s]) n.minZ[s] = b.lo.z;
    if (b.hi.x > n.maxX[s]) n.maxX[s] = b.hi.x;
    if (b.hi.y > n.maxY[s]) n.maxY[s] = b.hi.y;
    if (b.hi.z > n.maxZ[s]) n.maxZ[s] = b.hi.z;
    return expanded;
}

void refitLeaves(World &w)
{
    for (int leaf = 0; leaf < w.N; leaf++) {
        const Box &b = w.leaf_aabbs[leaf];
        uint32_t lp = w.leaf_parents[leaf];
        int32_t node_idx = (int32_t)(lp >> 2);
        if (!growSlot(w.nodes[node_idx], (int)(lp & 3), b)) continue;
        int32_t child = node_idx;
        node_idx = w.nodes[node_idx].parent;
        while (node_idx != kSentinel) {
            BVHNode &n = w.nodes[node_idx];
            int slot = -1;
            for (int j = 0; j < 4; j++) {
                if (n.children[j] == child) { slot = j; break; }
            }
            if (!growSlot(n, slot, b)) break;
            child = node_idx;
            node_idx = n.parent;
        }
    }
}

// findOverlappingEntry + BVH::findOverlaps, broadphase.cpp:897-932,
// include/madrona/physics.inl:61-100. Entity ids are creation-ordered, so
// `e.id < other.id` is `row < other_row`; leaf index == row.
void findOverlapping(World &w)
{
    for (int a = 0; a < w.N; a++) {
        bool a_static = w.response[a] == 2;
        const Box &q = w.leaf_aabbs[a];
        int32_t stack[128];
        stack[0] = 0;
        int sz = 1;
        while (sz > 0) {
            const BVHNode &n = w.nodes[stack[--sz]];
            for (int i = 0; i < 4; i++) {
                if (n.children[i] == kSentinel) continue;
                Box cb { { n.minX[i], n.minY[i], n.minZ[i] },
                         { n.maxX[i], n.maxY[i], n.maxZ[i] } };
                if (!overlaps(q, cb)) continue;
                if (n.children[i] & 0x80000000) {
                    int32_t b = n.children[i] & ~0x80000000;
                    if (a < b) {
                        if (a_static && w.response[b] == 2) continue;
                        w.candidates.push_back({ a, b });
                    }
                } else {
                    stack[sz++] = n.children[i];
                }
            }
        }
    }
}

// -------------------------------------------------------------- solver ----
// solver::substepRigidBodies, src/physics/physics.cpp:79-164
void substepBodies(const Sim &s, World &w)
{
    float h = s.h;
    for (int i = 0; i < w.N; i++) {
        V3 x = w.pos[i];
        Q4 q = w.rot[i];
        V3 v = w.vel_lin[i], omega = w.vel_ang[i];
        uint32_t rt = w.response[i];
        w.prev_pos[i] = x;
        w.prev_rot[i] = q;
        if (rt == 2) {
            w.pre_x[i] = x; w.pre_q[i] = q;
            w.pre_v[i] = { 0, 0, 0 }; w.pre_omega[i] = { 0, 0, 0 };
            continue;
        }
        const Object &o = s.objs[w.obj[i]];
        float inv_m = o.inv_mass;
        V3 inv_I = o.inv_inertia;
        if (rt == 0) v = v + h * s.g;
        v = v + (h * inv_m) * w.ext_force[i];
        x = x + h * v;
        V3 I { inv_I.x == 0 ? 0.0f : 1.0f / inv_I.x,
               inv_I.y == 0 ? 0.0f : 1.0f / inv_I.y,
               inv_I.z == 0 ? 0.0f : 1.0f / inv_I.z };
        Q4 to_local = qinv(q);
        V3 tau_local = rotate(to_local, w.ext_torque[i]);
        V3 omega_local = rotate(to_local, omega);
        V3 I_omega = mulDiag(I, omega_local);
        omega_local = omega_local + h * mulDiag(inv_I, tau_local - cross(omega_local, I_omega));
        omega = rotate(q, omega_local);
        Q4 apply = fromAngular((0.5f * h) * omega);
        q = q + apply * q;
        q = qnormalize(q);
        w.pos[i] = x; w.rot[i] = q;
        w.pre_x[i] = x; w.pre_q[i] = q;
        w.pre_v[i] = v; w.pre_omega[i] = omega;
    }
}

// physics.cpp:185-211
inline void applyPositional(V3 &x1, V3 &x2, Q4 &q1, Q4 &q2, V3 ra1, V3 ra2,
                            float inv_m1, float inv_m2, V3 n, float dl)
{
    x1 = x1 + (dl * inv_m1) * n;
    x2 = x2 - (dl * inv_m2) * n;
    float half = 0.5f * dl;
    V3 a1 = rotate(q1, half * ra1);
    V3 a2 = rotate(q2, half * ra2);
    q1 = q1 + fromAngular(a1) * q1;
    q2 = q2 - fromAngular(a2) * q2;
    q1 = qnormalize(q1);
    q2 = qnormalize(q2);
}

// physics.cpp:166-183
inline float positionalLambda(V3 ta1, V3 ta2, V3 ra1, V3 ra2,
                              float inv_m1, float inv_m2, float c, float alpha)
{
    float w1 = inv_m1 + dot(ta1, ra1);
    float w2 = inv_m2 + dot(ta2, ra2);
    return -c / (w1 + w2 + alpha);
}

// physics.cpp:213-245
inline float applyPositionalFull(V3 &x1, V3 &x2, Q4 &q1, Q4 &q2, V3 r1, V3 r2,
                                 float inv_m1, float inv_m2, V3 inv_I1, V3 inv_I2,
                                 V3 n, float c, float alpha)
{
    V3 n1 = rotate(qinv(q1), n), n2 = rotate(qinv(q2), n);
    V3 ta1 = cross(r1, n1), ta2 = cross(r2, n2);
    V3 ra1 = mulDiag(inv_I1, ta1), ra2 = mulDiag(inv_I2, ta2);
    float lambda = positionalLambda(ta1, ta2, ra1, ra2, inv_m1, inv_m2, c, alpha);
    applyPositional(x1, x2, q1, q2, ra1, ra2, inv_m1, inv_m2, n, lambda);
    return lambda;
}

// physics.cpp:365-382
inline void localContacts(V3 pre_x1, Q4 pre_q1, V3 pre_x2, Q4 pre_q2,
                          const Contact &c, int i, V3 &r1, V3 &r2)
{
    V3 c1 { c.pts[i][0], c.pts[i][1], c.pts[i][2] };
    float depth = c.pts[i][3];
    V3 c2 = c1 - c.normal * depth;
    r1 = rotate(qinv(pre_q1), c1 - pre_x1);
    r2 = rotate(qinv(pre_q2), c2 - pre_x2);
}

void handleJoint(const Sim &s, World &w, const orc_joint &j);

// handleContact + handleContactConstraint, physics.cpp:281-363,387-476
void solvePositions(const Sim &s, World &w)
{
    for (size_t ci = 0; ci < w.contacts.size(); ci++) {
        Contact c = w.contacts[ci];
        int a = c.ref, b = c.alt;
        V3 x1 = w.pos[a], x2 = w.pos[b];
        Q4 q1 = w.rot[a], q2 = w.rot[b];
        const Object &o1 = s.objs[w.obj[a]], &o2 = s.objs[w.obj[b]];
        float inv_m1 = o1.inv_mass, inv_m2 = o2.inv_mass;
        V3 inv_I1 = o1.inv_inertia, inv_I2 = o2.inv_inertia;
        if (w.response[a] == 2) { inv_m1 = 0.f; inv_I1 = { 0, 0, 0 }; }
        if (w.response[b] == 2) { inv_m2 = 0.f; inv_I2 = { 0, 0, 0 }; }
        float avg_mu_s = 0.5f * (o1.mu_s + o2.mu_s);

        for (int i = 0; i < 4; i++) {
            if (i >= c.num_points) continue;
            V3 r1, r2;
            localContacts(w.pre_x[a], w.pre_q[a], w.pre_x[b], w.pre_q[b], c, i, r1, r2);
            float lambda_n = 0.f;

            V3 p1 = rotate(q1, r1) + x1;
            V3 p2 = rotate(q2, r2) + x2;
            float d = dot(p1 - p2, c.normal);
            if (d > 0) {
                lambda_n = applyPositionalFull(x1, x2, q1, q2, r1, r2, inv_m1, inv_m2,
                                               inv_I1, inv_I2, c.normal, d, 0);
                V3 p1_hat = rotate(w.prev_rot[a], r1) + w.prev_pos[a];
                V3 p2_hat = rotate(w.prev_rot[b], r2) + w.prev_pos[b];
                p1 = rotate(q1, r1) + x1;
                p2 = rotate(q2, r2) + x2;
                V3 delta_p = (p1 - p1_hat) - (p2 - p2_hat);
                V3 delta_p_t = delta_p - dot(delta_p, c.normal) * c.normal;
                float tang = len(delta_p_t);
                if (tang > 0.f) {
                    V3 t = delta_p_t / tang;
                    V3 t1 = rotate(qinv(q1), t), t2 = rotate(qinv(q2), t);
                    V3 fta1 = cross(r1, t1), fta2 = cross(r2, t2);
                    V3 fra1 = mulDiag(inv_I1, fta1), fra2 = mulDiag(inv_I2, fta2);
                    float lambda_t = positionalLambda(fta1, fta2, fra1, fra2,
                                                      inv_m1, inv_m2, tang, 0);
                    float threshold = lambda_n * avg_mu_s;
                    if (lambda_t > threshold) {
                        applyPositional(x1, x2, q1, q2, fra1, fra2, inv_m1, inv_m2, t, lambda_t);
                    }
                }
            }
            w.contacts[ci].lambda_n[i] = lambda_n;
        }
        w.pos[a] = x1; w.pos[b] = x2;
        w.rot[a] = q1; w.rot[b] = q2;
    }
    // joints after the contacts, in collection (= table row) order, physics.cpp:663-668
    for (const orc_joint &j : w.joints) handleJoint(s, w, j);
}

// computeAngularUpdate + applyAngularUpdate, physics.cpp:247-279 (alpha_tilde = 0)
inline void angularUpdate(Q4 &q1, Q4 &q2, V3 inv_I1, V3 inv_I2, V3 n1, V3 n2, float theta)
{
    V3 ra1 = mulDiag(inv_I1, n1), ra2 = mulDiag(inv_I2, n2);
    float w1 = dot(n1, ra1), w2 = dot(n2, ra2);
    float dl = -theta / (w1 + w2 + 0.f);
    float half = 0.5f * dl;
    Q4 u1 = fromAngular(rotate(q1, half * ra1));
    Q4 u2 = fromAngular(rotate(q2, half * ra2));
    q1 = qnormalize(q1 + u1 * q1);
    q2 = qnormalize(q2 - u2 * q2);
}

// applyJointOrientationConstraint / applyJointAxisConstraint (physics.cpp:478-535)
// share everything after the error vector delta_q.
inline void angularConstraint(Q4 &q1, Q4 &q2, V3 delta_q, V3 inv_I1, V3 inv_I2)
{
    float mag = len(delta_q);
    if (mag > 0) {
        delta_q = delta_q / mag;
        V3 l1 = rotate(qinv(q1), delta_q), l2 = rotate(qinv(q2), delta_q);
        angularUpdate(q1, q2, inv_I1, inv_I2, l1, l2, mag);
    }
}

// handleJointConstraint, physics.cpp:537-648
void handleJoint(const Sim &s, World &w, const orc_joint &j)
{
    int a = j.row1, b = j.row2;
    V3 x1 = w.pos[a], x2 = w.pos[b];
    Q4 q1 = w.rot[a], q2 = w.rot[b];
    const Object &o1 = s.objs[w.obj[a]], &o2 = s.objs[w.obj[b]];
    float inv_m1 = o1.inv_mass, inv_m2 = o2.inv_mass;
    V3 inv_I1 = o1.inv_inertia, inv_I2 = o2.inv_inertia;
    if (w.response[a] == 2) { inv_m1 = 0.f; inv_I1 = { 0, 0, 0 }; }
    if (w.response[b] == 2) { inv_m2 = 0.f; inv_I2 = { 0, 0, 0 }; }
    V3 r1 { j.r1[0], j.r1[1], j.r1[2] }, r2 { j.r2[0], j.r2[1], j.r2[2] };
    const float *d = j.data;
    V3 corr;
    if (j.type == 0) {
        Q4 at1 { d[0], d[1], d[2], d[3] }, at2 { d[4], d[5], d[6], d[7] };
        float separation = d[8];
        Q4 o1q = qnormalize(q1 * at1), o2q = qnormalize(q2 * at2);
        Q4 diff = o1q * qinv(o2q);
        angularConstraint(q1, q2, 2.f * V3 { diff.x, diff.y, diff.z }, inv_I1, inv_I2);
        V3 r1w = rotate(q1, r1) + x1, r2w = rotate(q2, r2) + x2;
        V3 delta_r = r2w - r1w;
        Q4 axes = qnormalize(q1 * at1);
        V3 a1 = rotate(axes, V3 { 0, 1, 0 });      // math::fwd
        V3 b1 = rotate(axes, V3 { 1, 0, 0 });      // math::right
        V3 c1 = cross(a1, b1);
        corr = { 0, 0, 0 };
        corr = corr - (dot(delta_r, a1) - separation) * a1;
        corr = corr - dot(delta_r, b1) * b1;
        corr = corr - dot(delta_r, c1) * c1;
    } else {
        V3 a1l { d[0], d[1], d[2] }, a2l { d[3], d[4], d[5] };
        V3 ax1 = rotate(q1, a1l), ax2 = rotate(q2, a2l);
        angularConstraint(q1, q2, cross(ax1, ax2), inv_I1, inv_I2);
        V3 r1w = rotate(q1, r1) + x1, r2w = rotate(q2, r2) + x2;
        corr = r2w - r1w;
    }
    float mag = len(corr);
    if (mag > 0.f) {
        corr = corr / mag;
        applyPositionalFull(x1, x2, q1, q2, r1, r2, inv_m1, inv_m2, inv_I1, inv_I2, corr, mag, 0);
    }
    w.pos[a] = x1; w.pos[b] = x2;
    w.rot[a] = q1; w.rot[b] = q2;
}

// solver::setVelocities, physics.cpp:673-714
void setVelocities(const Sim &s, World &w)
{
    float h = s.h;
    for (int i = 0; i < w.N; i++) {
        Q4 q = w.rot[i], qp = w.prev_rot[i];
        Q4 dq;
        if (q.w != qp.w || q.x != qp.x || q.y != qp.y || q.z != qp.z) dq = q * qinv(qp);
        else dq = { 1, 0, 0, 0 };
        V3 new_omega = (2.f / h) * V3 { dq.x, dq.y, dq.z };
        w.vel_lin[i] = (w.pos[i] - w.prev_pos[i]) / h;
        w.vel_ang[i] = dq.w > 0.f ? new_omega : -new_omega;
    }
}

// physics.cpp:716-750
inline V3 relVel(V3 v1, V3 v2, V3 o1, V3 o2, V3 d1, V3 d2)
{
    return (v1 + cross(o1, d1)) - (v2 + cross(o2, d2));
}
inline void applyVelUpdate(V3 &v1, V3 &v2, V3 &o1, V3 &o2, Q4 q1, Q4 q2, V3 ta1, V3 ta2,
                           float inv_m1, float inv_m2, V3 inv_I1, V3 inv_I2,
                           V3 dv, float mag)
{
    V3 ra1 = mulDiag(inv_I1, ta1), ra2 = mulDiag(inv_I2, ta2);
    float w1 = inv_m1 + dot(ta1, ra1);
    float w2 = inv_m2 + dot(ta2, ra2);
    mag *= 1.f / (w1 + w2);
    v1 = v1 + (mag * inv_m1) * dv;
    v2 = v2 - (mag * inv_m2) * dv;
    o1 = o1 + rotate(q1, mag * ra1);
    o2 = o2 - rotate(q2, mag * ra2);
}

// solveVelocitiesForContact, physics.cpp:865-993 (+ :752-863)
void solveVelocities(const Sim &s, World &w)
{
    float h = s.h;
    for (size_t ci = 0; ci < w.contacts.size(); ci++) {
        const Contact &c = w.contacts[ci];
        int a = c.ref, b = c.alt;
        V3 v1 = w.vel_lin[a], o1 = w.vel_ang[a];
        V3 v2 = w.vel_lin[b], o2 = w.vel_ang[b];
        Q4 q1 = w.rot[a], q2 = w.rot[b];
        const Object &ob1 = s.objs[w.obj[a]], &ob2 = s.objs[w.obj[b]];
        float inv_m1 = ob1.inv_mass, inv_m2 = ob2.inv_mass;
        V3 inv_I1 = ob1.inv_inertia, inv_I2 = ob2.inv_inertia;
        if (w.response[a] == 2) { inv_m1 = 0.f; inv_I1 = { 0, 0, 0 }; }
        if (w.response[b] == 2) { inv_m2 = 0.f; inv_I2 = { 0, 0, 0 }; }
        float mu_d = 0.5f * (ob1.mu_d + ob2.mu_d);

        V3 r1l[4], r2l[4], r1w[4], r2w[4], rt1[4], rt2[4];
        float vn_bars[4];
        for (int i = 0; i < 4; i++) {
            if (i >= c.num_points) continue;
            V3 r1, r2;
            localContacts(w.pre_x[a], w.pre_q[a], w.pre_x[b], w.pre_q[b], c, i, r1, r2);
            V3 r1p = rotate(w.pre_q[a], r1), r2p = rotate(w.pre_q[b], r2);
            V3 v_bar = relVel(w.pre_v[a], w.pre_v[b], w.pre_omega[a], w.pre_omega[b], r1p, r2p);
            vn_bars[i] = dot(c.normal, v_bar);
            r1l[i] = r1; r2l[i] = r2;
            r1w[i] = rotate(q1, r1); r2w[i] = rotate(q2, r2);
            rt1[i] = cross(r1, rotate(qinv(q1), c.normal));
            rt2[i] = cross(r2, rotate(qinv(q2), c.normal));
        }
        for (int it = 0; it < 2; it++) {
            for (int i = 0; i < 4; i++) {
                if (i >= c.num_points) continue;
                V3 v = relVel(v1, v2, o1, o2, r1w[i], r2w[i]);
                float vn = dot(c.normal, v);
                float e = 0.3f;
                if (fabsf(vn_bars[i]) <= s.restitution_threshold) e = 0.f;
                float mag = fminf(-e * vn_bars[i], 0) - vn;
                applyVelUpdate(v1, v2, o1, o2, q1, q2, rt1[i], rt2[i],
                               inv_m1, inv_m2, inv_I1, inv_I2, c.normal, mag);
            }
        }
        for (int i = 0; i < 4; i++) {
            if (i >= c.num_points) continue;
            V3 v = relVel(v1, v2, o1, o2, r1w[i], r2w[i]);
            float fric = mu_d * fabsf(c.lambda_n[i]) / h;
            float vn = dot(c.normal, v);
            V3 vt = v - c.normal * vn;
            float vt_len = len(vt);
            if (vt_len != 0 && fric != 0.f) {
                float corrected = -fminf(fric, vt_len);
                V3 dw = vt / vt_len;
                V3 d1 = rotate(qinv(q1), dw), d2 = rotate(qinv(q2), dw);
                V3 fta1 = cross(r1l[i], d1), fta2 = cross(r2l[i], d2);
                applyVelUpdate(v1, v2, o1, o2, q1, q2, fta1, fta2,
                               inv_m1, inv_m2, inv_I1, inv_I2, dw, corrected);
            }
        }
        w.vel_lin[a] = v1; w.vel_ang[a] = o1;
        w.vel_lin[b] = v2; w.vel_ang[b] = o2;
    }
    w.contacts.clear();
}

// --------------------------------------------------------- narrowphase ----
struct HullState {
    const V3 *verts;
    const Plane *planes;
    const Hull *mesh;
    V3 center;
};

// makeHullState, src/physics/narrowphase.cpp:139-212
HullState makeHullState(const Hull &m, V3 t, Q4 r, V3 s, V3 *dst_v, Plane *dst_p)
{
    M3 rot = matFromQuat(r);
    M3 vtx = matMulDiag(rot, s);
    M3 nrm = matMulDiag(rot, V3 { 1.f / s.x, 1.f / s.y, 1.f / s.z });
    for (size_t i = 0; i < m.verts.size(); i++) dst_v[i] = matMul(vtx, m.verts[i]) + t;
    for (size_t i = 0; i < m.planes.size(); i++) {
        Plane p = m.planes[i];
        V3 origin = matMul(vtx, p.n * p.d) + t;
        V3 n = normalize(matMul(nrm, p.n));
        dst_p[i] = { n, dot(n, origin) };
    }
    return { dst_v, dst_p, &m, t };
}

// narrowphase.cpp:309-350
inline float hullDistanceFromPlane(const Plane &p, const HullState &h)
{
    float m = FLT_MAX;
    for (size_t i = 0; i < h.mesh->verts.size(); i++) {
        float d = dot(p.n, h.verts[i]);
        if (d < m) m = d;
    }
    return m - p.d;
}

struct FaceQuery { float sep; int face; Plane plane; };
// narrowphase.cpp:352-378
FaceQuery queryFaces(const HullState &a, const HullState &b)
{
    FaceQuery r { -FLT_MAX, -1, {} };
    for (size_t f = 0; f < a.mesh->planes.size(); f++) {
        Plane p = a.planes[f];
        float d = hullDistanceFromPlane(p, b);
        if (d > r.sep) {
            r.sep = d; r.face = (int)f; r.plane = p;
            if (r.sep > 0) break;
        }
    }
    return r;
}

// narrowphase.cpp:380-393
inline bool minkowskiFace(V3 a, V3 b, V3 c, V3 d)
{
    V3 bxa = cross(b, a), dxc = cross(d, c);
    float cba = dot(c, bxa), dba = dot(d, bxa), adc = dot(a, dxc), bdc = dot(b, dxc);
    return cba * dba < 0.0f && adc * bdc < 0.0f && cba * bdc > 0.0f;
}

struct EdgeQuery { float sep; V3 normal; int ea, eb; };
// queryEdgeDirections (CPU branch), narrowphase.cpp:474-576 with
// buildsMinkowskiFace :418-426 and edgeDistance :433-472
EdgeQuery queryEdges(const HullState &a, const HullState &b)
{
    EdgeQuery r { -FLT_MAX, { 0, 0, 0 }, 0, 0 };
    const Hull &ma = *a.mesh, &mb = *b.mesh;
    for (size_t ia = 0; ia < ma.edges.size(); ia++) {
        uint32_t hea = ma.edges[ia];
        const HEdge &ea = ma.hedges[hea];
        for (size_t ib = 0; ib < mb.edges.size(); ib++) {
            uint32_t heb = mb.edges[ib];
            const HEdge &eb = mb.hedges[heb];
            V3 an1 = a.planes[ea.poly].n, an2 = a.planes[ma.hedges[ea.twin].poly].n;
            V3 bn1 = b.planes[eb.poly].n, bn2 = b.planes[mb.hedges[eb.twin].poly].n;
            float sep = -FLT_MAX;
            V3 normal { 0, 0, 0 };
            if (minkowskiFace(an1, an2, -bn1, -bn2)) {
                V3 a1 = a.verts[ea.root], a2 = a.verts[ma.hedges[ea.next].root];
                V3 b1 = b.verts[eb.root], b2 = b.verts[mb.hedges[eb.next].root];
                V3 da = a2 - a1, db = b2 - b1;
                V3 cr = cross(da, db);
                float l2 = len2(cr);
                if (l2 != 0) {
                    float inv = 1.f / sqrtf(l2);
                    normal = cr * inv;
                    if (dot(normal, a1 - a.center) < 0.0f) normal = -normal;
                    sep = dot(normal, b1 - a1);
                }
            }
            if (sep > r.sep) {
                r.sep = sep; r.normal = normal; r.ea = (int)hea; r.eb = (int)heb;
                if (r.sep > 0) return r;
            }
        }
    }
    return r;
}

// narrowphase.cpp:578-624
inline int incidentFace(const HullState &h, V3 ref_n)
{
    float m = FLT_MAX;
    int best = -1;
    for (size_t f = 0; f < h.mesh->planes.size(); f++) {
        float d = dot(h.planes[f].n, ref_n);
        if (d < m) { m = d; best = (int)f; }
    }
    return best;
}

inline float planeDist(const Plane &p, V3 a) { return dot(a, p.n) - p.d; }

// clipPolygon, narrowphase.cpp:626-661; planeIntersection :246-250
int clipPolygon(V3 *dst, Plane cp, const V3 *in, int n_in)
{
    int n = 0;
    V3 v1 = in[n_in - 1];
    float d1 = planeDist(cp, v1);
    for (int i = 0; i < n_in; i++) {
        V3 v2 = in[i];
        float d2 = planeDist(cp, v2);
        if (d1 <= 0.0f && d2 <= 0.0f) {
            dst[n++] = v2;
        } else if (d1 <= 0.0f && d2 > 0.0f) {
            float dist = planeDist(cp, v1);
            dst[n++] = v1 + (v2 - v1) * (-dist / dot(cp.n, v2 - v1));
        } else if (d2 <= 0.0f && d1 > 0.0f) {
            float dist = planeDist(cp, v1);
            dst[n++] = v1 + (v2 - v1) * (-dist / dot(cp.n, v2 - v1));
            dst[n++] = v2;
        }
        v1 = v2;
        d1 = d2;
    }
    return n;
}

struct Manifold { V3 pts[4]; float depths[4]; int n; V3 normal; };

// buildFaceContactManifold, narrowphase.cpp:790-864. Reproduces the
// reference's reduction as written: `largestArea` is never updated, so point 2
// is the LAST candidate with positive signed area and point 3 the LAST with
// negative area. When no candidate qualifies the reference reads
// a slot of `Manifold manifold;` it never wrote (undefined behaviour: the
// stock build returns stale stack bytes). The slot is ZERO here, which is what
// the unmodified reference computes when its locals are defined:
// oracle/_ref/libmadrona_ref_defined.so = the same sources compiled with
// -ftrivial-auto-var-init=zero -fno-elide-constructors (without the second
// flag `manifold` is the function's return slot, which the first does not
// touch). `*indeterminate` is still bumped so tests against the STOCK build
// can exclude the world.
Manifold buildManifold(V3 n, V3 *contacts, float *depths, int num, int64_t *indeterminate)
{
    Manifold m;
    memset(&m, 0, sizeof(m));
    if (num <= 4) {
        m.n = num;
        for (int i = 0; i < num; i++) { m.pts[i] = contacts[i]; m.depths[i] = depths[i]; }
    } else {
        m.n = 4;
        m.pts[0] = contacts[0];
        m.depths[0] = depths[0];
        V3 p0 = m.pts[0];
        float largest_d2 = 0.0f;
        int far_idx = 0;
        for (int i = 1; i < num; i++) {
            float d2 = len2(p0 - contacts[i]);
            if (d2 > largest_d2) {
                largest_d2 = d2;
                m.pts[1] = contacts[i];
                m.depths[1] = depths[i];
                far_idx = i;
            }
        }
        contacts[far_idx] = m.pts[0];
        V3 diff0 = m.pts[1] - p0;
        const float largest_area = 0.0f;
        int area_idx = 0;
        bool found2 = false, found3 = false;
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area > largest_area) {
                m.pts[2] = contacts[i];
                m.depths[2] = depths[i];
                area_idx = i;
                found2 = true;
            }
        }
        contacts[area_idx] = m.pts[0];
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area < largest_area) {
                m.pts[3] = contacts[i];
                m.depths[3] = depths[i];
                found3 = true;
            }
        }
        if (!found2 || !found3) *indeterminate += 1;
    }
    const Q4 ident { 1, 0, 0, 0 };
    const V3 zero { 0, 0, 0 };
    for (int i = 0; i < m.n; i++) m.pts[i] = rotate(ident, m.pts[i]) + zero;
    m.normal = rotate(ident, n);
    return m;
}

inline int faceVerts(const Hull &m, const V3 *verts, int face, V3 *out)
{
    uint32_t he = m.polys[face], start = he;
    int n = 0;
    do {
        const HEdge &e = m.hedges[he];
        he = e.next;
        out[n++] = verts[e.root];
    } while (he != start);
    return n;
}

// createFaceContact, narrowphase.cpp:866-972
Manifold faceContact(Plane ref_plane, int ref_face, int inc_face,
                     const HullState &ref, const HullState &other, int64_t *indet)
{
    V3 buf_a[64], buf_b[64];
    V3 *in = buf_a, *dst = buf_b;
    int n = faceVerts(*other.mesh, other.verts, inc_face, in);

    const Hull &rm = *ref.mesh;
    uint32_t he = rm.polys[ref_face], start = he;
    V3 cur = ref.verts[rm.hedges[he].root];
    do {
        he = rm.hedges[he].next;
        V3 next = ref.verts[rm.hedges[he].root];
        V3 edge = next - cur;
        V3 pn = cross(edge, ref_plane.n);
        float d = dot(pn, cur);
        cur = next;
        n = clipPolygon(dst, Plane { pn, d }, in, n);
        std::swap(in, dst);
    } while (he != start);

    float depths[64];
    int below = 0;
    for (int i = 0; i < n; i++) {
        V3 v = in[i];
        float d = planeDist(ref_plane, v);
        if (d < 0.0f) {
            in[below] = v - d * ref_plane.n;
            depths[below] = -d;
            below++;
        }
    }
    return buildManifold(ref_plane.n, in, depths, below, indet);
}

// createFacePlaneContact, narrowphase.cpp:974-1017
Manifold facePlaneContact(Plane plane, int inc_face, const HullState &h, int64_t *indet)
{
    V3 pts[64];
    float depths[64];
    const Hull &m = *h.mesh;
    uint32_t he = m.polys[inc_face], start = he;
    int n = 0;
    do {
        const HEdge &e = m.hedges[he];
        he = e.next;
        V3 v = h.verts[e.root];
        float d = planeDist(plane, v);
        if (d < 0.0f) {
            pts[n] = v - d * plane.n;
            depths[n] = -d;
            n++;
        }
    } while (he != start);
    return buildManifold(plane.n, pts, depths, n, indet);
}

// shortestSegmentBetween + createEdgeContact, narrowphase.cpp:1020-1121
Manifold edgeContact(V3 normal, float sep, int ea, int eb,
                     const HullState &a, const HullState &b)
{
    const Hull &ma = *a.mesh, &mb = *b.mesh;
    V3 a1 = a.verts[ma.hedges[ea].root], a2 = a.verts[ma.hedges[ma.hedges[ea].next].root];
    V3 b1 = b.verts[mb.hedges[eb].root], b2 = b.verts[mb.hedges[mb.hedges[eb].next].root];
    V3 v1 = a2 - a1, v2 = b2 - b1, v21 = b1 - a1;
    float d22 = dot(v2, v2), d11 = dot(v1, v1), d21 = dot(v2, v1);
    float d211 = dot(v21, v1), d212 = dot(v21, v2);
    float denom = d21 * d21 - d22 * d11;
    float s, t;
    if (fabsf(denom) < 0.00001f) {
        s = 0.0f;
        t = (d11 * s - d211) / d21;
    } else {
        s = (d212 * d21 - d22 * d211) / denom;
        t = (-d211 * d21 + d11 * d212) / denom;
    }
    s = fmaxf(fminf(s, 1.0f), 0.0f);
    t = fmaxf(fminf(t, 1.0f), 0.0f);
    (void)t;
    V3 contact = a1 + s * v1;
    Manifold m;
    memset(&m, 0, sizeof(m));
    const Q4 ident { 1, 0, 0, 0 };
    const V3 zero { 0, 0, 0 };
    m.pts[0] = rotate(ident, contact) + zero;
    m.depths[0] = -sep;
    m.n = 1;
    m.normal = rotate(ident, normal);
    return m;
}

inline void pushContact(World &w, const Manifold &m, int ref, int alt)
{
    if (m.n <= 0) return;
    Contact c;
    memset(&c, 0, sizeof(c));
    c.ref = ref; c.alt = alt;
    for (int i = 0; i < m.n; i++) {
        c.pts[i][0] = m.pts[i].x; c.pts[i][1] = m.pts[i].y; c.pts[i][2] = m.pts[i].z;
        c.pts[i][3] = m.depths[i];
    }
    c.num_points = m.n;
    c.normal = m.normal;
    w.contacts.push_back(c);
}

// runNarrowphase (CPU branch) + narrowphaseDispatch + generateContacts,
// narrowphase.cpp:1515-1728,1182-1364,1366-1513; doSAT :678-758; doSATPlane :760-788
void runNarrowphase(const Sim &s, World &w)
{
    V3 vbuf[1024];
    Plane pbuf[1024];
    for (auto [ca, cb] : w.candidates) {
        int a = ca, b = cb;
        const Object *oa = &s.objs[w.obj[a]], *ob = &s.objs[w.obj[b]];
        if ((uint32_t)oa->type > (uint32_t)ob->type) { std::swap(a, b); std::swap(oa, ob); }
        Box wa = applyTRS(oa->aabb, w.pos[a], w.rot[a], w.scale[a]);
        Box wb = applyTRS(ob->aabb, w.pos[b], w.rot[b], w.scale[b]);
        if (!overlaps(wa, wb)) continue;
        int test = oa->type | ob->type;
        if (test == 2) {
            HullState ha = makeHullState(oa->hull, w.pos[a], w.rot[a], w.scale[a], vbuf, pbuf);
            HullState hb = makeHullState(ob->hull, w.pos[b], w.rot[b], w.scale[b],
                                         vbuf + oa->hull.verts.size(),
                                         pbuf + oa->hull.planes.size());
            FaceQuery fa = queryFaces(ha, hb);
            if (fa.sep > 0.0f) continue;
            FaceQuery fb = queryFaces(hb, ha);
            if (fb.sep > 0.0f) continue;
            EdgeQuery eq = queryEdges(ha, hb);
            if (eq.sep > 0.0f) continue;
            bool face_a = fa.sep > eq.sep, face_b = fb.sep > eq.sep;
            if (face_a || face_b) {
                bool a_ref = fa.sep >= fb.sep;
                Plane rp = a_ref ? fa.plane : fb.plane;
                int rf = a_ref ? fa.face : fb.face;
                const HullState &inc = a_ref ? hb : ha;
                int inc_face = incidentFace(inc, rp.n);
                if (a_ref) pushContact(w, faceContact(rp, rf, inc_face, ha, hb, &w.indeterminate), a, b);
                else pushContact(w, faceContact(rp, rf, inc_face, hb, ha, &w.indeterminate), b, a);
            } else {
                pushContact(w, edgeContact(eq.normal, eq.sep, eq.ea, eq.eb, ha, hb), a, b);
            }
        } else if (test == 6) {
            HullState ha = makeHullState(oa->hull, w.pos[a], w.rot[a], w.scale[a], vbuf, pbuf);
            V3 pn = rotate(w.rot[b], V3 { 0, 0, 1 });
            Plane plane { pn, dot(pn, w.pos[b]) };
            float sep = hullDistanceFromPlane(plane, ha);
            if (sep > 0.0f) continue;
            int inc_face = incidentFace(ha, plane.n);
            pushContact(w, facePlaneContact(plane, inc_face, ha, &w.indeterminate), b, a);
        } else {
            fprintf(stderr, "port: unsupported narrowphase test %d\n", test);
            abort();
        }
    }
}

// ---------------------------------------------------------------- step ----
// Node order of the reference graph: physics.cpp:1142-1205,
// broadphase.cpp:934-991, narrowphase.cpp:1768-1782.
void stepWorld(const Sim &s, World &w, orc_probe *probe)
{
    updateLeafPositions(s, w);
    if (w.force_rebuild) { w.force_rebuild = false; rebuildBVH(w); }
    refitLeaves(w);
    findOverlapping(w);
    if (probe) {
        probe->num_pairs = (int32_t)w.candidates.size();
        int np = probe->num_pairs < probe->max_pairs ? probe->num_pairs : probe->max_pairs;
        for (int i = 0; i < np; i++) {
            probe->pairs[2 * i] = w.candidates[i].first;
            probe->pairs[2 * i + 1] = w.candidates[i].second;
        }
    }
    for (int sub = 0; sub < s.cfg.num_substeps; sub++) {
        substepBodies(s, w);
        runNarrowphase(s, w);
        if (probe) {
            probe->num_contacts[sub] = (int32_t)w.contacts.size();
            for (size_t i = 0; i < w.contacts.size(); i++) {
                const Contact &c = w.contacts[i];
                orc_contact &o = probe->contacts[(size_t)sub * s.cfg.max_contacts + i];
                memset(&o, 0, sizeof(o));
                o.ref_row = c.ref; o.alt_row = c.alt;
                memcpy(o.points, c.pts, sizeof(o.points));
                o.num_points = c.num_points;
                memcpy(o.normal, &c.normal, 12);
            }
        }
        solvePositions(s, w);
        setVelocities(s, w);
        solveVelocities(s, w);
    }
    w.candidates.clear();
    updateLeafPositions(s, w);
    refitLeaves(w);
}

}  // namespace

extern "C" {

const char *orc_kind(void) { return "port"; }

void *orc_create(const orc_config *cfg, const orc_object *objs, int32_t num_objs,
                 const float *verts, const uint32_t *poly_words,
                 const float *pos, const float *rot, const float *scale,
                 const float *vel, const int32_t *obj_ids, const uint32_t *response,
                 const float *ext_force, const float *ext_torque)
{
    Sim *s = new Sim;
    s->cfg = *cfg;
    // SolverData ctor, physics.cpp:15-32; BVH expansion factors :1023-1027
    s->h = cfg->delta_t / (float)cfg->num_substeps;
    s->g = { cfg->gravity[0], cfg->gravity[1], cfg->gravity[2] };
    s->restitution_threshold = 2.f * len(s->g) * s->h;
    s->vel_expand = 2.f * cfg->delta_t;
    s->accel_expand = 100.f * cfg->delta_t * cfg->delta_t;
    s->objs.resize(num_objs);
    for (int i = 0; i < num_objs; i++) {
        const orc_object &o = objs[i];
        Object &d = s->objs[i];
        d.type = o.type;
        d.inv_inertia = { o.inv_inertia[0], o.inv_inertia[1], o.inv_inertia[2] };
        d.inv_mass = o.inv_mass; d.mu_s = o.mu_s; d.mu_d = o.mu_d;
        d.aabb = { { o.aabb_min[0], o.aabb_min[1], o.aabb_min[2] },
                   { o.aabb_max[0], o.aabb_max[1], o.aabb_max[2] } };
        if (o.type == 2) {
            buildHull(d.hull, verts + 3 * o.vert_offset, o.num_verts,
                      poly_words + o.poly_offset, o.num_poly_words);
        }
    }
    const int W = cfg->num_worlds, N = cfg->bodies_per_world;
    s->worlds.resize(W);
    for (int wi = 0; wi < W; wi++) {
        World &w = s->worlds[wi];
        w.N = N;
        size_t b0 = (size_t)wi * N;
        auto v3 = [&](const float *p, std::vector<V3> &dst) {
            dst.resize(N);
            memcpy(dst.data(), p + 3 * b0, sizeof(V3) * N);
        };
        v3(pos, w.pos); v3(scale, w.scale); v3(ext_force, w.ext_force); v3(ext_torque, w.ext_torque);
        w.rot.resize(N);
        memcpy(w.rot.data(), rot + 4 * b0, sizeof(Q4) * N);
        w.vel_lin.resize(N); w.vel_ang.resize(N);
        for (int i = 0; i < N; i++) {
            const float *v = vel + 6 * (b0 + i);
            w.vel_lin[i] = { v[0], v[1], v[2] };
            w.vel_ang[i] = { v[3], v[4], v[5] };
        }
        w.obj.assign(obj_ids + b0, obj_ids + b0 + N);
        w.response.assign(response + b0, response + b0 + N);
        w.prev_pos = w.pos; w.prev_rot = w.rot;
        w.pre_x = w.pos; w.pre_q = w.rot;
        w.pre_v.assign(N, { 0, 0, 0 }); w.pre_omega.assign(N, { 0, 0, 0 });
        w.nodes.resize(numBVHNodes(N));
        w.leaf_aabbs.resize(N); w.leaf_parents.resize(N); w.sorted_leaves.resize(N);
        w.force_rebuild = true;
    }
    return s;
}

void orc_destroy(void *h) { delete (Sim *)h; }

void orc_flag_rebuild(void *h)
{
    for (World &w : ((Sim *)h)->worlds) w.force_rebuild = true;
}

void orc_flag_rebuild_world(void *h, int32_t world) { ((Sim *)h)->worlds[world].force_rebuild = true; }

void orc_set_joints(void *h, int32_t max_joints, const int32_t *counts, const orc_joint *joints)
{
    Sim *s = (Sim *)h;
    for (size_t wi = 0; wi < s->worlds.size(); wi++) {
        const orc_joint *src = joints + wi * (size_t)max_joints;
        s->worlds[wi].joints.assign(src, src + counts[wi]);
    }
}

void orc_step(void *hp, int32_t num_steps, int32_t num_threads)
{
    Sim *s = (Sim *)hp;
    const int W = (int)s->worlds.size();
    if (num_threads <= 1) {
        for (int st = 0; st < num_steps; st++)
            for (int i = 0; i < W; i++) stepWorld(*s, s->worlds[i], nullptr);
        return;
    }
    // Same schedule as the reference harness (and src/mw/cpu_exec.cpp:244-284):
    // workers pull world indices from an atomic counter.
    std::atomic<int> next { 0 };
    std::vector<std::thread> pool;
    for (int t = 0; t < num_threads; t++) {
        pool.emplace_back([&]() {
            while (true) {
                int i = next.fetch_add(1, std::memory_order_relaxed);
                if (i >= W) break;
                for (int st = 0; st < num_steps; st++) stepWorld(*s, s->worlds[i], nullptr);
            }
        });
    }
    for (auto &t : pool) t.join();
}

static void *colPtr(World &w, int col, int row, char *scratch)
{
    switch (col) {
    case ORC_COL_POSITION: return &w.pos[row];
    case ORC_COL_ROTATION: return &w.rot[row];
    case ORC_COL_SCALE: return &w.scale[row];
    case ORC_COL_OBJECT_ID: return &w.obj[row];
    case ORC_COL_RESPONSE_TYPE: return &w.response[row];
    case ORC_COL_EXT_FORCE: return &w.ext_force[row];
    case ORC_COL_EXT_TORQUE: return &w.ext_torque[row];
    default: return scratch;
    }
}

void orc_read(void *hp, int32_t col, void *out)
{
    Sim *s = (Sim *)hp;
    const int N = s->cfg.bodies_per_world, nb = orc_col_bytes[col];
    char *o = (char *)out;
    for (size_t wi = 0; wi < s->worlds.size(); wi++) {
        World &w = s->worlds[wi];
        for (int i = 0; i < N; i++) {
            char tmp[32] = {};
            char *dst = o + (wi * N + i) * nb;
            switch (col) {
            case ORC_COL_ENTITY: { int32_t e[2] = { 0, i }; memcpy(dst, e, 8); } break;
            case ORC_COL_VELOCITY: memcpy(dst, &w.vel_lin[i], 12); memcpy(dst + 12, &w.vel_ang[i], 12); break;
            case ORC_COL_SUBSTEP_PREV: memcpy(dst, &w.prev_pos[i], 12); memcpy(dst + 12, &w.prev_rot[i], 16); break;
            case ORC_COL_PRESOLVE_POS: memcpy(dst, &w.pre_x[i], 12); memcpy(dst + 12, &w.pre_q[i], 16); break;
            case ORC_COL_PRESOLVE_VEL: memcpy(dst, &w.pre_v[i], 12); memcpy(dst + 12, &w.pre_omega[i], 12); break;
            case ORC_COL_LEAF_ID: { int32_t l = i; memcpy(dst, &l, 4); } break;
            default: memcpy(dst, colPtr(w, col, i, tmp), nb);
            }
        }
    }
}

void orc_write(void *hp, int32_t col, const void *in)
{
    Sim *s = (Sim *)hp;
    const int N = s->cfg.bodies_per_world, nb = orc_col_bytes[col];
    const char *o = (const char *)in;
    for (size_t wi = 0; wi < s->worlds.size(); wi++) {
        World &w = s->worlds[wi];
        for (int i = 0; i < N; i++) {
            const char *src = o + (wi * N + i) * nb;
            char tmp[32];
            switch (col) {
            case ORC_COL_ENTITY: case ORC_COL_LEAF_ID: break;
            case ORC_COL_VELOCITY: memcpy(&w.vel_lin[i], src, 12); memcpy(&w.vel_ang[i], src + 12, 12); break;
            case ORC_COL_SUBSTEP_PREV: memcpy(&w.prev_pos[i], src, 12); memcpy(&w.prev_rot[i], src + 12, 16); break;
            case ORC_COL_PRESOLVE_POS: memcpy(&w.pre_x[i], src, 12); memcpy(&w.pre_q[i], src + 12, 16); break;
            case ORC_COL_PRESOLVE_VEL: memcpy(&w.pre_v[i], src, 12); memcpy(&w.pre_omega[i], src + 12, 12); break;
            default: memcpy(colPtr(w, col, i, tmp), src, nb);
            }
        }
    }
}

int32_t orc_probe_step(void *hp, int32_t world, orc_probe *out)
{
    Sim *s = (Sim *)hp;
    World &w = s->worlds[world];
    stepWorld(*s, w, out);
    for (int i = 0; i < w.N; i++) {
        uint32_t lp = w.leaf_parents[i];
        const BVHNode &n = w.nodes[lp >> 2];
        int sl = lp & 3;
        float *sb = out->slot_boxes + 6 * i;
        sb[0] = n.minX[sl]; sb[1] = n.minY[sl]; sb[2] = n.minZ[sl];
        sb[3] = n.maxX[sl]; sb[4] = n.maxY[sl]; sb[5] = n.maxZ[sl];
        memcpy(out->leaf_aabbs + 6 * i, &w.leaf_aabbs[i], 24);
    }
    int32_t stack[128];
    int sz = 1, rank = 0;
    stack[0] = 0;
    while (sz > 0) {
        const BVHNode &n = w.nodes[stack[--sz]];
        for (int i = 0; i < 4; i++) {
            if (n.children[i] == kSentinel) continue;
            if (n.children[i] & 0x80000000) out->ref_rank[n.children[i] & ~0x80000000] = rank++;
            else stack[sz++] = n.children[i];
        }
    }
    return out->num_pairs > out->max_pairs ? -1 : 0;
}

// ------------------------------------------------------------- traceRay ---
// AABB::rayIntersects, include/madrona/math.hpp:1043-1069
static bool rayHitsBox(const Box &b, V3 o, V3 inv_d, float t_min, float t_max)
{
    V3 tl = { inv_d.x * (b.lo.x - o.x), inv_d.y * (b.lo.y - o.y), inv_d.z * (b.lo.z - o.z) };
    V3 tu = { inv_d.x * (b.hi.x - o.x), inv_d.y * (b.hi.y - o.y), inv_d.z * (b.hi.z - o.z) };
    float mn[4] = { fminf(tl.x, tu.x), fminf(tl.y, tu.y), fminf(tl.z, tu.z), t_min };
    float mx[4] = { fmaxf(tl.x, tu.x), fmaxf(tl.y, tu.y), fmaxf(tl.z, tu.z), t_max };
    float box_min = fmaxf(mn[0], fmaxf(mn[1], fmaxf(mn[2], mn[3])));
    float box_max = fminf(mx[0], fminf(mx[1], fminf(mx[2], mx[3])));
    return box_min <= box_max;
}

// BVH::traceRayIntoLeaf + traceRayIntoPlane + traceRayIntoConvexPolyhedron,
// src/physics/broadphase.cpp:720-856. The leaf transform is the pose the last
// updateLeafPosition saw (= the current one after a step).
static bool rayIntoLeaf(const Sim &s, const World &w, int leaf, V3 wo, V3 wd, float t_min, float t_max,
                        float *hit_t, V3 *hit_normal)
{
    const Object &obj = s.objs[w.obj[leaf]];
    const Q4 to_local = qinv(w.rot[leaf]);
    const V3 sc = w.scale[leaf];
    V3 o = rotate(to_local, wo - w.pos[leaf]);
    o = { o.x / sc.x, o.y / sc.y, o.z / sc.z };
    V3 d = rotate(qinv(w.rot[leaf]), wd);
    d = { d.x / sc.x, d.y / sc.y, d.z / sc.z };
    V3 n = { 0.f, 0.f, 0.f };
    bool hit;
    if (obj.type == 2) {
        float tfirst = t_min, tlast = t_max;
        hit = true;
        for (size_t f = 0; f < obj.hull.planes.size() && hit; f++) {
            const Plane &pl = obj.hull.planes[f];
            float denom = dot(pl.n, d);
            float neg_dist = pl.d - dot(pl.n, o);
            if (denom == 0.0f) {
                if (neg_dist < 0.0f) hit = false;
            } else {
                float t = neg_dist / denom;
                if (denom < 0.0f) {
                    if (t >= tfirst) { tfirst = t; n = pl.n; }
                } else {
                    if (t <= tlast) tlast = t;
                }
                if (tfirst > tlast) hit = false;
            }
        }
        if (hit && n.x == 0 && n.y == 0 && n.z == 0) hit = false;
        if (hit) *hit_t = tfirst;
    } else {
        float denom = d.z;
        hit = denom != 0;
        if (hit) {
            float t = -o.z / denom;
            hit = !(t < t_min || t > t_max);
            if (hit) { *hit_t = t; n = { 0.f, 0.f, 1.f }; }
        }
    }
    // the reference rotates whatever normal the primitive test left, hit or not
    *hit_normal = rotate(w.rot[leaf], n);
    return hit;
}

// BVH::traceRay, src/physics/broadphase.cpp:653-718
void orc_trace_rays(void *hp, int32_t world, int32_t n, const float *origins, const float *dirs,
                    const float *t_maxs, float *out_t, float *out_normal, int32_t *out_row)
{
    Sim *s = (Sim *)hp;
    const World &w = s->worlds[world];
    for (int r = 0; r < n; r++) {
        const V3 o = { origins[3*r], origins[3*r+1], origins[3*r+2] };
        const V3 d = { dirs[3*r], dirs[3*r+1], dirs[3*r+2] };
        const V3 inv_d = { 1.f / d.x, 1.f / d.y, 1.f / d.z };
        float t_max = t_maxs[r];
        int32_t closest = -1;
        V3 closest_n = { 0.f, 0.f, 0.f };
        int32_t stack[128];
        stack[0] = 0;
        int sz = 1;
        while (sz > 0) {
            const BVHNode &nd = w.nodes[stack[--sz]];
            for (int i = 0; i < 4; i++) {
                if (nd.children[i] == kSentinel) continue;
                Box cb { { nd.minX[i], nd.minY[i], nd.minZ[i] }, { nd.maxX[i], nd.maxY[i], nd.maxZ[i] } };
                if (!rayHitsBox(cb, o, inv_d, 0.f, t_max)) continue;
                if (nd.children[i] & 0x80000000) {
                    const int32_t leaf = nd.children[i] & ~0x80000000;
                    float hit_t;
                    V3 hn;
                    if (rayIntoLeaf(*s, w, leaf, o, d, 0.f, t_max, &hit_t, &hn)) {
                        t_max = hit_t;
                        closest = leaf;
                        closest_n = hn;
                    }
                } else {
                    stack[sz++] = nd.children[i];
                }
            }
        }
        out_t[r] = t_max;
        out_row[r] = closest;
        out_normal[3*r] = closest_n.x; out_normal[3*r+1] = closest_n.y; out_normal[3*r+2] = closest_n.z;
    }
}

// instanceTransformSetup / updateViewData, src/mw/render/interop.cpp:10-96;
// Mat3x4::fromTRS, include/madrona/math.hpp:929-962
void orc_feeder(void *hp, const float *world_offset, float *out_transforms, int32_t V,
                const int32_t *view_rows, const float *camera_offset, const float *y_scale, float *out_views)
{
    Sim *s = (Sim *)hp;
    const V3 off = { world_offset[0], world_offset[1], world_offset[2] };
    for (size_t wi = 0; wi < s->worlds.size(); wi++) {
        const World &w = s->worlds[wi];
        for (int i = 0; i < w.N; i++) {
            const V3 t = w.pos[i] + off, sc = w.scale[i];
            const Q4 r = w.rot[i];
            float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
            float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
            float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
            V3 ds = { 2.f * sc.x, 2.f * sc.y, 2.f * sc.z };
            V3 c0 = { sc.x - ds.x * (y2 + z2), ds.x * (xy + wz), ds.x * (xz - wy) };
            V3 c1 = { ds.y * (xy - wz), sc.y - ds.y * (x2 + z2), ds.y * (yz + wx) };
            V3 c2 = { ds.z * (xz + wy), ds.z * (yz - wx), sc.z - ds.z * (x2 + y2) };
            float *m = out_transforms + 12 * (wi * w.N + i);
            m[0] = c0.x; m[1] = c1.x; m[2] = c2.x; m[3] = t.x;
            m[4] = c0.y; m[5] = c1.y; m[6] = c2.y; m[7] = t.y;
            m[8] = c0.z; m[9] = c1.z; m[10] = c2.z; m[11] = t.z;
        }
        for (int k = 0; k < V; k++) {
            const size_t vi = wi * V + k;
            const int row = view_rows[vi];
            const V3 cam = w.pos[row] + V3 { camera_offset[3*vi], camera_offset[3*vi+1], camera_offset[3*vi+2] } + off;
            float *o = out_views + 8 * vi;
            o[0] = w.rot[row].w; o[1] = w.rot[row].x; o[2] = w.rot[row].y; o[3] = w.rot[row].z;
            o[4] = cam.x; o[5] = cam.y; o[6] = cam.z; o[7] = y_scale[vi];
        }
    }
}

void orc_indeterminate(void *hp, int64_t *out)
{
    Sim *s = (Sim *)hp;
    for (size_t i = 0; i < s->worlds.size(); i++) out[i] = s->worlds[i].indeterminate;
}

int32_t orc_get_hull(void *hp, int32_t obj, uint32_t *half_edges, uint32_t *edges,
                     uint32_t *polys, float *planes, int32_t *counts)
{
    Sim *s = (Sim *)hp;
    const Hull &m = s->objs[obj].hull;
    counts[0] = (int32_t)m.hedges.size(); counts[1] = (int32_t)m.edges.size();
    counts[2] = (int32_t)m.polys.size(); counts[3] = (int32_t)m.verts.size();
    memcpy(half_edges, m.hedges.data(), 16 * m.hedges.size());
    memcpy(edges, m.edges.data(), 4 * m.edges.size());
    memcpy(polys, m.polys.data(), 4 * m.polys.size());
    memcpy(planes, m.planes.data(), 16 * m.planes.size());
    return 0;
}

}
