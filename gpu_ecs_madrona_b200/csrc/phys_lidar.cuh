// Ray casts against the broadphase: BVH::traceRay (src/physics/broadphase.cpp:
// 653-718) with traceRayIntoLeaf / traceRayIntoPlane / traceRayIntoConvexPolyhedron
// (:720-856) and AABB::rayIntersects (include/madrona/math.hpp:1043-1069), the
// query behind CameraMode::Lidar observations (include/madrona/mw_render.hpp:9-13).
//
// The step keeps no node tree, only what the tree decides: the order in which
// the reference's LIFO walk meets the leaves and each leaf's grow-only slot
// box. That is enough for the same answers: an inner node's box contains the
// boxes below it (rebuild merges, refit only grows), the slab test is monotone
// in the box, so a leaf whose own slot box passes the test is reached by the
// reference and one whose box fails is not; pruning removes leaves from the
// walk without reordering the rest. A ray visits the leaves in rank order,
// shrinking t_max at every hit; on equal t the later leaf wins, as in the
// reference. (Exception: a NaN slab, origin exactly on a box face with a zero
// direction component, is not monotone; such rays may differ.)
#pragma once
#include "phys_world.cuh"

namespace mdr {

struct RayJob {
    const float *origins;      // [W][R][3]
    const float *dirs;         // [W][R][3]
    const float *t_max;        // [W][R] or nullptr (infinity)
    float *out_t;              // [W][R]   t_max when nothing is hit
    float *out_normal;         // [W][R][3] zeros when nothing is hit
    int32_t *out_entity;       // [W][R][2] (gen, id) or Entity::none() = all ones; may be nullptr
    int32_t *out_row;          // [W][R] row of the hit body or -1; may be nullptr
    int32_t rays_per_world;
};

// One leaf as the ray loop wants it, in rank order (20 words = 80 bytes).
struct RayLeaf {
    float lo[3], hi[3];        // slot box
    float pos[3];
    float rot[4];
    float scale[3];
    int32_t obj;               // object id | row << 16
    int32_t pad[3];
};

MDR_HD bool rayHitsBox(const float *lo, const float *hi, V3 o, V3 inv_d, float t_min, float t_max)
{
    const float tlx = inv_d.x * (lo[0] - o.x), tly = inv_d.y * (lo[1] - o.y), tlz = inv_d.z * (lo[2] - o.z);
    const float tux = inv_d.x * (hi[0] - o.x), tuy = inv_d.y * (hi[1] - o.y), tuz = inv_d.z * (hi[2] - o.z);
    const float box_min = fmaxf(fminf(tlx, tux), fmaxf(fminf(tly, tuy), fmaxf(fminf(tlz, tuz), t_min)));
    const float box_max = fminf(fmaxf(tlx, tux), fminf(fmaxf(tly, tuy), fminf(fmaxf(tlz, tuz), t_max)));
    return box_min <= box_max;
}

// traceRayIntoLeaf: the ray goes to the leaf's unscaled object space.
MDR_HD bool rayIntoLeaf(const AssetsDev &a, const RayLeaf &lf, V3 wo, V3 wd, float t_min, float t_max,
                        float *hit_t, V3 *hit_normal)
{
    const Q4 q = mk4(lf.rot[0], lf.rot[1], lf.rot[2], lf.rot[3]);
    const Q4 to_local = qinv(q);
    V3 o = rotate(to_local, wo - mk3(lf.pos[0], lf.pos[1], lf.pos[2]));
    o = mk3(o.x / lf.scale[0], o.y / lf.scale[1], o.z / lf.scale[2]);
    V3 d = rotate(to_local, wd);
    d = mk3(d.x / lf.scale[0], d.y / lf.scale[1], d.z / lf.scale[2]);
    const ObjDev *obj = a.objs + (lf.obj & 0xffff);
    V3 n = mk3(0.f, 0.f, 0.f);
    if (MDR_LDG(&obj->type) == 2) {
        // RTCD 5.3.8 on the hull's face planes (outward normals)
        const float *planes = a.planes + 4 * MDR_LDG(&obj->face_off);
        const int nf = MDR_LDG(&obj->num_faces);
        float tfirst = t_min, tlast = t_max;
        for (int f = 0; f < nf; f++) {
            const V3 pn = mk3(MDR_LDG(planes + 4 * f), MDR_LDG(planes + 4 * f + 1), MDR_LDG(planes + 4 * f + 2));
            const float pd = MDR_LDG(planes + 4 * f + 3);
            const float denom = dot(pn, d);
            const float neg_dist = pd - dot(pn, o);
            if (denom == 0.0f) {
                if (neg_dist < 0.0f) return false;
            } else {
                const float t = neg_dist / denom;
                if (denom < 0.0f) {
                    if (t >= tfirst) { tfirst = t; n = pn; }
                } else {
                    if (t <= tlast) tlast = t;
                }
                if (tfirst > tlast) return false;
            }
        }
        if (n.x == 0 && n.y == 0 && n.z == 0) return false;     // only back faces
        *hit_t = tfirst;
    } else {
        // plane z = 0 in its own space
        const float denom = d.z;
        if (denom == 0) return false;
        const float t = -o.z / denom;
        if (t < t_min || t > t_max) return false;
        *hit_t = t;
        n = mk3(0.f, 0.f, 1.f);
    }
    *hit_normal = rotate(q, n);
    return true;
}

// One ray against the world's leaves in rank order.
MDR_HD void traceRay(const AssetsDev &a, const RayLeaf *leaves, int n, V3 o, V3 d, float t_max,
                     float *out_t, V3 *out_normal, int *out_row)
{
    const V3 inv_d = mk3(1.f / d.x, 1.f / d.y, 1.f / d.z);
    int closest = -1;
    V3 closest_n = mk3(0.f, 0.f, 0.f);
    for (int r = 0; r < n; r++) {
        const RayLeaf &lf = leaves[r];
        if (!rayHitsBox(lf.lo, lf.hi, o, inv_d, 0.f, t_max)) continue;
        float hit_t;
        V3 hn;
        if (rayIntoLeaf(a, lf, o, d, 0.f, t_max, &hit_t, &hn)) {
            t_max = hit_t;
            closest = lf.obj >> 16;
            closest_n = hn;
        }
    }
    *out_t = t_max;
    *out_normal = closest_n;
    *out_row = closest;
}

#if defined(__CUDACC__)
// One block per world: the leaves go to SMEM in rank order once, then one ray
// per thread (looping when a world casts more rays than the block has threads).
__global__ void __launch_bounds__(128) physTraceRaysKernel(PhysTables t, AssetsDev a, RayJob job)
{
    extern __shared__ __align__(16) char ray_smem[];
    RayLeaf *leaves = reinterpret_cast<RayLeaf *>(ray_smem);
    const int w = blockIdx.x;
    const size_t r0 = (size_t)w * t.stride;
    int n = t.row_count[w];
    if (n > t.stride) n = t.stride;
    for (int r = threadIdx.x; r < n; r += blockDim.x) {
        const int leaf = t.order[r0 + r];
        const size_t i = r0 + leaf;
        RayLeaf lf;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            lf.lo[k] = t.slot_box[6 * i + k];
            lf.hi[k] = t.slot_box[6 * i + 3 + k];
            lf.pos[k] = t.pos[3 * i + k];
            lf.scale[k] = t.scale[3 * i + k];
        }
#pragma unroll
        for (int k = 0; k < 4; k++) lf.rot[k] = t.rot[4 * i + k];
        lf.obj = (t.obj[i] & 0xffff) | (leaf << 16);
        lf.pad[0] = lf.pad[1] = lf.pad[2] = 0;
        leaves[r] = lf;
    }
    __syncthreads();
    const int R = job.rays_per_world;
    for (int k = threadIdx.x; k < R; k += blockDim.x) {
        const size_t ri = (size_t)w * R + k;
        const V3 o = mk3(job.origins[3 * ri], job.origins[3 * ri + 1], job.origins[3 * ri + 2]);
        const V3 d = mk3(job.dirs[3 * ri], job.dirs[3 * ri + 1], job.dirs[3 * ri + 2]);
        const float t_max = job.t_max ? job.t_max[ri] : __int_as_float(0x7f800000);
        float hit_t;
        V3 hn;
        int row;
        traceRay(a, leaves, n, o, d, t_max, &hit_t, &hn, &row);
        job.out_t[ri] = hit_t;
        job.out_normal[3 * ri] = hn.x; job.out_normal[3 * ri + 1] = hn.y; job.out_normal[3 * ri + 2] = hn.z;
        if (job.out_row) job.out_row[ri] = row;
        if (job.out_entity) {
            job.out_entity[2 * ri] = row >= 0 ? t.entity[2 * (r0 + row)] : -1;
            job.out_entity[2 * ri + 1] = row >= 0 ? t.entity[2 * (r0 + row) + 1] : -1;
        }
    }
}
#endif

}  // namespace mdr
