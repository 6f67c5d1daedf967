// TEST INFRASTRUCTURE ONLY. Host emulation of the CUDA step kernel: compiles the
// SAME device source (gpu_ecs_madrona_b200/csrc/phys_*.cuh) with g++ as a
// 1-thread "block" per world, over malloc'd tables, behind the mdr_phys_* names
// of include/madrona_b200.h. It exists so that the algorithmic restructuring of
// the step (brute-force rank-ordered broadphase, level-scheduled solver,
// per-body hull transforms) can be checked bit-for-bit against the oracle on a
// machine without a GPU (`pytest -m "not gpu"`). It is built into
// tests/emu/_build/, is never loaded by the package, bench.py or smoke(), and
// is not a fallback: gpu_ecs_madrona_b200 refuses to run without its CUDA library.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../gpu_ecs_madrona_b200/csrc/phys_step.cuh"
#include "../../gpu_ecs_madrona_b200/csrc/phys_lidar.cuh"
#include "../../gpu_ecs_madrona_b200/csrc/phys_feeder.cuh"
#include "../../gpu_ecs_madrona_b200/csrc/phys_assets.hpp"
#include "../../gpu_ecs_madrona_b200/csrc/phys_cols.hpp"

using namespace mdr;

struct mdr_phys {
    mdr_phys_config cfg;
    HostAssets assets;
    StepParams params;
    std::vector<char> smem;
    std::vector<SolverContact> overflow;
    std::vector<void *> allocs;
    int32_t err = 0;
    int stride = 0;
    template <typename T> T *alloc(size_t n)
    {
        void *p = calloc(n ? n : 1, sizeof(T));
        allocs.push_back(p);
        return (T *)p;
    }
};

extern "C" {

const char *mdr_version(void) { return "emu"; }
int32_t mdr_device_count(void) { return 0; }
int32_t mdr_phys_column_bytes(int32_t col) { return col >= 0 && col < MDR_NUM_PHYS_COLS ? kColBytes[col] : 0; }

int32_t mdr_phys_create(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                        const float *verts, const uint32_t *poly_words, mdr_phys **out)
{
    if (!cfg || !out || cfg->num_worlds <= 0 || cfg->max_bodies_per_world <= 0 ||
        cfg->max_bodies_per_world > MDR_MAX_BODIES || cfg->num_substeps <= 0 || cfg->max_contacts <= 0)
        return MDR_EINVAL;
    mdr_phys *p = new mdr_phys;
    p->cfg = *cfg;
    if (!buildAssets(p->assets, objs, num_objs, verts, poly_words)) { delete p; return MDR_EINVAL; }
    memset(&p->params, 0, sizeof(p->params));
    StepParams &sp = p->params;
    fillStepConstants(sp, *cfg);
    const int W = cfg->num_worlds;
    const int stride = (cfg->max_bodies_per_world + 3) / 4 * 4;
    p->stride = stride;
    const size_t R = (size_t)W * stride;
    PhysTables &t = sp.t;
    t.stride = stride; t.num_worlds = W;
    for (int c = 0; c < MDR_NUM_PHYS_COLS; c++) *colSlot(t, c) = p->alloc<char>(R * kColBytes[c]);
    t.row_count = p->alloc<int32_t>(W);
    t.leaf_aabb = p->alloc<float>(R * 6);
    t.slot_box = p->alloc<float>(R * 6);
    t.order = p->alloc<uint16_t>(R);
    t.sorted = p->alloc<uint16_t>(R);
    t.rebuild_flag = p->alloc<uint8_t>(W);
    sp.a.objs = p->assets.objs.data();
    sp.a.verts = p->assets.verts.data();
    sp.a.planes = p->assets.planes.data();
    sp.a.hedges = p->assets.hedges.data();
    sp.a.edges = p->assets.edges.data();
    sp.a.polys = p->assets.polys.data();
    sp.a.num_objs = num_objs;
    sp.a.max_verts = p->assets.max_verts;
    sp.a.max_faces = p->assets.max_faces;
    sp.max_pairs = cfg->max_contacts;
    sp.smem_contacts = cfg->smem_contacts > 0 ? cfg->smem_contacts : 8;  // exercise the overflow path
    if (sp.smem_contacts > sp.max_pairs) sp.smem_contacts = sp.max_pairs;
    p->overflow.resize(sp.max_pairs - sp.smem_contacts + 1);
    sp.overflow = p->overflow.data();
    sp.world_pairs = p->alloc<int32_t>(W);
    sp.world_contacts = p->alloc<int32_t>(W);
    sp.error_flags = &p->err;
    sp.probe.world = -1;
    WorldView v;
    size_t bytes = carveWorld(v, nullptr, stride, sp.max_pairs, sp.smem_contacts);
    p->smem.resize(bytes + 16);
    *out = p;
    return MDR_OK;
}

void mdr_phys_destroy(mdr_phys *p)
{
    if (!p) return;
    for (void *a : p->allocs) free(a);
    delete p;
}

int32_t mdr_phys_set_row_counts(mdr_phys *p, const int32_t *rows)
{
    PhysTables &t = p->params.t;
    for (int w = 0; w < t.num_worlds; w++) {
        if (rows[w] < 0 || rows[w] > p->cfg.max_bodies_per_world) return MDR_EINVAL;
        t.row_count[w] = rows[w];
        for (int i = 0; i < t.stride; i++) {
            size_t r = (size_t)w * t.stride + i;
            t.entity[2 * r] = 0;
            t.entity[2 * r + 1] = i < rows[w] ? i : -1;
            t.leaf_id[r] = i < rows[w] ? i : -1;
        }
        t.rebuild_flag[w] = 1;
    }
    return MDR_OK;
}

int32_t mdr_phys_set_joints(mdr_phys *p, int32_t max_joints, const int32_t *counts, const mdr_joint *joints)
{
    if (!p || max_joints < 0 || (max_joints > 0 && (!counts || !joints))) return MDR_EINVAL;
    const int W = p->cfg.num_worlds;
    for (int w = 0; w < W && max_joints > 0; w++) {
        if (counts[w] < 0 || counts[w] > max_joints) return MDR_EINVAL;
        for (int k = 0; k < counts[w]; k++) {
            const mdr_joint &j = joints[(size_t)w * max_joints + k];
            if (j.type < 0 || j.type > 1 || j.row1 < 0 || j.row2 < 0 ||
                j.row1 >= p->cfg.max_bodies_per_world || j.row2 >= p->cfg.max_bodies_per_world)
                return MDR_EINVAL;
        }
    }
    JointDev *dj = p->alloc<JointDev>((size_t)W * max_joints);
    int32_t *dc = p->alloc<int32_t>(W);
    if (max_joints > 0) {
        memcpy(dj, joints, (size_t)W * max_joints * sizeof(JointDev));
        memcpy(dc, counts, (size_t)W * 4);
    }
    p->params.joints = dj;
    p->params.joint_count = dc;
    p->params.joint_stride = max_joints;
    return MDR_OK;
}

int32_t mdr_phys_write_column(mdr_phys *p, int32_t col, const void *src)
{
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    const int n = p->cfg.max_bodies_per_world, nb = kColBytes[col];
    for (int w = 0; w < p->cfg.num_worlds; w++)
        memcpy((char *)*s + (size_t)w * p->stride * nb, (const char *)src + (size_t)w * n * nb, (size_t)n * nb);
    return MDR_OK;
}

int32_t mdr_phys_read_column(mdr_phys *p, int32_t col, void *dst)
{
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    const int n = p->cfg.max_bodies_per_world, nb = kColBytes[col];
    for (int w = 0; w < p->cfg.num_worlds; w++)
        memcpy((char *)dst + (size_t)w * n * nb, (char *)*s + (size_t)w * p->stride * nb, (size_t)n * nb);
    return MDR_OK;
}

int32_t mdr_phys_flag_rebuild(mdr_phys *p)
{
    memset(p->params.t.rebuild_flag, 1, p->cfg.num_worlds);
    return MDR_OK;
}

static void rebuildFlagged(mdr_phys *p)
{
    StepParams &sp = p->params;
    PhysTables &t = sp.t;
    for (int w = 0; w < t.num_worlds; w++) {
        if (!t.rebuild_flag[w]) continue;
        t.rebuild_flag[w] = 0;
        const size_t r0 = (size_t)w * t.stride;
        const int n = t.row_count[w];
        for (int i = 0; i < n; i++) {
            const ObjDev *o = sp.a.objs + t.obj[r0 + i];
            Box wb = applyTRS(objAABB(o), ld3(t.pos + 3 * r0, i), ld4(t.rot + 4 * r0, i), ld3(t.scale + 3 * r0, i));
            stBox(t.leaf_aabb + 6 * r0, i, expandWithMotion(wb, ldVelLin(t.vel + 6 * r0, i), sp.vel_expand, sp.accel_expand));
        }
        bvhRebuildWorld(t.leaf_aabb + 6 * r0, t.slot_box + 6 * r0, t.order + r0, t.sorted + r0, n);
    }
}

static void stepAll(mdr_phys *p)
{
    rebuildFlagged(p);
    StepParams &sp = p->params;
    char *base = (char *)(((uintptr_t)p->smem.data() + 15) & ~(uintptr_t)15);
    BlockCtx ctx;
    for (int w = 0; w < sp.t.num_worlds; w++) {
        WorldView v;
        carveWorld(v, base, p->stride, sp.max_pairs, sp.smem_contacts);
        v.overflow = sp.overflow;
        v.active = true;
        int group_level = 0;
        stepWorld(ctx, v, sp, w, &group_level);
    }
}

int32_t mdr_phys_step(mdr_phys *p, int32_t num_steps)
{
    for (int s = 0; s < num_steps; s++) stepAll(p);
    return (p->err & MDR_ERR_PAIR_OVERFLOW) ? MDR_EOVERFLOW : MDR_OK;
}
int32_t mdr_phys_step_async(mdr_phys *p, int32_t n) { return mdr_phys_step(p, n); }
int32_t mdr_phys_sync(mdr_phys *) { return MDR_OK; }

int32_t mdr_phys_probe_step(mdr_phys *p, int32_t world, int32_t max_pairs, int32_t *num_pairs,
                            int32_t *pairs, int32_t *num_contacts, mdr_contact *contacts)
{
    static_assert(sizeof(mdr_contact) == sizeof(ContactDev), "contact layout");
    ProbeDev &pr = p->params.probe;
    pr.world = world; pr.max_pairs = max_pairs; pr.max_contacts = p->cfg.max_contacts;
    pr.num_pairs = num_pairs; pr.pairs = pairs; pr.num_contacts = num_contacts;
    pr.contacts = (ContactDev *)contacts;
    memset(contacts, 0, sizeof(ContactDev) * (size_t)p->cfg.num_substeps * p->cfg.max_contacts);
    int32_t rc = mdr_phys_step(p, 1);
    pr.world = -1;
    for (int s = 0; s < p->cfg.num_substeps; s++) {
        ContactDev *row = (ContactDev *)contacts + (size_t)s * p->cfg.max_contacts;
        int out = 0;
        for (int i = 0; i < p->cfg.max_contacts; i++) {
            if (row[i].num_points <= 0) continue;
            ContactDev c = row[i];
            row[out++] = c;
        }
        for (int i = out; i < p->cfg.max_contacts; i++) memset(&row[i], 0, sizeof(ContactDev));
        num_contacts[s] = out;
    }
    return rc;
}

int32_t mdr_phys_stats(mdr_phys *p, int32_t *pairs, int32_t *contacts)
{
    if (pairs) memcpy(pairs, p->params.world_pairs, 4 * p->cfg.num_worlds);
    if (contacts) memcpy(contacts, p->params.world_contacts, 4 * p->cfg.num_worlds);
    return MDR_OK;
}

int32_t mdr_phys_read_bvh(mdr_phys *p, float *slot, float *leaf, uint16_t *order)
{
    const int n = p->cfg.max_bodies_per_world;
    PhysTables &t = p->params.t;
    for (int w = 0; w < p->cfg.num_worlds; w++) {
        size_t r0 = (size_t)w * t.stride;
        if (slot) memcpy(slot + (size_t)w * n * 6, t.slot_box + 6 * r0, 24 * (size_t)n);
        if (leaf) memcpy(leaf + (size_t)w * n * 6, t.leaf_aabb + 6 * r0, 24 * (size_t)n);
        if (order) memcpy(order + (size_t)w * n, t.order + r0, 2 * (size_t)n);
    }
    return MDR_OK;
}

int32_t mdr_phys_get_hull(mdr_phys *p, int32_t obj, uint32_t *half_edges, uint32_t *edges,
                          uint32_t *polys, float *planes, int32_t *counts)
{
    if (obj < 0 || obj >= (int)p->assets.objs.size()) return MDR_EINVAL;
    const ObjDev &o = p->assets.objs[obj];
    counts[0] = o.num_hedges; counts[1] = o.num_edges; counts[2] = o.num_faces; counts[3] = o.num_verts;
    if (o.type != MDR_PRIM_HULL) return MDR_OK;
    memcpy(half_edges, p->assets.hedges.data() + o.hedge_off, 16 * (size_t)o.num_hedges);
    memcpy(edges, p->assets.edges.data() + o.edge_off, 4 * (size_t)o.num_edges);
    memcpy(polys, p->assets.polys.data() + o.face_off, 4 * (size_t)o.num_faces);
    memcpy(planes, p->assets.planes.data() + 4 * (size_t)o.face_off, 16 * (size_t)o.num_faces);
    return MDR_OK;
}

// physTraceRaysKernel's body, one world after the other.
int32_t mdr_phys_trace_rays(mdr_phys *p, int32_t R, const float *origins, const float *dirs, const float *t_max,
                            float *out_t, float *out_normal, int32_t *out_entity, int32_t *out_row)
{
    PhysTables &t = p->params.t;
    std::vector<RayLeaf> leaves(t.stride);
    for (int w = 0; w < p->cfg.num_worlds; w++) {
        const size_t r0 = (size_t)w * t.stride;
        const int n = t.row_count[w];
        for (int r = 0; r < n; r++) {
            const int leaf = t.order[r0 + r];
            const size_t i = r0 + leaf;
            RayLeaf lf;
            for (int k = 0; k < 3; k++) {
                lf.lo[k] = t.slot_box[6 * i + k]; lf.hi[k] = t.slot_box[6 * i + 3 + k];
                lf.pos[k] = t.pos[3 * i + k]; lf.scale[k] = t.scale[3 * i + k];
            }
            for (int k = 0; k < 4; k++) lf.rot[k] = t.rot[4 * i + k];
            lf.obj = (t.obj[i] & 0xffff) | (leaf << 16);
            leaves[r] = lf;
        }
        for (int k = 0; k < R; k++) {
            const size_t ri = (size_t)w * R + k;
            float hit_t;
            V3 hn;
            int row;
            traceRay(p->params.a, leaves.data(), n, mk3(origins[3 * ri], origins[3 * ri + 1], origins[3 * ri + 2]),
                     mk3(dirs[3 * ri], dirs[3 * ri + 1], dirs[3 * ri + 2]), t_max ? t_max[ri] : INFINITY,
                     &hit_t, &hn, &row);
            out_t[ri] = hit_t;
            out_normal[3 * ri] = hn.x; out_normal[3 * ri + 1] = hn.y; out_normal[3 * ri + 2] = hn.z;
            if (out_row) out_row[ri] = row;
            if (out_entity) {
                out_entity[2 * ri] = row >= 0 ? t.entity[2 * (r0 + row)] : -1;
                out_entity[2 * ri + 1] = row >= 0 ? t.entity[2 * (r0 + row) + 1] : -1;
            }
        }
    }
    return MDR_OK;
}
int32_t mdr_phys_trace_rays_device(mdr_phys *p, int32_t R, const float *o, const float *d, const float *tm,
                                   float *ot, float *on, int32_t *oe, int32_t *orow)
{
    return mdr_phys_trace_rays(p, R, o, d, tm, ot, on, oe, orow);
}

// the feeder kernels' bodies over host buffers
int32_t mdr_phys_pack_instances(mdr_phys *p, const float *world_offset, float *out_transforms, int32_t *out_object_ids)
{
    FeederJob job;
    memset(&job, 0, sizeof(job));
    for (int k = 0; k < 3; k++) job.world_offset[k] = world_offset ? world_offset[k] : 0.f;
    job.transforms = out_transforms;
    job.object_ids = out_object_ids;
    for (int w = 0; w < p->cfg.num_worlds; w++)
        for (int i = 0; i < p->params.t.stride; i++) feedInstance(p->params.t, job, w, i);
    return MDR_OK;
}
int32_t mdr_phys_pack_views(mdr_phys *p, int32_t V, const int32_t *view_rows, const float *camera_offset,
                            const float *y_scale, const float *world_offset, float *out_views)
{
    FeederJob job;
    memset(&job, 0, sizeof(job));
    for (int k = 0; k < 3; k++) job.world_offset[k] = world_offset ? world_offset[k] : 0.f;
    job.views_per_world = V; job.view_rows = view_rows; job.camera_offset = camera_offset;
    job.y_scale = y_scale; job.views = out_views;
    for (int w = 0; w < p->cfg.num_worlds; w++)
        for (int k = 0; k < V; k++) feedView(p->params.t, job, w, k);
    return MDR_OK;
}
int32_t mdr_phys_row_stride(mdr_phys *p) { return p->params.t.stride; }

}
