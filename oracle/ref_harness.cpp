// TEST INFRASTRUCTURE ONLY — drives the UNMODIFIED reference physics
// (/root/reference/src/physics + src/core, single-world StateManager)
// behind the orc_* interface of oracle_api.h. No reference source is copied: the
// reference .cpp files are compiled where they lie by oracle/Makefile and linked
// with this file into oracle/_ref/libmadrona_ref.so.
//
// One StateManager / Context / TaskGraph per world, mirroring
// include/madrona/mw_cpu.inl:5-68 (TaskGraphExecutor) which cannot be used
// directly with src/physics in this snapshot.
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <functional>
#include <map>
#include <memory>
#include <thread>
#include <vector>
#include <array>
#include <tuple>
#include <utility>
#include <type_traits>
#include <new>
#include <string>
#include <mutex>
#include <fcntl.h>
#include <unistd.h>

// The harness reads intermediates (task-graph node list, BVH boxes) that the
// reference keeps private: oracle/Makefile compiles with -fno-access-control.
#include <madrona/taskgraph.hpp>
#include <madrona/physics.hpp>
#include <madrona/components.hpp>
#include <madrona/custom_context.hpp>

#include "physics_impl.hpp"   // -I/root/reference/src/physics
#include "worker_init.hpp"    // -I/root/reference/src/core

#include "oracle_api.h"

using namespace madrona;
using namespace madrona::math;
using namespace madrona::base;
using namespace madrona::phys;
using madrona::phys::broadphase::LeafID;
using madrona::phys::solver::SubstepPrevState;
using madrona::phys::solver::PreSolvePositional;
using madrona::phys::solver::PreSolveVelocity;

namespace {

// Column order is the ABI of phys::Cols (include/madrona/physics.hpp:449-464).
struct RigidBody : Archetype<Position, Rotation, Scale, Velocity, ObjectID,
    ResponseType, SubstepPrevState, PreSolvePositional, PreSolveVelocity,
    ExternalForce, ExternalTorque, LeafID> {};

struct Sim;
class Engine : public CustomContext<Engine, Sim> {
    using CustomContext::CustomContext;
};

struct Sim : WorldBase {
    Sim(Engine &ctx) : WorldBase(ctx) {}
};

struct SharedObjects {
    ObjectManager mgr;
    std::vector<RigidBodyMetadata> metadata;
    std::vector<AABB> aabbs;
    std::vector<CollisionPrimitive> prims;
};

struct World {
    StateManager *sm;
    StateCache *cache;
    Sim *sim;
    Engine *ctx;
    TaskGraph *graph;
    std::vector<Entity> bodies;
    bool has_hinge = false;
};

struct Handle {
    orc_config cfg;
    SharedObjects objs;
    std::vector<World> worlds;
};

void buildGraph(World &w, int substeps)
{
    TaskGraph::Builder builder(*w.ctx);
    auto bp = RigidBodyPhysicsSystem::setupBroadphaseTasks(builder, {});
    auto ss = RigidBodyPhysicsSystem::setupSubstepTasks(builder, {bp}, substeps);
    RigidBodyPhysicsSystem::setupCleanupTasks(builder, {ss});
    // TaskGraph is move-only through its private ctor; keep it on the heap.
    w.graph = new TaskGraph(builder.build());
}

}

extern "C" {

const char *orc_kind(void) { return "reference"; }

void *orc_create(const orc_config *cfg,
                 const orc_object *objs, int32_t num_objs,
                 const float *verts, const uint32_t *poly_words,
                 const float *pos, const float *rot, const float *scale,
                 const float *vel, const int32_t *obj_ids,
                 const uint32_t *response,
                 const float *ext_force, const float *ext_torque)
{
    Handle *h = new Handle;
    h->cfg = *cfg;
    SharedObjects &so = h->objs;
    so.metadata.resize(num_objs);
    so.aabbs.resize(num_objs);
    so.prims.resize(num_objs);
    for (int i = 0; i < num_objs; i++) {
        const orc_object &o = objs[i];
        so.metadata[i] = RigidBodyMetadata {
            { o.inv_inertia[0], o.inv_inertia[1], o.inv_inertia[2] },
            o.inv_mass, o.mu_s, o.mu_d,
        };
        so.aabbs[i] = AABB {
            { o.aabb_min[0], o.aabb_min[1], o.aabb_min[2] },
            { o.aabb_max[0], o.aabb_max[1], o.aabb_max[2] },
        };
        memset(&so.prims[i], 0, sizeof(CollisionPrimitive));
        so.prims[i].type = (CollisionPrimitive::Type)o.type;
        if (o.type == (int)CollisionPrimitive::Type::Hull) {
            geometry::FastPolygonList polys {};
            polys.allocate(o.num_poly_words);
            const uint32_t *p = poly_words + o.poly_offset;
            int32_t k = 0;
            while (k < o.num_poly_words) {
                uint32_t n = p[k];
                polys.addPolygon(Span<const uint32_t>(p + k + 1, n));
                k += n + 1;
            }
            so.prims[i].hull.halfEdgeMesh.construct(polys, o.num_verts,
                (const Vector3 *)(verts + 3 * o.vert_offset));
            polys.free();
        }
    }
    so.mgr = ObjectManager {
        so.metadata.data(), so.aabbs.data(), so.prims.data(),
        nullptr, nullptr, nullptr, nullptr, nullptr,
    };

    const int W = cfg->num_worlds, N = cfg->bodies_per_world;
    h->worlds.resize(W);
    for (int wi = 0; wi < W; wi++) {
        World &w = h->worlds[wi];
        w.sm = new StateManager();
        w.cache = new StateCache();
        ECSRegistry reg(w.sm, nullptr);
        base::registerTypes(reg);
        RigidBodyPhysicsSystem::registerTypes(reg);
        reg.registerArchetype<RigidBody>();

        WorkerInit wi_init { w.sm, w.cache };
        w.sim = (Sim *)malloc(sizeof(Sim));
        w.ctx = new Engine(w.sim, wi_init);
        new (w.sim) Sim(*w.ctx);

        RigidBodyPhysicsSystem::init(*w.ctx, &so.mgr, cfg->delta_t,
            cfg->num_substeps,
            Vector3 { cfg->gravity[0], cfg->gravity[1], cfg->gravity[2] },
            N, cfg->max_contacts, 64);

        w.bodies.resize(N);
        for (int i = 0; i < N; i++) {
            size_t b = (size_t)wi * N + i;
            Vector3 p { pos[3*b], pos[3*b+1], pos[3*b+2] };
            Quat q { rot[4*b], rot[4*b+1], rot[4*b+2], rot[4*b+3] };
            Diag3x3 s { scale[3*b], scale[3*b+1], scale[3*b+2] };
            Velocity v {
                { vel[6*b], vel[6*b+1], vel[6*b+2] },
                { vel[6*b+3], vel[6*b+4], vel[6*b+5] },
            };
            Vector3 f { ext_force[3*b], ext_force[3*b+1], ext_force[3*b+2] };
            Vector3 t { ext_torque[3*b], ext_torque[3*b+1], ext_torque[3*b+2] };
            Entity e = w.ctx->makeEntityNow<RigidBody>(
                Position(p), Rotation(q), Scale(s), v,
                ObjectID { obj_ids[b] }, (ResponseType)response[b],
                SubstepPrevState { p, q }, PreSolvePositional { p, q },
                PreSolveVelocity { Vector3::zero(), Vector3::zero() },
                ExternalForce(f), ExternalTorque(t), LeafID { -1 });
            w.ctx->getUnsafe<LeafID>(e) = RigidBodyPhysicsSystem::registerEntity(
                *w.ctx, e, ObjectID { obj_ids[b] });
            w.bodies[i] = e;
        }
        w.ctx->getSingleton<broadphase::BVH>().rebuildOnUpdate();
        buildGraph(w, cfg->num_substeps);
    }
    return h;
}

void orc_destroy(void *hp)
{
    // The reference's per-world allocations (rawAlloc'd BVH/solver buffers,
    // StateManager tables) have no matching teardown path; leak them as the
    // reference's own executors do at process exit.
    delete (Handle *)hp;
}

void orc_flag_rebuild(void *hp)
{
    Handle *h = (Handle *)hp;
    for (World &w : h->worlds) {
        w.ctx->getSingleton<broadphase::BVH>().rebuildOnUpdate();
    }
}

// RigidBodyPhysicsSystem::reset is per world: only this context's BVH is flagged (physics.cpp:1038-1043).
void orc_flag_rebuild_world(void *hp, int32_t world)
{
    ((Handle *)hp)->worlds[world].ctx->getSingleton<broadphase::BVH>().rebuildOnUpdate();
}

// applyJointAxisConstraint (src/physics/physics.cpp:529-531) prints both axes of
// every hinge in every substep: stdout goes to /dev/null while such worlds step.
struct QuietStdout {
    int saved = -1;
    explicit QuietStdout(bool on)
    {
        if (!on) return;
        fflush(stdout);
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        dup2(nul, 1);
        close(nul);
    }
    ~QuietStdout()
    {
        if (saved < 0) return;
        fflush(stdout);
        dup2(saved, 1);
        close(saved);
    }
};

static bool anyHinge(Handle *h)
{
    for (World &w : h->worlds) if (w.has_hinge) return true;
    return false;
}

void orc_set_joints(void *hp, int32_t max_joints, const int32_t *counts, const orc_joint *joints)
{
    Handle *h = (Handle *)hp;
    for (size_t wi = 0; wi < h->worlds.size(); wi++) {
        World &w = h->worlds[wi];
        for (int k = 0; k < counts[wi]; k++) {
            const orc_joint &j = joints[wi * (size_t)max_joints + k];
            const float *d = j.data;
            Vector3 r1 { j.r1[0], j.r1[1], j.r1[2] }, r2 { j.r2[0], j.r2[1], j.r2[2] };
            JointConstraint jc = j.type == 0 ?
                JointConstraint::setupFixed(w.bodies[j.row1], w.bodies[j.row2],
                    Quat { d[0], d[1], d[2], d[3] }, Quat { d[4], d[5], d[6], d[7] }, r1, r2, d[8]) :
                JointConstraint::setupHinge(w.bodies[j.row1], w.bodies[j.row2],
                    Vector3 { d[0], d[1], d[2] }, Vector3 { d[3], d[4], d[5] },
                    Vector3 { d[6], d[7], d[8] }, Vector3 { d[9], d[10], d[11] }, r1, r2);
            w.ctx->makeEntityNow<ConstraintData>(jc);
            if (j.type != 0) w.has_hinge = true;
        }
    }
}

void orc_step(void *hp, int32_t num_steps, int32_t num_threads)
{
    Handle *h = (Handle *)hp;
    const int W = (int)h->worlds.size();
    QuietStdout quiet(anyHinge(h));
    if (num_threads <= 1) {
        for (int s = 0; s < num_steps; s++)
            for (int i = 0; i < W; i++)
                h->worlds[i].graph->run(h->worlds[i].ctx);
        return;
    }
    // Mirrors src/mw/cpu_exec.cpp:244-284: workers pull world indices from an
    // atomic counter; one world is only ever touched by one thread. Worlds are
    // independent, so each worker runs all num_steps of the world it pulled
    // (no per-step barrier: the most favourable schedule for the CPU baseline).
    std::atomic<int> next { 0 };
    std::vector<std::thread> pool;
    for (int t = 0; t < num_threads; t++) {
        pool.emplace_back([&]() {
            while (true) {
                int i = next.fetch_add(1, std::memory_order_relaxed);
                if (i >= W) break;
                for (int s = 0; s < num_steps; s++)
                    h->worlds[i].graph->run(h->worlds[i].ctx);
            }
        });
    }
    for (auto &t : pool) t.join();
}

static void *colPtr(World &w, int col, int row)
{
    // Row i of the RigidBody table; single-world StateManager: column 0 is
    // Entity, user columns start at 1 (include/madrona/state.hpp:382-387).
    Loc l = w.ctx->getLoc(w.bodies[row]);
    auto &store = *w.sm->archetype_stores_[l.archetype];
    char *base = (char *)store.tblStorage.tbl.data(col);
    return base + (size_t)l.row * orc_col_bytes[col];
}

void orc_read(void *hp, int32_t col, void *out)
{
    Handle *h = (Handle *)hp;
    const int N = h->cfg.bodies_per_world, nb = orc_col_bytes[col];
    char *o = (char *)out;
    for (size_t wi = 0; wi < h->worlds.size(); wi++)
        for (int i = 0; i < N; i++)
            memcpy(o + (wi * N + i) * nb, colPtr(h->worlds[wi], col, i), nb);
}

void orc_write(void *hp, int32_t col, const void *in)
{
    Handle *h = (Handle *)hp;
    const int N = h->cfg.bodies_per_world, nb = orc_col_bytes[col];
    const char *o = (const char *)in;
    for (size_t wi = 0; wi < h->worlds.size(); wi++)
        for (int i = 0; i < N; i++)
            memcpy(colPtr(h->worlds[wi], col, i), o + (wi * N + i) * nb, nb);
}


// Leaf `leaf`'s slot box inside its parent node (the grown box refitLeaf
// maintains, src/physics/broadphase.cpp:545-594).
static void slotBox(broadphase::BVH &bvh, int leaf, float *out)
{
    uint32_t lp = bvh.leaf_parents_[leaf];
    auto &n = bvh.nodes_[lp >> 2];
    int s = lp & 3;
    out[0] = n.minX[s]; out[1] = n.minY[s]; out[2] = n.minZ[s];
    out[3] = n.maxX[s]; out[4] = n.maxY[s]; out[5] = n.maxZ[s];
}

int32_t orc_probe_step(void *hp, int32_t world, orc_probe *out)
{
    Handle *h = (Handle *)hp;
    World &w = h->worlds[world];
    const int N = h->cfg.bodies_per_world;
    const int S = h->cfg.num_substeps;
    TaskGraph &tg = *w.graph;
    int32_t rc = 0;
    QuietStdout quiet(w.has_hinge);

    // Node layout (src/physics/physics.cpp:1142-1205, broadphase.cpp:934-991,
    // narrowphase.cpp:1768-1782): 0 updateLeaf, 1 updateBVH, 2 refit,
    // 3 findOverlapping, then per substep 8 nodes
    // [collect, substep, narrowphase, resetTmp, solvePos, setVel, solveVel, resetTmp].
    const uint32_t cand_arch = TypeTracker::typeID<CandidateTemporary>();
    auto &solver = w.ctx->getSingleton<SolverData>();

    for (CountT ni = 0; ni < tg.sorted_nodes_.size(); ni++) {
        const auto &node = tg.sorted_nodes_[ni];
        node.fn((NodeBase *)(&tg.node_datas_[node.dataIDX].userData[0]), w.ctx);

        if (ni == 3) {
            auto &store = *w.sm->archetype_stores_[cand_arch];
            int np = (int)store.tblStorage.tbl.numRows();
            const CandidateCollision *cc =
                (const CandidateCollision *)store.tblStorage.tbl.data(1);
            out->num_pairs = np;
            if (np > out->max_pairs) { rc = -1; np = out->max_pairs; }
            for (int i = 0; i < np; i++) {
                out->pairs[2*i] = cc[i].a.row;
                out->pairs[2*i+1] = cc[i].b.row;
            }
        }
        if (ni >= 4 && ni < CountT(4 + 8 * S) && ((ni - 4) % 8) == 2) {
            int sub = (int)(ni - 4) / 8;
            int nc = (int)solver.numContacts.load_relaxed();
            out->num_contacts[sub] = nc;
            static_assert(sizeof(Contact) == sizeof(orc_contact));
            for (int i = 0; i < nc; i++) {
                Contact c = solver.contacts[i];
                orc_contact &o = out->contacts[(size_t)sub * h->cfg.max_contacts + i];
                o.ref_row = c.ref.row; o.ref_pad = 0;
                o.alt_row = c.alt.row; o.alt_pad = 0;
                memcpy(o.points, c.points, sizeof(o.points));
                o.num_points = c.numPoints;
                memcpy(o.normal, &c.normal, 12);
                memset(o.lambda_n, 0, 16);
                // points beyond num_points are indeterminate in the reference
                for (int k = c.numPoints; k < 4; k++)
                    memset(o.points[k], 0, 16);
            }
        }
    }

    auto &bvh = w.ctx->getSingleton<broadphase::BVH>();
    for (int i = 0; i < N; i++) {
        slotBox(bvh, i, out->slot_boxes + 6 * i);
        memcpy(out->leaf_aabbs + 6 * i, &bvh.leaf_aabbs_[i], 24);
    }
    // Full (unpruned) traversal in findOverlaps order (physics.inl:61-100).
    {
        int32_t stack[128]; int sz = 1; stack[0] = 0; int rank = 0;
        while (sz > 0) {
            const auto &node = bvh.nodes_[stack[--sz]];
            for (int i = 0; i < 4; i++) {
                if (!node.hasChild(i)) continue;
                if (node.isLeaf(i)) out->ref_rank[node.leafIDX(i)] = rank++;
                else stack[sz++] = node.children[i];
            }
        }
    }
    return rc;
}

void orc_trace_rays(void *hp, int32_t world, int32_t n, const float *origins, const float *dirs,
                    const float *t_max, float *out_t, float *out_normal, int32_t *out_row)
{
    Handle *h = (Handle *)hp;
    World &w = h->worlds[world];
    auto &bvh = w.ctx->getSingleton<broadphase::BVH>();
    for (int i = 0; i < n; i++) {
        float t = 0.f;
        Vector3 nrm { 0.f, 0.f, 0.f };
        Entity e = bvh.traceRay(Vector3 { origins[3*i], origins[3*i+1], origins[3*i+2] },
                                Vector3 { dirs[3*i], dirs[3*i+1], dirs[3*i+2] }, &t, &nrm, t_max[i]);
        out_t[i] = t;
        if (e == Entity::none()) {
            out_row[i] = -1;
            out_normal[3*i] = out_normal[3*i+1] = out_normal[3*i+2] = 0.f;
        } else {
            out_row[i] = w.ctx->getLoc(e).row;
            out_normal[3*i] = nrm.x; out_normal[3*i+1] = nrm.y; out_normal[3*i+2] = nrm.z;
        }
    }
}

void orc_feeder(void *hp, const float *world_offset, float *out_transforms, int32_t V,
                const int32_t *view_rows, const float *camera_offset, const float *y_scale, float *out_views)
{
    Handle *h = (Handle *)hp;
    const int N = h->cfg.bodies_per_world;
    const Vector3 off { world_offset[0], world_offset[1], world_offset[2] };
    for (size_t wi = 0; wi < h->worlds.size(); wi++) {
        World &w = h->worlds[wi];
        for (int i = 0; i < N; i++) {
            const Position &pos = *(const Position *)colPtr(w, ORC_COL_POSITION, i);
            const Rotation &rot = *(const Rotation *)colPtr(w, ORC_COL_ROTATION, i);
            const Scale &scale = *(const Scale *)colPtr(w, ORC_COL_SCALE, i);
            // the statements of instanceTransformSetup, interop.cpp:25-41
            Mat3x4 o2w = Mat3x4::fromTRS(pos + off, rot, scale);
            float (*m)[4] = (float (*)[4])(out_transforms + 12 * (wi * N + i));
            m[0][0] = o2w.cols[0].x; m[0][1] = o2w.cols[1].x; m[0][2] = o2w.cols[2].x; m[0][3] = o2w.cols[3].x;
            m[1][0] = o2w.cols[0].y; m[1][1] = o2w.cols[1].y; m[1][2] = o2w.cols[2].y; m[1][3] = o2w.cols[3].y;
            m[2][0] = o2w.cols[0].z; m[2][1] = o2w.cols[1].z; m[2][2] = o2w.cols[2].z; m[2][3] = o2w.cols[3].z;
        }
        for (int k = 0; k < V; k++) {
            const size_t vi = wi * V + k;
            const int row = view_rows[vi];
            const Position &pos = *(const Position *)colPtr(w, ORC_COL_POSITION, row);
            const Rotation &rot = *(const Rotation *)colPtr(w, ORC_COL_ROTATION, row);
            const Vector3 cam_off { camera_offset[3*vi], camera_offset[3*vi+1], camera_offset[3*vi+2] };
            // updateViewData, interop.cpp:69-79
            auto camera_pos = pos + cam_off + off;
            float *o = out_views + 8 * vi;
            Quat q = rot;
            o[0] = q.w; o[1] = q.x; o[2] = q.y; o[3] = q.z;
            o[4] = camera_pos.x; o[5] = camera_pos.y; o[6] = camera_pos.z; o[7] = y_scale[vi];
        }
    }
}

void orc_indeterminate(void *hp, int64_t *out)
{
    Handle *h = (Handle *)hp;
    for (size_t i = 0; i < h->worlds.size(); i++) out[i] = -1;
}

int32_t orc_get_hull(void *hp, int32_t obj,
                     uint32_t *half_edges, uint32_t *edges, uint32_t *polys,
                     float *planes, int32_t *counts)
{
    Handle *h = (Handle *)hp;
    const auto &m = h->objs.prims[obj].hull.halfEdgeMesh;
    counts[0] = m.mHalfEdgeCount; counts[1] = m.mEdgeCount;
    counts[2] = m.mPolygonCount; counts[3] = m.mVertexCount;
    for (uint32_t i = 0; i < m.mHalfEdgeCount; i++) {
        half_edges[4*i] = m.mHalfEdges[i].next;
        half_edges[4*i+1] = m.mHalfEdges[i].twin;
        half_edges[4*i+2] = m.mHalfEdges[i].rootVertex;
        half_edges[4*i+3] = m.mHalfEdges[i].polygon;
    }
    memcpy(edges, m.mEdges, 4 * m.mEdgeCount);
    memcpy(polys, m.mPolygons, 4 * m.mPolygonCount);
    memcpy(planes, m.mFacePlanes, 16 * m.mPolygonCount);
    return 0;
}

}
