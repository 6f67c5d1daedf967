/* madrona_b200 -- C ABI of the B200-native batched world step.
 *
 * This is the drop-in boundary for the hot path of shacklettbp/gpu-ecs-madrona
 * (SURVEY.md section 8b). The reference has no C ABI: its surface is C++20
 * templates (include/madrona/mw_gpu.hpp:20-76 MWCudaExecutor,
 * include/madrona/mw_cpu.hpp:9-81 TaskGraphExecutor, include/madrona/state.hpp:
 * 109-134 ECSRegistry, include/madrona/taskgraph.hpp:15-113 TaskGraph::Builder,
 * include/madrona/physics.hpp:419-447 RigidBodyPhysicsSystem). Each entry point
 * below names the reference interface it stands in for. Plain pointers and
 * sizes only; all functions return MDR_OK (0) or a negative MDR_E* code for
 * argument errors, and abort on CUDA errors like the reference's REQ_CUDA
 * (include/madrona/cuda_utils.hpp:46-54). The library is libmadrona_b200.so;
 * it has no CPU fallback: creation fails with MDR_ENODEVICE without a GPU.
 */
#ifndef MADRONA_B200_H
#define MADRONA_B200_H
#include <stdint.h>

#include "madrona_b200_apps.h"

#ifdef __cplusplus
extern "C" {
#endif

enum {
    MDR_OK = 0,
    MDR_EINVAL = -1,      /* bad argument                                   */
    MDR_ENODEVICE = -2,   /* no CUDA device / wrong architecture            */
    MDR_ECAPACITY = -3,   /* a configured capacity does not fit the device  */
    MDR_EOVERFLOW = -4,   /* a per-world capacity overflowed during a step  */
    MDR_ESTATE = -5       /* call made in the wrong phase (e.g. after build) */
};

/* CollisionPrimitive::Type, include/madrona/physics.hpp:251-256 */
enum { MDR_PRIM_SPHERE = 1, MDR_PRIM_HULL = 2, MDR_PRIM_PLANE = 4 };
/* ResponseType, include/madrona/physics.hpp:162-166 */
enum { MDR_RESPONSE_DYNAMIC = 0, MDR_RESPONSE_KINEMATIC = 1, MDR_RESPONSE_STATIC = 2 };

/* One shared collision object: a row of the reference's ObjectManager
 * (physics.hpp:275-290: metadata, aabbs, primitives) plus the source mesh the
 * reference's PhysicsLoader::loadObjects consumes (src/physics/
 * physics_assets.cpp:205-391). Hull polygons use the FastPolygonList encoding
 * [n, i0 .. i(n-1)]* (physics.hpp:12-35). */
typedef struct mdr_object {
    int32_t type;
    float inv_inertia[3];
    float inv_mass;
    float mu_s;
    float mu_d;
    float aabb_min[3];
    float aabb_max[3];
    int32_t vert_offset;      /* first vertex in the shared vertex array       */
    int32_t num_verts;
    int32_t poly_offset;      /* first word in the shared polygon-word array   */
    int32_t num_poly_words;
} mdr_object;

/* RigidBodyPhysicsSystem::init arguments (src/physics/physics.cpp:1012-1036)
 * plus the executor's StateConfig (include/madrona/mw_gpu.hpp:20-34). */
typedef struct mdr_phys_config {
    int32_t num_worlds;
    int32_t max_bodies_per_world;   /* row capacity of the rigid-body archetype  */
    int32_t num_substeps;
    int32_t max_contacts;           /* candidate pairs / contacts per world      */
    float delta_t;
    float gravity[3];
    int32_t gpu_id;
    int32_t block_threads;          /* threads stepping one world; 0: default     */
    int32_t smem_contacts;          /* contact slots kept in shared memory; 0: default */
    int32_t worlds_per_block;       /* worlds stepped side by side by one block; 0: default */
} mdr_phys_config;

/* Columns of the rigid-body archetype: 0 = Entity, 1..12 = phys::Cols
 * (include/madrona/physics.hpp:449-464), in the reference's packed layout. */
enum {
    MDR_COL_ENTITY = 0, MDR_COL_POSITION = 1, MDR_COL_ROTATION = 2, MDR_COL_SCALE = 3,
    MDR_COL_VELOCITY = 4, MDR_COL_OBJECT_ID = 5, MDR_COL_RESPONSE_TYPE = 6,
    MDR_COL_SUBSTEP_PREV = 7, MDR_COL_PRESOLVE_POS = 8, MDR_COL_PRESOLVE_VEL = 9,
    MDR_COL_EXT_FORCE = 10, MDR_COL_EXT_TORQUE = 11, MDR_COL_LEAF_ID = 12,
    MDR_NUM_PHYS_COLS = 13
};

/* Contact, include/madrona/physics.hpp:185-192 (Loc reduced to a row). 112 B. */
typedef struct mdr_contact {
    int32_t ref_row, ref_pad, alt_row, alt_pad;
    float points[4][4];
    int32_t num_points;
    float normal[3];
    float lambda_n[4];
} mdr_contact;

typedef struct mdr_phys mdr_phys;

/* ---- rigid-body world step ------------------------------------------------
 * Stands in for MWCudaExecutor(StateConfig, CompileConfig) of a world whose
 * registerTypes/setupTasks are RigidBodyPhysicsSystem::{registerTypes,
 * setupBroadphaseTasks, setupSubstepTasks, setupCleanupTasks}
 * (include/madrona/physics.hpp:419-447; src/mw/cuda_exec.cpp:1692-1763). */
int32_t mdr_phys_create(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                        const float *verts, const uint32_t *poly_words, mdr_phys **out);
void mdr_phys_destroy(mdr_phys *p);

/* World constructor work (user WorldT ctor + RigidBodyPhysicsSystem::
 * registerEntity, physics.cpp:1045-1053): sets how many rows each world has
 * (host array [num_worlds]); rows get entity ids and leaf ids 0..n-1 in order
 * and a BVH rebuild is flagged, as bvh.rebuildOnUpdate() does. */
int32_t mdr_phys_set_row_counts(mdr_phys *p, const int32_t *rows_per_world);
/* Host <-> device copy of one whole column, [num_worlds][max_bodies][bytes]
 * world-major in the reference's packed layout (the StateManager export
 * layout of a fixed-size archetype, src/core/state.cpp:481-483). */
int32_t mdr_phys_write_column(mdr_phys *p, int32_t col, const void *host_src);
int32_t mdr_phys_read_column(mdr_phys *p, int32_t col, void *host_dst);
/* The same copy of a whole table tile-for-tile, [num_worlds][row_stride][bytes]
 * (row_stride >= max_bodies, mdr_phys_row_stride): one contiguous asynchronous
 * transfer on the executor's stream, meant for pinned host buffers
 * (mdr_host_alloc) -- the action-in / observation-out path of a trainer, the
 * reference's exportCopyIn/exportCopyOut nodes (src/mw/cuda_exec.cpp:1562-1579).
 * Ordered with the steps on the stream; mdr_phys_sync waits for it. */
int32_t mdr_phys_row_stride(mdr_phys *p);
int32_t mdr_phys_write_column_raw_async(mdr_phys *p, int32_t col, const void *host_src);
int32_t mdr_phys_read_column_raw_async(mdr_phys *p, int32_t col, void *host_dst);
/* Page-locked host memory (cudaHostAlloc / cudaFreeHost). */
void *mdr_host_alloc(uint64_t bytes);
void mdr_host_free(void *ptr);
/* Device pointer of a column: MWCudaExecutor::getExported (mw_gpu.hpp:70). */
void *mdr_phys_column_ptr(mdr_phys *p, int32_t col);
int32_t mdr_phys_column_bytes(int32_t col);

/* RigidBodyPhysicsSystem::reset semantics for the BVH (physics.cpp:1038-1043,
 * include/madrona/physics.inl:109-112): rebuild on the next step. */
int32_t mdr_phys_flag_rebuild(mdr_phys *p);
/* The same for the listed worlds only: RigidBodyPhysicsSystem::reset is per world (it flags the
 * context's own BVH, physics.cpp:1038-1043, include/madrona/physics.inl:109-112). Worlds that are
 * not named keep their grown slot boxes and traversal order, hence their pair order. */
int32_t mdr_phys_reset_worlds(mdr_phys *p, const int32_t *worlds, int32_t n);
/* Per-world halves of an episode reset: one world's rows of a column ([rows][bytes] on the host;
 * the reference writes components through ctx.getUnsafe in the world's own reset code), and one
 * world's row count (ids 0..rows-1; flags that world's BVH). */
int32_t mdr_phys_write_world_column(mdr_phys *p, int32_t col, int32_t world, int32_t rows, const void *src);
int32_t mdr_phys_read_world_column(mdr_phys *p, int32_t col, int32_t world, int32_t rows, void *dst);
int32_t mdr_phys_set_world_rows(mdr_phys *p, int32_t world, int32_t rows);
/* The rigid-body step over columns that live elsewhere (an ECS archetype's tables, so that physics
 * is a node of the same task graph as user systems: RigidBodyPhysicsSystem::setup*Tasks add their
 * nodes to the user's builder, src/physics/physics.cpp:1142-1205). columns[MDR_NUM_PHYS_COLS] are
 * device pointers [num_worlds][row_stride][column bytes] in phys::Cols order (include/madrona/
 * physics.hpp:449-464), 16-byte aligned; row_counts [num_worlds] on the device. The owner creates
 * entities and sets row counts; mdr_phys_set_row_counts answers MDR_ESTATE here. */
int32_t mdr_phys_create_on_columns(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                                   const float *verts, const uint32_t *poly_words, int32_t row_stride,
                                   void *const *columns, int32_t *row_counts, mdr_phys **out);

/* Joint constraints: the ConstraintData archetype (one JointConstraint per
 * entity, include/madrona/physics.hpp:196-240), collected every substep by
 * collectConstraintsSystem (src/physics/physics.cpp:34-40,1160-1161) and solved
 * after the substep's contacts by solvePositions -> handleJointConstraint
 * (:537-648,663-670), in table row order. The two entities are given as rows of
 * the joint's world. type 0 = Fixed (JointConstraint::setupFixed,
 * include/madrona/physics.inl:151-169): data = attachRot1 w,x,y,z, attachRot2
 * w,x,y,z, separation. type 1 = Hinge (setupHinge, :171-191): data = a1Local,
 * a2Local, b1Local, b2Local. 84 bytes. */
typedef struct mdr_joint {
    int32_t type;
    int32_t row1, row2;
    float r1[3], r2[3];
    float data[12];
} mdr_joint;
/* Replaces the joint table: joints_per_world[w] joints for world w taken from
 * joints[w][0 .. max_joints_per_world). max_joints_per_world 0 removes them all.
 * MDR_EINVAL for a row outside the world's row capacity or an unknown type. */
int32_t mdr_phys_set_joints(mdr_phys *p, int32_t max_joints_per_world, const int32_t *joints_per_world,
                            const mdr_joint *joints);

/* MWCudaExecutor::run() (src/mw/cuda_exec.cpp:1777-1794), `num_steps` times.
 * mdr_phys_step blocks like run(); the _async form only enqueues on the
 * executor's stream (mdr_phys_sync waits). Returns MDR_EOVERFLOW after the
 * sync if any world overflowed max_contacts pairs. */
int32_t mdr_phys_step(mdr_phys *p, int32_t num_steps);
int32_t mdr_phys_step_async(mdr_phys *p, int32_t num_steps);
int32_t mdr_phys_sync(mdr_phys *p);
/* One run() with host-side actions and observations, the trainer's loop around
 * MWCudaExecutor::run() (copy actions into the exported columns, run, copy the
 * observation columns out; src/mw/cuda_exec.cpp:1562-1579,1777-1794) as ONE
 * call: the worlds are cut into `chunks` ranges and the copy-in of range k+1 and
 * the copy-out of range k-1 overlap the step of range k. Host buffers are whole
 * columns in the raw layout (num_worlds x row_stride elements), page-locked
 * (mdr_host_alloc) for the copies to overlap. Enqueues only; mdr_phys_sync waits.
 * A range is a whole number of resident waves of worlds (chunks is clamped to
 * the number of waves; one range per wave exposes the least copy-out). The
 * ranges' launches rotate over MDR_PHYS_IO_LANES streams (default 3, at most 4),
 * so that the next range's blocks fill the SMs a finishing launch frees.
 * MDR_PHYS_IO_TRACE=1 prints the device timeline of every call at the next sync. */
int32_t mdr_phys_step_io_async(mdr_phys *p, int32_t num_in, const int32_t *in_cols, const void *const *in_host,
                               int32_t num_out, const int32_t *out_cols, void *const *out_host, int32_t chunks);
void *mdr_phys_stream(mdr_phys *p);           /* cudaStream_t */

/* broadphase::BVH::traceRay (src/physics/broadphase.cpp:653-856; declared
 * include/madrona/physics.hpp:316-320), batched: `rays_per_world` rays in every
 * world against the broadphase state the last step left (call after >= 1 step;
 * the query behind CameraMode::Lidar, include/madrona/mw_render.hpp:9-13).
 * All arrays are world-major: origins / dirs [W][R][3], t_max [W][R] or NULL
 * (infinity); out_t [W][R] (t_max when nothing is hit), out_normal [W][R][3]
 * (zeros when nothing is hit), out_entity [W][R][2] = (gen, id), Entity::none()
 * = all ones when nothing is hit, out_row [W][R] = row of the hit body or -1
 * (out_entity / out_row may be NULL). The _device form takes device pointers
 * and only enqueues on the executor's stream; the other takes host pointers
 * and blocks. */
/* The observation feeder: RenderingSystem's two nodes (src/mw/render/interop.cpp:
 * 10-96, setupTasks :112-139) writing to plain DEVICE buffers instead of a Vulkan /
 * Metal renderer's (graphics APIs: unavailable on this box). Enqueued on the
 * executor's stream. `world_offset` is a host float[3] or NULL (zero).
 *  pack_instances = instanceTransformSetup: out_transforms [W][row_stride][12] is the
 *   row-major 3x4 object-to-world matrix Mat3x4::fromTRS(pos + world_offset, rot,
 *   scale) of every body (translation in column 3, interop.cpp:27-41),
 *   out_object_ids [W][row_stride] its ObjectID (-1 and a zero matrix past a world's
 *   rows). World-major row order (the reference appends by an atomic counter).
 *  pack_views = updateViewData: out_views [W][V][8] = PackedViewData {rotation
 *   w,x,y,z; pos + camera_offset + world_offset; y_scale} of the body in row
 *   view_rows[w][v] (ViewSettings, include/madrona/mw_render.hpp:19-25). */
int32_t mdr_phys_pack_instances(mdr_phys *p, const float *world_offset, float *out_transforms,
                                int32_t *out_object_ids);
int32_t mdr_phys_pack_views(mdr_phys *p, int32_t views_per_world, const int32_t *view_rows,
                            const float *camera_offset, const float *y_scale, const float *world_offset,
                            float *out_views);

int32_t mdr_phys_trace_rays(mdr_phys *p, int32_t rays_per_world, const float *origins, const float *dirs,
                            const float *t_max, float *out_t, float *out_normal, int32_t *out_entity,
                            int32_t *out_row);
int32_t mdr_phys_trace_rays_device(mdr_phys *p, int32_t rays_per_world, const float *origins, const float *dirs,
                                   const float *t_max, float *out_t, float *out_normal, int32_t *out_entity,
                                   int32_t *out_row);

/* One blocking step that also records the intermediates of `world`: the
 * candidate pair list (reference order) and every substep's contact list.
 * pairs[2*max_pairs], num_contacts[substeps], contacts[substeps*max_contacts].*/
int32_t mdr_phys_probe_step(mdr_phys *p, int32_t world, int32_t max_pairs, int32_t *num_pairs,
                            int32_t *pairs, int32_t *num_contacts, mdr_contact *contacts);
/* Per-world statistics of the last step: candidate pairs and contacts of the
 * last substep, host arrays [num_worlds] (either may be NULL). */
int32_t mdr_phys_stats(mdr_phys *p, int32_t *pairs_per_world, int32_t *contacts_per_world);
/* BVH singleton state for parity checks: grown slot boxes and current leaf
 * boxes [W][max_bodies][6], traversal order [W][max_bodies] (any may be NULL). */
int32_t mdr_phys_read_bvh(mdr_phys *p, float *slot_boxes, float *leaf_aabbs, uint16_t *order);
/* Half-edge tables built for hull object `obj` (HalfEdgeMesh, physics.hpp:
 * 54-146): counts[4] = half-edges, edges, faces, vertices. */
int32_t mdr_phys_get_hull(mdr_phys *p, int32_t obj, uint32_t *half_edges, uint32_t *edges,
                          uint32_t *polys, float *planes, int32_t *counts);
/* Collision hull from a Wavefront OBJ: PhysicsLoader::loadHullFromDisk (src/physics/
 * physics_assets.cpp:205-252) over imp::ImportedObject::importObject (src/common/importer.cpp:
 * 35-437). Host only, no device needed. Produces what an mdr_object of type MDR_PRIM_HULL takes:
 * vertices (merged by byte equality and numbered in order of first mention by the faces, as the
 * reference's importer does), FastPolygonList words and the object-space AABB {min, max}.
 * Two-call pattern: verts / poly_words may be NULL to query the counts; with arrays, *num_verts
 * and *num_poly_words hold their capacities on entry (MDR_EOVERFLOW if too small) and the counts
 * on return. Parse errors (everything the reference answers with FATAL / assert): MDR_EINVAL,
 * with a message in err (may be NULL). `mdr_load_hull_obj_file` reads the file at `path`. */
int32_t mdr_load_hull_obj(const char *obj_text, int64_t num_bytes, float *verts, int32_t *num_verts,
                          uint32_t *poly_words, int32_t *num_poly_words, float aabb[6],
                          char *err, int32_t err_cap);
int32_t mdr_load_hull_obj_file(const char *path, float *verts, int32_t *num_verts,
                               uint32_t *poly_words, int32_t *num_poly_words, float aabb[6],
                               char *err, int32_t err_cap);
/* Quaternion arithmetic of the step (math::Quat, include/madrona/math.hpp:478-730), host evaluation of the
 * library's own source, for the reference's known-answer tests (tests/math.cpp:26-51). op 0: angleAxis(a[0],
 * axis a[1..3]); 1: a * b; 2: a.rotateVec(b[0..2]) (in out[1..3]); 3: normalize; 4: inv. Quats are w,x,y,z. */
int32_t mdr_math_quat(int32_t op, const float *a, const float *b, float *out);
/* Stream-ordered hand-off to a consumer on another stream (py::CudaSync of the reference,
 * include/madrona/python.hpp:23-40, src/python/utils.cpp:15-44): the event recorded after the last step
 * enqueued on the executor's stream, and "make `stream` wait for `event`" (no host synchronisation). */
void *mdr_phys_done_event(mdr_phys *p);
int32_t mdr_stream_wait_event(void *stream, void *event);
/* Launch geometry chosen at creation: [0] blocks, [1] threads per block,
 * [2] dynamic shared memory bytes per block, [3] resident blocks per SM,
 * [4] threads per world, [5] worlds per block. */
int32_t mdr_phys_launch_info(mdr_phys *p, int32_t info[6]);
/* Average device time of a step, in ms, over the steps since the last call
 * (CUDA events on the executor's stream); <0 if none were timed. `launches`
 * receives that number of steps (world ranges for mdr_phys_step_io_async). */
float mdr_phys_kernel_ms(mdr_phys *p, int32_t *launches);
/* Step kernels the last mdr_phys_step_async call launched. A call of n > 1 steps is a BATCH: the host cannot see the
 * state between its steps, so one launch runs all n (a world's step s starts when its step s - 1 is done, wherever
 * that ran); results are those of n single steps, bit for bit. MDR_PHYS_NO_BATCH=1 (read at creation): n launches. */
int32_t mdr_phys_last_launches(mdr_phys *p);

/* ---- ECS state layer + task graph -------------------------------------------
 * Stands in for the registration / graph-building / executor surface a world
 * reaches through ECSRegistry (include/madrona/state.hpp:109-134), TaskGraph::
 * Builder (include/madrona/taskgraph.hpp:15-97) and MWCudaExecutor
 * (include/madrona/mw_gpu.hpp:55-76). Every world has its own entity id space
 * (one StateManager per world, src/mw/device/include/madrona/mw_gpu_entry.hpp:
 * 29-31). Archetype tables are [num_worlds][row_stride][bytes] per column in
 * HBM; ids, generations, row order and row counts follow the reference's
 * sequential semantics exactly. System functions are compiled ahead of time
 * and named by func id (gpu_ecs_madrona_b200/csrc/ecs_apps.cuh), where the
 * reference JIT-compiles them and assigns ids by scanning PTX
 * (src/mw/cuda_exec.cpp:596-675). Phases: register / add nodes -> build ->
 * make entities, write columns -> step. */
typedef struct mdr_ecs_config {
    int32_t num_worlds;
    int32_t first_world;          /* global index of world 0 (multi-GPU sharding; seeds)   */
    int32_t gpu_id;
    int32_t ids_per_world;        /* entity-id capacity per world; 0: sized from the tables */
    int32_t max_script_handles;   /* handle slots per world for mdr_ecs_script; 0: none      */
    int32_t app;                  /* set by mdr_app_setup                                     */
} mdr_ecs_config;
typedef struct mdr_ecs mdr_ecs;
enum { MDR_ARCH_FIXED = 1 };      /* registerFixedSizeArchetype: exports alias the table    */
enum { MDR_COMP_ENTITY = 0 };     /* component id of Entity in queries                       */

int32_t mdr_ecs_create(const mdr_ecs_config *cfg, mdr_ecs **out);
void mdr_ecs_destroy(mdr_ecs *e);
/* ECSRegistry::registerComponent<T>: returns the component id (registration
 * order, Entity = 0). Sizes are multiples of 4 bytes. */
int32_t mdr_ecs_register_component(mdr_ecs *e, int32_t bytes, int32_t align);
/* registerArchetype<Archetype<Cs...>> / registerFixedSizeArchetype: returns the
 * archetype id; max_rows is the per-world row capacity. */
int32_t mdr_ecs_register_archetype(mdr_ecs *e, const int32_t *comps, int32_t n, int32_t max_rows, int32_t flags);
/* registerSingleton<T> (include/madrona/state.inl:171-187): component + one-row
 * archetype + one entity per world from the init id cache. Returns the archetype id. */
int32_t mdr_ecs_register_singleton(mdr_ecs *e, int32_t bytes, int32_t align);
/* exportColumn<Archetype, Component>(slot) */
int32_t mdr_ecs_export_column(mdr_ecs *e, int32_t arch, int32_t comp, int32_t slot);
/* builder.addToGraph<ParallelForNode<Ctx, fn, Cs...>>(deps); returns the node id */
int32_t mdr_ecs_add_parallel_for(mdr_ecs *e, int32_t func, const int32_t *comps, int32_t n, const int32_t *deps,
                                 int32_t ndeps);
/* The same for a system compiled in the caller's translation unit: `device_fn` is the device address of
 * an mdr::ecs::NodeFn (include/madrona/b200/ecs_device.cuh) -- the whole row loop of the node for one
 * world, systems inlined. Needs mdr_ecs_set_step_launcher: a device function can only be called from a
 * kernel of its own module (the reference links user code and megakernel into one cubin at run time,
 * src/mw/cuda_exec.cpp:553-818; here they are compiled together ahead of time). */
int32_t mdr_ecs_add_parallel_for_fn(mdr_ecs *e, const void *device_fn, const int32_t *comps, int32_t n,
                                    const int32_t *deps, int32_t ndeps);
/* Launches the caller's instantiation of mdr::ecs::ecsGraphKernel over nodes [first_node, last_node)
 * on `stream`; called while the run graph is captured. Returns MDR_OK. */
typedef int32_t (*mdr_ecs_launch_fn)(const void *device_descriptor, int32_t first_node, int32_t last_node,
                                     int32_t num_worlds, void *stream, void *user);
int32_t mdr_ecs_set_step_launcher(mdr_ecs *e, mdr_ecs_launch_fn fn, void *user);
/* RigidBodyPhysicsSystem::setup*Tasks as ONE node: the rigid-body step kernel runs on the archetype's
 * own columns between the megakernel launches of the nodes before and after it. The physics object is
 * attached after mdr_ecs_build (it needs the column addresses): mdr_phys_create_on_ecs. */
int32_t mdr_ecs_add_physics(mdr_ecs *e, const int32_t *deps, int32_t ndeps);
int32_t mdr_ecs_attach_physics(mdr_ecs *e, int32_t node, mdr_phys *p);
int32_t mdr_phys_create_on_ecs(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                               const float *verts, const uint32_t *poly_words, mdr_ecs *e, int32_t arch,
                               mdr_phys **out);
/* Context::tmpAlloc arena (src/core/state.cpp:83-127) and per-world user data (WorldT, Context::data()):
 * bytes per world, before mdr_ecs_build. */
int32_t mdr_ecs_set_tmp_bytes(mdr_ecs *e, int32_t bytes_per_world);
int32_t mdr_ecs_set_world_data_bytes(mdr_ecs *e, int32_t bytes_per_world);
/* For kernels the caller launches itself on mdr_ecs_stream (world constructors): the device-side
 * descriptor (const mdr::ecs::EcsDev *), and the note that tables changed (exports are refreshed). */
const void *mdr_ecs_device_descriptor(mdr_ecs *e);
int32_t mdr_ecs_num_worlds(mdr_ecs *e);
int32_t mdr_ecs_tables_written(mdr_ecs *e);
/* ClearTmpNode<Archetype>, ResetTmpAllocNode (include/madrona/taskgraph.inl:83-101) */
int32_t mdr_ecs_add_clear_tmp(mdr_ecs *e, int32_t arch, const int32_t *deps, int32_t ndeps);
int32_t mdr_ecs_add_reset_tmp_alloc(mdr_ecs *e, const int32_t *deps, int32_t ndeps);
/* SortArchetypeNode (src/mw/device/sort_archetype.cpp:1281-1392): stable sort of
 * every world's rows by a 4-byte unsigned key component; remap_entities patches
 * the id map for archetypes whose rows carry entity ids. */
int32_t mdr_ecs_add_sort(mdr_ecs *e, int32_t arch, int32_t key_comp, int32_t remap_entities, const int32_t *deps,
                         int32_t ndeps);
/* Builder::build() + state allocation (src/mw/cuda_exec.cpp:1090-1355) */
int32_t mdr_ecs_build(mdr_ecs *e);
/* World-constructor loops: `count` x makeEntityNow<arch>() in every world. */
int32_t mdr_ecs_make_entities(mdr_ecs *e, int32_t arch, int32_t count);
/* Immediate-mode ops per world (script app): ops of world w are
 * ops[offsets[w] .. offsets[w+1]). */
int32_t mdr_ecs_script(mdr_ecs *e, const mdr_script_op *ops, const int32_t *offsets);
/* Host copy of a column, [num_worlds][rows][bytes] (rows <= row stride). */
int32_t mdr_ecs_write_column(mdr_ecs *e, int32_t arch, int32_t col, int32_t rows, const void *host);
int32_t mdr_ecs_read_column(mdr_ecs *e, int32_t arch, int32_t col, int32_t rows, void *host);
int32_t mdr_ecs_row_counts(mdr_ecs *e, int32_t arch, int32_t *host);
int32_t mdr_ecs_num_components(mdr_ecs *e);     /* registered so far, Entity included */
int32_t mdr_ecs_num_archetypes(mdr_ecs *e);
int32_t mdr_ecs_num_columns(mdr_ecs *e, int32_t arch);
int32_t mdr_ecs_column_bytes(mdr_ecs *e, int32_t arch, int32_t col);
int32_t mdr_ecs_row_stride(mdr_ecs *e, int32_t arch);
void *mdr_ecs_column_ptr(mdr_ecs *e, int32_t arch, int32_t col);
int32_t mdr_ecs_id_capacity(mdr_ecs *e);
/* Entity id map of one world (arrays of mdr_ecs_id_capacity entries); state[10] =
 * init cache {free head, free count, overflow head, overflow count}, world cache
 * {same}, global free head, number of ids. Returns the number of ids. */
int32_t mdr_ecs_read_idmap(mdr_ecs *e, int32_t world, int32_t *a, int32_t *b, uint32_t *gen, int32_t state[10]);
/* MWCudaExecutor::run(): export copy-in, the task-graph megakernel, export offsets,
 * export copy-out -- one CUDA graph launch per step. */
int32_t mdr_ecs_step(mdr_ecs *e, int32_t num_steps);
int32_t mdr_ecs_step_async(mdr_ecs *e, int32_t num_steps);
int32_t mdr_ecs_sync(mdr_ecs *e);
void *mdr_ecs_stream(mdr_ecs *e);
void *mdr_ecs_done_event(mdr_ecs *e);
float mdr_ecs_step_ms(mdr_ecs *e, int32_t *steps);
int32_t mdr_ecs_launches_per_step(mdr_ecs *e);
/* Kernels the last mdr_ecs_step_async call launched. A call with num_steps > 1 is a BATCH: nothing on the host can
 * see the state between its steps, so the num_steps walks of a world's graph run inside one launch (staged exports
 * copied in once before, offsets + copy-out once after); results are those of num_steps single steps, bit for bit.
 * Graphs with a physics node or a user-side megakernel keep one graph launch per step. MDR_ECS_NO_BATCH=1 (read at
 * mdr_ecs_build) turns batches off. */
int32_t mdr_ecs_last_launches(mdr_ecs *e);
/* Archetype column map, SURVEY 8 row a2: StaticIntegerMap<128> (include/madrona/hashmap.hpp:26-58, hashmap.inl:
 * 14-74, src/common/hashmap.cpp:17-116), kept per archetype as `ColumnMap columnLookup` (state.hpp:245-293).
 * `mdr_static_map_build` (host) fills both 128-word storage arrays exactly as the reference's constructor does
 * (perfect multiplicative hash, shift and multiplier in slots 14 / 15 of both); MDR_ECAPACITY where the reference
 * FATALs (no perfect hash within 128 slots, or more than 126 keys), MDR_EINVAL for key 0xffffffff.
 * `mdr_static_map_lookup` = lookup(): the value, or -1. `mdr_ecs_column_map` hands out the map built for archetype
 * `arch` at mdr_ecs_build (component id -> column; Entity = component 0 -> column 0); `mdr_ecs_lookup_columns`
 * resolves component ids ON THE DEVICE through the uploaded map, as Context::get<T>(Loc) does (state.inl:240-262). */
int32_t mdr_static_map_build(const uint32_t *keys, const uint32_t *values, int32_t n,
                             uint32_t *key_storage, uint32_t *value_storage);
int32_t mdr_static_map_lookup(const uint32_t *key_storage, const uint32_t *value_storage, uint32_t key);
int32_t mdr_ecs_column_map(mdr_ecs *e, int32_t arch, uint32_t *key_storage, uint32_t *value_storage);
int32_t mdr_ecs_lookup_columns(mdr_ecs *e, int32_t arch, const int32_t *comp_ids, int32_t n, int32_t *out_cols);
/* getExported(slot): device pointer; staged exports hold mdr_ecs_export_rows rows
 * (world-major concatenation), fixed ones alias the table. */
void *mdr_ecs_get_exported(mdr_ecs *e, int32_t slot);
int64_t mdr_ecs_export_rows(mdr_ecs *e, int32_t slot);
int32_t mdr_ecs_export_bytes_per_row(mdr_ecs *e, int32_t slot);
/* Host copy of an exported column; returns the bytes copied. */
int64_t mdr_ecs_read_exported(mdr_ecs *e, int32_t slot, void *host, int64_t max_bytes);
/* run() + the exported columns `slots` in host memory (page-locked for full speed) as one call: the trainer's loop
 * around MWCudaExecutor::run() + Tensor reads (src/mw/cuda_exec.cpp:1777-1794). out_bytes[k] = bytes written to
 * host[k] (rows x bytes per row); MDR_ECAPACITY if a buffer is too small (that column is not copied). Blocking. */
int32_t mdr_ecs_step_read(mdr_ecs *e, int32_t num_slots, const int32_t *slots, void *const *host,
                          const int64_t *max_bytes, int64_t *out_bytes);
/* registerTypes + setupTasks + world constructors of a built-in world
 * (MDR_APP_* of madrona_b200_apps.h). */
int32_t mdr_app_setup(mdr_ecs *e, int32_t app, const int32_t *params, int32_t nparams);

const char *mdr_version(void);
int32_t mdr_device_count(void);

/* ---- global sort / compaction of a shared archetype table ---------------------
 * SortArchetypeNodeBase::addToGraph (src/mw/device/sort_archetype.cpp:1267-1392) and the nodes
 * it chains (:960-1265) as one operator: rows of a table stored as one global SoA (not per-world
 * tiles) are stably sorted by the low 8 * passes bits of a u32 key column, passes =
 * ceil(bits(num_worlds + 1) / 8) when the key is the WorldID column (num_worlds > 0) and 4
 * otherwise (:1290-1299); every column is permuted alike; after a WorldID sort the table keeps
 * the rows whose top digit is below 255 (resizeTable :1216-1221: destroyed rows carry WorldID -1
 * and sort last) and entity_row[entity id] is rewritten for them (rearrangeEntities :1243-1263).
 * Column 0 is the Entity column {gen, id} (8 bytes); column sizes are multiples of 4 bytes.
 * columns_in / columns_out are DEVICE pointers (16-byte aligned), one per column; a column may
 * be sorted in place (same pointer in both: staged and copied back, as the reference does) or
 * into a second buffer (one read + one write per byte). Blocking, like MWCudaExecutor::run(). */
typedef struct mdr_sort mdr_sort;
int32_t mdr_sort_num_passes(int32_t num_worlds);
int32_t mdr_sort_create(int32_t gpu_id, int32_t max_rows, int32_t num_columns, const int32_t *column_bytes,
                        int32_t key_column, int32_t num_worlds, mdr_sort **out);
void mdr_sort_destroy(mdr_sort *s);
int32_t mdr_sort_run(mdr_sort *s, const void *const *columns_in, void *const *columns_out, int32_t num_rows,
                     int32_t *entity_row, int32_t max_entity_ids, int32_t *out_num_rows);
/* Allocation-free key sort. NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it). The key sort of mdr_sort_run (histogram, scan, one Onesweep pass per 8-bit digit) with all of its scratch in
 * a fixed device array inside the library instead of buffers from mdr_sort_create: keys_in [num_rows] u32 (device,
 * 16-byte aligned) -> keys_out [num_rows] (stably sorted by the low 8 * ceil(num_key_bits / 8) bits) and perm_out
 * [num_rows] i32 (source row of every output row); keys_out must not alias keys_in. Work is ENQUEUED on `stream`
 * (a cudaStream_t, NULL = the default stream) and not waited for. Calls no cudaMalloc / cudaFree or any allocating
 * routine. LIMITS: num_rows <= MDR_SORT_STATIC_MAX_ROWS (more: MDR_ECAPACITY and a message); the array
 * (mdr_sort_static_scratch_bytes(), ~0.6 MB) is held by every GPU that loads the library until exit; every call on
 * a device shares it, so calls that overlap on the GPU race -- order them on one stream or wait in between. The
 * array keeps the last call's contents; each call clears the words that must start at zero. */
#define MDR_SORT_STATIC_MAX_ROWS 65536
int32_t mdr_sort_pairs_static(int32_t gpu_id, const uint32_t *keys_in, uint32_t *keys_out, int32_t *perm_out,
                              int32_t num_rows, int32_t num_key_bits, void *stream);
int32_t mdr_sort_static_max_rows(void);
int64_t mdr_sort_static_scratch_bytes(void);
/* Device time of the last mdr_sort_run (CUDA events on its stream) and its kernel launches. */
float mdr_sort_last_ms(mdr_sort *s, int32_t *launches);

#ifdef __cplusplus
}
#endif
#endif
