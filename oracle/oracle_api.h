/* TEST INFRASTRUCTURE ONLY — never linked into or imported by the product path.
 *
 * Common C interface of the two CPU checkers for the collisions hot path:
 *   oracle/_ref/libmadrona_ref.so   the UNMODIFIED reference sources under
 *                                   /root/reference compiled by oracle/Makefile
 *                                   (kind "reference")
 *   oracle/libmadrona_port.so       our scalar C++ restatement (oracle/port.cpp,
 *                                   kind "port"), pinned bit-for-bit against the
 *                                   former (tests/test_oracle_pin.py, tests/golden/)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load these libraries.
 *
 * Both libraries export the same symbols with a different prefix-free name set
 * (orc_*), so they are loaded from separate .so files, never together by name.
 */
#pragma once
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One collision object shared by all worlds (reference: ObjectManager,
 * include/madrona/physics.hpp:275-290). type follows
 * CollisionPrimitive::Type (physics.hpp:252-256): 2 = Hull, 4 = Plane. */
typedef struct {
    int32_t type;
    float inv_inertia[3];
    float inv_mass;
    float mu_s;
    float mu_d;
    float aabb_min[3];
    float aabb_max[3];
    int32_t vert_offset;   /* first vertex in the shared vertex array   */
    int32_t num_verts;
    int32_t poly_offset;   /* first word in the shared polygon-list array */
    int32_t num_poly_words;/* FastPolygonList words: [n, i0..in-1]*      */
} orc_object;

typedef struct {
    int32_t num_worlds;
    int32_t bodies_per_world;
    int32_t num_substeps;
    int32_t max_contacts;
    float delta_t;
    float gravity[3];
} orc_config;

/* Column ids for orc_read / orc_write: 0 = Entity, 1..12 = phys::Cols
 * (include/madrona/physics.hpp:449-464). */
enum {
    ORC_COL_ENTITY = 0, ORC_COL_POSITION = 1, ORC_COL_ROTATION = 2, ORC_COL_SCALE = 3,
    ORC_COL_VELOCITY = 4, ORC_COL_OBJECT_ID = 5, ORC_COL_RESPONSE_TYPE = 6,
    ORC_COL_SUBSTEP_PREV = 7, ORC_COL_PRESOLVE_POS = 8, ORC_COL_PRESOLVE_VEL = 9,
    ORC_COL_EXT_FORCE = 10, ORC_COL_EXT_TORQUE = 11, ORC_COL_LEAF_ID = 12,
    ORC_NUM_COLS = 13
};
static const int32_t orc_col_bytes[ORC_NUM_COLS] =
    { 8, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4 };

/* One contact as the reference stores it (physics.hpp:185-192), with Loc
 * reduced to the row inside the rigid-body archetype. 112 bytes. */
typedef struct {
    int32_t ref_row;
    int32_t ref_pad;
    int32_t alt_row;
    int32_t alt_pad;
    float points[4][4];
    int32_t num_points;
    float normal[3];
    float lambda_n[4];
} orc_contact;

/* Intermediates of ONE step of ONE world. Caller sizes the arrays:
 * pairs[2*max_pairs], contacts[substeps*max_contacts], num_contacts[substeps],
 * slot_boxes[6*N] (grown leaf slot boxes, min xyz / max xyz), leaf_aabbs[6*N],
 * ref_rank[N] (position of each leaf in the reference BVH's full traversal). */
typedef struct {
    int32_t max_pairs;
    int32_t num_pairs;
    int32_t *pairs;
    int32_t *num_contacts;
    orc_contact *contacts;
    float *slot_boxes;
    float *leaf_aabbs;
    int32_t *ref_rank;
} orc_probe;

/* Body arrays are world-major [W][N][...]; rot is (w,x,y,z); vel is
 * (linear xyz, angular xyz); response: 0 Dynamic, 1 Kinematic, 2 Static. */
void *orc_create(const orc_config *cfg,
                 const orc_object *objs, int32_t num_objs,
                 const float *verts, const uint32_t *poly_words,
                 const float *pos, const float *rot, const float *scale,
                 const float *vel, const int32_t *obj_ids,
                 const uint32_t *response,
                 const float *ext_force, const float *ext_torque);
void orc_destroy(void *h);
/* Steps every world num_steps times, worlds striped over num_threads threads. */
void orc_step(void *h, int32_t num_steps, int32_t num_threads);
/* Gathers column col of all worlds into out ([W][N][orc_col_bytes[col]]). */
void orc_read(void *h, int32_t col, void *out);
void orc_write(void *h, int32_t col, const void *in);
/* Steps world w once, recording intermediates. Returns 0, or -1 on overflow. */
int32_t orc_probe_step(void *h, int32_t world, orc_probe *out);
/* Flags a BVH rebuild for every world (RigidBodyPhysicsSystem::reset semantics
 * without clearing leaves: bvh.rebuildOnUpdate()). */
void orc_flag_rebuild(void *h);
/* The same for one world (an episode reset of a subset of worlds). */
void orc_flag_rebuild_world(void *h, int32_t world);
/* Half-edge mesh tables the checker built for object obj (for pinning our own
 * HalfEdgeMesh construction): returns counts, copies up to cap entries. */
int32_t orc_get_hull(void *h, int32_t obj,
                     uint32_t *half_edges /*[4*n] next,twin,root,poly*/,
                     uint32_t *edges, uint32_t *polys, float *planes /*[4*n]*/,
                     int32_t *counts /*[4] hedges,edges,polys,verts*/);
/* Per-world count of manifold reductions whose 3rd/4th point the reference
 * never writes (narrowphase.cpp:830-853, undefined behaviour): worlds with a
 * non-zero count are excluded from bit-exact comparison against the STOCK
 * reference build (_ref/libmadrona_ref.so); against the build whose locals are
 * defined (_ref/libmadrona_ref_defined.so, oracle/Makefile) every world is
 * compared -- the restatement uses that build's value, zero. The compiled
 * reference cannot know and reports -1. */
void orc_indeterminate(void *h, int64_t *out /*[W]*/);
/* BVH::traceRay (src/physics/broadphase.cpp:653-856) for n rays in world `world`,
 * against the broadphase state the last step left (leaf transforms and slot
 * boxes are refreshed at the end of every step; call after >= 1 step).
 * origins / dirs [n][3], t_max [n]; out_t [n] (t_max when nothing is hit),
 * out_normal [n][3] (zeros when nothing is hit: the reference leaves it
 * unset), out_row [n] = row of the hit entity or -1. */
void orc_trace_rays(void *h, int32_t world, int32_t n, const float *origins, const float *dirs,
                    const float *t_max, float *out_t, float *out_normal, int32_t *out_row);
/* The renderer feeder's arithmetic (src/mw/render/interop.cpp:10-96) on the
 * checker's own columns: out_transforms [W][N][12] = row-major 3x4
 * Mat3x4::fromTRS(pos + world_offset, rot, scale) as interop.cpp:27-41 lays it
 * out; out_views [W][V][8] = {rot w,x,y,z; pos + camera_offset + world_offset;
 * y_scale} of the body in view_rows[w][v] (interop.cpp:69-79). */
void orc_feeder(void *h, const float *world_offset, float *out_transforms, int32_t views_per_world,
                const int32_t *view_rows, const float *camera_offset, const float *y_scale, float *out_views);
/* One joint constraint of one world (reference: JointConstraint,
 * include/madrona/physics.hpp:196-238, with the two entities reduced to rows of
 * the rigid-body table). type 0 = Fixed: data = attachRot1 (w,x,y,z),
 * attachRot2 (w,x,y,z), separation; type 1 = Hinge: data = a1Local, a2Local,
 * b1Local, b2Local (3 floats each). */
typedef struct {
    int32_t type;
    int32_t row1, row2;
    float r1[3], r2[3];
    float data[12];
} orc_joint;
/* Creates joints_per_world[w] ConstraintData entities in world w from
 * joints[w][0..], in that order (= the order collectConstraintsSystem,
 * src/physics/physics.cpp:34-40, appends them and solvePositions :663-668
 * solves them in, after the contacts of the substep). Call once, before stepping.
 * The compiled reference prints from applyJointAxisConstraint (:529-531, a
 * debug printf left in the snapshot): the harness points stdout at /dev/null
 * while it steps worlds that hold hinges. */
void orc_set_joints(void *h, int32_t max_joints_per_world, const int32_t *joints_per_world,
                    const orc_joint *joints /*[W][max_joints_per_world]*/);
const char *orc_kind(void);

#ifdef __cplusplus
}
#endif
