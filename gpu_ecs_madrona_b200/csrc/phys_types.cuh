// Device-visible tables of the rigid-body world step.
//
// HBM layout: every component of the rigid-body archetype is one allocation
// [num_worlds][stride][component bytes] in the reference's packed byte layout
// (Position 12 B, Rotation 16 B w,x,y,z, Velocity 24 B, ... include/madrona/
// physics.hpp:150-171,402-415, components.hpp:9-29), i.e. the reference's
// fixed-size-archetype table (include/madrona/state.inl:481-496) made the
// default: world w's rows are the contiguous tile [w*stride, w*stride+rows).
// Exported columns are handed out in place (zero copy).
#pragma once
#include "phys_math.cuh"

namespace mdr {

// One shared collision object (reference: ObjectManager rows,
// include/madrona/physics.hpp:243-290). Hull tables are flat arrays indexed by
// the offsets below (the reference's HalfEdgeMesh, physics.hpp:54-146).
struct ObjDev {
    int32_t type;            // 2 hull, 4 plane (CollisionPrimitive::Type)
    float inv_inertia[3];
    float inv_mass, mu_s, mu_d;
    float aabb[6];           // min xyz, max xyz
    int32_t vert_off, num_verts;
    int32_t face_off, num_faces;     // planes[] and polys[] share the face index
    int32_t hedge_off, num_hedges;
    int32_t edge_off, num_edges;
    float inertia[3];        // 1 / inv_inertia (0 where that is 0): substepRigidBodies divides per body and substep
                             // (physics.cpp:134-138); the quotient is a property of the object, IEEE division on the
                             // host gives the same bits
    int32_t face_pairs;      // 1: faces 2k and 2k + 1 have exactly opposite local normals (boxes, prisms): the world
                             // normal of one is the negation of the other's, bit for bit (see hullPlaneContact)
};

struct HEdgeDev { uint32_t next, twin, root, poly; };

struct AssetsDev {
    const ObjDev *objs;
    const float *verts;        // xyz packed
    const float *planes;       // nx ny nz d
    const HEdgeDev *hedges;    // indices local to the object
    const uint32_t *edges;     // half-edge index of each undirected edge
    const uint32_t *polys;     // entry half-edge of each face
    int32_t num_objs;
    int32_t max_verts, max_faces;   // over all hull objects (sizes the SMEM hull slots)
};

// The reference's Contact (physics.hpp:185-192) with Loc reduced to a row: 112 B.
struct ContactDev {
    int32_t ref, ref_pad, alt, alt_pad;
    float pts[4][4];
    int32_t num_points;
    float normal[3];
    float lambda_n[4];
};

// A contact in the form the solver consumes, produced directly by the
// narrowphase (the reference's Contact goes to the probe only). Local arms are
// computed once per substep instead of once per solver pass. For a body that is
// a stable static (see staticStable) the slot holds the world-space point
// instead of the local arm: that is all the solver needs of such a body.
// 36 words: with 16-byte accesses a quarter-warp of consecutive slots is
// bank-conflict free.
struct SolverContact {
    uint32_t hdr;            // ref row | alt row << 10 | points << 20 | s1 << 24 | s2 << 25
    float normal[3];
    float lambda[4];
    float vn_bar[4];
    float r1[4][3];
    float r2[4][3];
};

struct PhysTables {
    float *pos, *rot, *scale, *vel;
    int32_t *obj;
    uint32_t *resp;
    float *prev, *prex, *prev_vel;       // SubstepPrevState 28 B, PreSolvePositional 28 B, PreSolveVelocity 24 B
    float *ext_f, *ext_t;
    int32_t *leaf_id;
    int32_t *entity;                     // (gen, id) pairs
    int32_t *row_count;                  // [W]
    // broadphase::BVH singleton state, per world
    float *leaf_aabb;                    // [W][stride][6] current (motion-expanded) leaf boxes
    float *slot_box;                     // [W][stride][6] grow-only leaf slot boxes
    uint16_t *order;                     // [W][stride]   leaf at rank r of the reference traversal
    uint16_t *sorted;                    // [W][stride]   scratch permutation of the rebuild
    uint8_t *rebuild_flag;               // [W]
    int32_t stride;
    int32_t num_worlds;
};

// Optional per-launch probe of one world (parity tests): pair list and the
// contact list of every substep, in the reference's order.
struct ProbeDev {
    int32_t world;                // -1: off
    int32_t max_pairs;
    int32_t max_contacts;
    int32_t *num_pairs;           // [1]
    int32_t *pairs;               // [max_pairs][2]
    int32_t *num_contacts;        // [substeps]
    ContactDev *contacts;         // [substeps][max_contacts], indexed by pair slot
                                  // (num_points 0 where the pair made no contact)
};

// mdr_joint (include/madrona_b200.h), the reference's JointConstraint with the
// entities reduced to rows.
struct JointDev {
    int32_t type;                 // 0 Fixed, 1 Hinge
    int32_t row1, row2;
    float r1[3], r2[3];
    float data[12];
};

struct StepParams {
    PhysTables t;
    AssetsDev a;
    float h;                      // delta_t / substeps
    float vel_expand, accel_expand;
    float restitution_threshold;
    float g[3];
    int32_t substeps;
    int32_t max_pairs;            // SMEM pair-list capacity per world
    int32_t smem_contacts;        // contact slots held in SMEM; the rest go to `overflow`
    SolverContact *overflow;      // [resident world slots][max_pairs - smem_contacts]
    uint8_t *sat_hint;            // [W][max_pairs rounded up to 4] separating face seen last per pair slot (0: none);
                                  // null: no hints (host emulation)
    int32_t *world_pairs;         // [W] pairs found in the last step
    int32_t *world_contacts;      // [W] contacts of the last substep
    int32_t *error_flags;         // [1] bit 0: pair list overflow
    uint32_t *work;               // [kMaxLanes][2] per concurrent launch: next task of the running launch, blocks that have left it
    uint32_t *group_done;         // [W] batches only: steps (counted since creation) finished by each world group
    ProbeDev probe;
    const JointDev *joints;       // [W][joint_stride]; joint_stride 0: the batch has no joints
    const int32_t *joint_count;   // [W]
    int32_t joint_stride;
#ifdef MDR_PHASE_CLOCKS
    // profiling builds only (profiles/exp): clock64 ticks per phase of stepWorld, summed over blocks
    unsigned long long *phase_clocks;
#endif
};

constexpr int kMaxLanes = 4;       // launches of one executor that may run side by side (step_io world ranges)
enum { MDR_ERR_PAIR_OVERFLOW = 1, MDR_ERR_BAD_OBJECT = 2, MDR_ERR_STALL = 4 };

// Caps of the hull-hull path: 32 vertices / faces per slot of the warp SAT (phys_sat_warp.cuh), two slots in the
// big-hull kernel variant; also the per-thread arrays of the host emulation.
#ifndef MDR_MAX_HULL_VERTS
#define MDR_MAX_HULL_VERTS 64
#endif
#ifndef MDR_MAX_HULL_FACES
#define MDR_MAX_HULL_FACES 64
#endif
#define MDR_MAX_BODIES 1024      // rows are packed in 10 bits
#define MDR_MAX_LEVEL 2047        // 11 bits of the pair word

}  // namespace mdr
