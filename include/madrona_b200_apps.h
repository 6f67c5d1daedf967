/* madrona_b200_apps -- workload definitions of the ECS / task-graph configs
 * (BASELINE.json configs[0], [1], [3]).
 *
 * The reference's examples do not build in this snapshot, so the
 * three configs are restated as task-graph worlds. This header holds only the
 * per-row ARITHMETIC of their systems (plain structs, integer hashing, +,-,*,
 * compares, one IEEE sqrt) so that the reference-driven harness, the scalar
 * port and the CUDA build evaluate the same expressions; everything the hot
 * path is about -- tables, entity ids, queries, node order, create / destroy /
 * sort / export -- is implemented separately on each side.
 *
 * Sources restated: examples/simple_taskgraph/simple.cpp:22-35 (clampSystem),
 * examples/simple_jobs/simple.cpp:60-117 (AABB from TRS, N^2 overlap fan-out),
 * examples/fantasy_vs/fvs.cpp:111-240 (action / caster / archer / cleanup).
 */
#ifndef MADRONA_B200_APPS_H
#define MADRONA_B200_APPS_H
#include <stdint.h>

#if defined(__CUDACC__)
#define MDR_APP_FN __host__ __device__ static inline
#else
#define MDR_APP_FN static inline
#endif

enum { MDR_APP_SCRIPT = 0, MDR_APP_TASKGRAPH = 1, MDR_APP_FANOUT = 2, MDR_APP_FVS = 3,
       /* churn: not a built-in world of the library -- examples/churn/churn.cu (out of tree) and the
        * reference-driven harness only: any number of makeEntityNow / destroyEntityNow per invocation */
       MDR_APP_CHURN = 4, MDR_NUM_APPS = 5 };

/* ---- component payloads (byte layouts shared by all three sides) ---------- */
typedef struct { float x, y, z; } mdr_vec3;                 /* Position, 12 B  */
typedef struct { float w, x, y, z; } mdr_quat;              /* Rotation, 16 B  */
typedef struct { float mn[3], mx[3]; } mdr_box;             /* 24 B            */
typedef struct { uint32_t gen; int32_t id; } mdr_entity;    /* Entity, 8 B (include/madrona/ecs.hpp:12-22) */
typedef struct { mdr_entity a, b; } mdr_candidate;          /* 16 B            */
typedef struct { mdr_entity e; int32_t kind; } mdr_cleanup; /* 12 B            */

/* ---- counter-based random numbers (replaces fvs.cpp:19-25's thread_local
 * mt19937, whose stream depends on thread assignment) ----------------------- */
MDR_APP_FN uint32_t mdr_fmix(uint32_t h)
{
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}
MDR_APP_FN uint32_t mdr_hash4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    uint32_t h = mdr_fmix(a * 0x9e3779b1u + 0x7f4a7c15u);
    h = mdr_fmix(h ^ (b * 0x85ebca77u + 0x165667b1u));
    h = mdr_fmix(h ^ (c * 0xc2b2ae3du + 0x27d4eb2fu));
    h = mdr_fmix(h ^ (d * 0x27d4eb2du + 0x9e3779b9u));
    return h;
}
MDR_APP_FN float mdr_u01(uint32_t h) { return (float)(h >> 8) * (1.0f / 16777216.0f); }
MDR_APP_FN float mdr_clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); }

/* World bounds of all three configs (fvs.cpp:52-55, simple_taskgraph init). */
#define MDR_BOUNDS_MIN_X (-10.f)
#define MDR_BOUNDS_MIN_Y (-10.f)
#define MDR_BOUNDS_MIN_Z (0.f)
#define MDR_BOUNDS_MAX_X (10.f)
#define MDR_BOUNDS_MAX_Y (10.f)
#define MDR_BOUNDS_MAX_Z (10.f)

MDR_APP_FN mdr_vec3 mdr_random_position(uint32_t h)
{
    mdr_vec3 p;
    p.x = MDR_BOUNDS_MIN_X + (MDR_BOUNDS_MAX_X - MDR_BOUNDS_MIN_X) * mdr_u01(mdr_fmix(h ^ 0x1111u));
    p.y = MDR_BOUNDS_MIN_Y + (MDR_BOUNDS_MAX_Y - MDR_BOUNDS_MIN_Y) * mdr_u01(mdr_fmix(h ^ 0x2222u));
    p.z = MDR_BOUNDS_MIN_Z + (MDR_BOUNDS_MAX_Z - MDR_BOUNDS_MIN_Z) * mdr_u01(mdr_fmix(h ^ 0x3333u));
    return p;
}

/* ================= config 1: simple_taskgraph ============================== */
/* One node that moves (so that clamping has work to do), one that clamps. */
MDR_APP_FN void mdr_tg_drift(mdr_vec3 *p)
{
    float x = p->x, y = p->y;
    p->x = x + 0.05f * y;
    p->y = y - 0.05f * x;
    p->z = p->z + 0.01f;
}
MDR_APP_FN void mdr_tg_clamp(mdr_vec3 *p)   /* simple.cpp:22-35 */
{
    p->x = mdr_clampf(p->x, MDR_BOUNDS_MIN_X, MDR_BOUNDS_MAX_X);
    p->y = mdr_clampf(p->y, MDR_BOUNDS_MIN_Y, MDR_BOUNDS_MAX_Y);
    p->z = mdr_clampf(p->z, MDR_BOUNDS_MIN_Z, MDR_BOUNDS_MAX_Z);
}

/* ================= churn (immediate-mode structural calls) ================== */
/* pairs of bullets a spawner makes this tick (0..2) and the life of bullet k of the pair list */
MDR_APP_FN int32_t mdr_churn_pairs(uint32_t world, uint32_t tick, int32_t spawner_id)
{
    return (int32_t)(mdr_hash4(world, tick, (uint32_t)spawner_id, 0x51u) % 3u);
}
MDR_APP_FN int32_t mdr_churn_life(uint32_t world, uint32_t tick, int32_t spawner_id, int32_t k)
{
    return 1 + (int32_t)(mdr_hash4(world, tick, (uint32_t)spawner_id, 0x100u + (uint32_t)k) % 6u);
}

/* ================= config 2: fan-out ======================================= */
MDR_APP_FN void mdr_fan_move(mdr_vec3 *p)
{
    float x = p->x, y = p->y;
    p->x = mdr_clampf(x + 0.02f * y, MDR_BOUNDS_MIN_X, MDR_BOUNDS_MAX_X);
    p->y = mdr_clampf(y - 0.02f * x, MDR_BOUNDS_MIN_Y, MDR_BOUNDS_MAX_Y);
}
MDR_APP_FN float mdr_absf(float v) { return v < 0.f ? -v : v; }
/* World box of the +-1 cube under (rotation, translation): simple_jobs
 * simple.cpp:60-94 restated with the rotation-matrix extents. */
MDR_APP_FN void mdr_fan_box(const mdr_vec3 *p, const mdr_quat *q, mdr_box *out)
{
    float x2 = q->x * q->x, y2 = q->y * q->y, z2 = q->z * q->z;
    float xy = q->x * q->y, xz = q->x * q->z, yz = q->y * q->z;
    float wx = q->w * q->x, wy = q->w * q->y, wz = q->w * q->z;
    float r00 = 1.f - 2.f * (y2 + z2), r01 = 2.f * (xy - wz), r02 = 2.f * (xz + wy);
    float r10 = 2.f * (xy + wz), r11 = 1.f - 2.f * (x2 + z2), r12 = 2.f * (yz - wx);
    float r20 = 2.f * (xz - wy), r21 = 2.f * (yz + wx), r22 = 1.f - 2.f * (x2 + y2);
    float ex = (mdr_absf(r00) + mdr_absf(r01)) + mdr_absf(r02);
    float ey = (mdr_absf(r10) + mdr_absf(r11)) + mdr_absf(r12);
    float ez = (mdr_absf(r20) + mdr_absf(r21)) + mdr_absf(r22);
    out->mn[0] = p->x - ex; out->mn[1] = p->y - ey; out->mn[2] = p->z - ez;
    out->mx[0] = p->x + ex; out->mx[1] = p->y + ey; out->mx[2] = p->z + ez;
}
MDR_APP_FN int mdr_box_overlap(const mdr_box *a, const mdr_box *b)   /* math.hpp:999-1007, strict */
{
    return a->mn[0] < b->mx[0] && b->mn[0] < a->mx[0] && a->mn[1] < b->mx[1] && b->mn[1] < a->mx[1] &&
           a->mn[2] < b->mx[2] && b->mn[2] < a->mx[2];
}
MDR_APP_FN uint32_t mdr_fan_key(int32_t row_a, int32_t row_b) { return ((uint32_t)row_a << 16) | (uint32_t)row_b; }

/* ================= config 4: fantasy_vs ==================================== */
#define MDR_FVS_DT (1.f / 60.f)
#define MDR_FVS_MOVE_SPEED 2.0f
#define MDR_FVS_MANA_REGEN 60.f
#define MDR_FVS_CAST_TIME 0.2f
#define MDR_FVS_SHOOT_TIME 0.1f
#define MDR_FVS_CAST_COST 20.f
#define MDR_FVS_BLAST_RADIUS 4.f
#define MDR_FVS_BLAST_DAMAGE 20
#define MDR_FVS_ARROW_DAMAGE 15
#define MDR_FVS_DRAGON_HP 300
#define MDR_FVS_KNIGHT_HP 100
enum { MDR_FVS_DRAGON = 0, MDR_FVS_KNIGHT = 1 };   /* archetype ids, registration order */
enum { MDR_FVS_S_ACTION = 1, MDR_FVS_S_MOVE = 2, MDR_FVS_S_CAST = 3, MDR_FVS_S_SHOOT = 4, MDR_FVS_S_SPAWN = 5 };

/* fvs.cpp:111-152. Returns nothing; updates pos / remaining time in place. */
MDR_APP_FN void mdr_fvs_action_select(uint32_t world, uint32_t tick, int32_t id, mdr_vec3 *pos, float *remaining)
{
    if (*remaining > 0.f) { *remaining = *remaining - MDR_FVS_DT; return; }
    if (mdr_u01(mdr_hash4(world, tick, (uint32_t)id, MDR_FVS_S_ACTION)) <= 0.5f) {
        uint32_t h = mdr_hash4(world, tick, (uint32_t)id, MDR_FVS_S_MOVE);
        mdr_vec3 np;
        np.x = pos->x + (2.f * mdr_u01(mdr_fmix(h ^ 0x1111u)) - 1.f);
        np.y = pos->y + (2.f * mdr_u01(mdr_fmix(h ^ 0x2222u)) - 1.f);
        np.z = pos->z + (2.f * mdr_u01(mdr_fmix(h ^ 0x3333u)) - 1.f);
        np.x = mdr_clampf(np.x, MDR_BOUNDS_MIN_X, MDR_BOUNDS_MAX_X);
        np.y = mdr_clampf(np.y, MDR_BOUNDS_MIN_Y, MDR_BOUNDS_MAX_Y);
        np.z = mdr_clampf(np.z, MDR_BOUNDS_MIN_Z, MDR_BOUNDS_MAX_Z);
        float dx = np.x - pos->x, dy = np.y - pos->y, dz = np.z - pos->z;
        *pos = np;
#if defined(__CUDA_ARCH__)
        float len = __fsqrt_rn((dx * dx + dy * dy) + dz * dz);
        *remaining = __fdiv_rn(len, MDR_FVS_MOVE_SPEED);
#else
        float len = __builtin_sqrtf((dx * dx + dy * dy) + dz * dz);
        *remaining = len / MDR_FVS_MOVE_SPEED;
#endif
    }
}
/* fvs.cpp:154-196, first half: mana bookkeeping; returns 1 when the caster
 * fires this tick and writes the blast centre. */
MDR_APP_FN int mdr_fvs_caster(uint32_t world, uint32_t tick, int32_t id, float *remaining, float *mp, mdr_vec3 *target)
{
    *mp = *mp + MDR_FVS_MANA_REGEN * MDR_FVS_DT;
    if (*remaining > 0.f) return 0;
    if (*mp < MDR_FVS_CAST_COST) return 0;
    *mp = *mp - MDR_FVS_CAST_COST;
    *target = mdr_random_position(mdr_hash4(world, tick, (uint32_t)id, MDR_FVS_S_CAST));
    *remaining = MDR_FVS_CAST_TIME;
    return 1;
}
MDR_APP_FN int mdr_fvs_in_blast(const mdr_vec3 *target, const mdr_vec3 *pos)
{
    float dx = target->x - pos->x, dy = target->y - pos->y, dz = target->z - pos->z;
    return ((dx * dx + dy * dy) + dz * dz) <= MDR_FVS_BLAST_RADIUS * MDR_FVS_BLAST_RADIUS;
}
/* fvs.cpp:198-223: returns the dragon row hit, or -1. */
MDR_APP_FN int32_t mdr_fvs_archer(uint32_t world, uint32_t tick, int32_t id, float *remaining, int32_t *arrows,
                              int32_t num_dragons)
{
    if (*remaining > 0.f || *arrows == 0 || num_dragons == 0) return -1;
    int32_t target = (int32_t)(mdr_hash4(world, tick, (uint32_t)id, MDR_FVS_S_SHOOT) % (uint32_t)num_dragons);
    *arrows = *arrows - 1;
    *remaining = MDR_FVS_SHOOT_TIME;
    return target;
}
MDR_APP_FN uint32_t mdr_fvs_cleanup_key(uint32_t archetype, int32_t row) { return (archetype << 16) | (uint32_t)row; }
/* Stats of a respawned unit (replaces Game::Game's distributions, fvs.cpp:86-109). */
MDR_APP_FN void mdr_fvs_spawn(uint32_t world, uint32_t tick, int32_t old_id, int32_t kind, mdr_vec3 *pos, int32_t *hp,
                          float *mp, int32_t *arrows)
{
    uint32_t h = mdr_hash4(world, tick, (uint32_t)old_id, MDR_FVS_S_SPAWN);
    *pos = mdr_random_position(h);
    *hp = kind == MDR_FVS_DRAGON ? MDR_FVS_DRAGON_HP : MDR_FVS_KNIGHT_HP;
    *mp = 50.f * mdr_u01(mdr_fmix(h ^ 0x4444u));
    *arrows = 20 + (int32_t)(mdr_fmix(h ^ 0x5555u) % 21u);
}

/* ================= script app (ECS semantics under test) ==================== */
/* Immediate-mode operations applied in order to one world; the host-side
 * analogue of tests/state.cpp's create / destroy / clear sequences. */
typedef struct { int32_t op, a, b; } mdr_script_op;
enum {
    MDR_OP_CREATE = 1,      /* a = archetype 0..2, b = value tag -> handle appended   */
    MDR_OP_DESTROY = 2,     /* a = handle index (creation order)                      */
    MDR_OP_CLEAR = 3,       /* a = archetype 0..2: Context::clearArchetype             */
    MDR_OP_TMP = 4,         /* b = value tag: Context::makeTemporary<T>               */
    MDR_OP_CLEAR_TMP = 5    /* Context::clearTemporaries<T>                            */
};
MDR_APP_FN void mdr_script_payload(int32_t v, float b[3], float c[4])
{
    b[0] = (float)v; b[1] = (float)v + 0.5f; b[2] = -(float)v;
    c[0] = (float)v * 2.f; c[1] = 1.f; c[2] = (float)v * 0.25f; c[3] = -1.f;
}

#endif
