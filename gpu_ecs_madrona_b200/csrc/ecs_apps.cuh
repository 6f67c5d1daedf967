// The task-graph worlds of BASELINE.json configs[0], [1], [3] written against the
// device Context the way a user of the reference writes them: components,
// archetypes, system functions, and the func-id dispatch that the reference
// generates at JIT time (src/mw/cuda_exec.cpp:553-675) and we compile ahead of
// time. Arithmetic comes from include/madrona_b200_apps.h.
#pragma once
#include "ecs_dev.cuh"

namespace mdr {
namespace ecs {
namespace apps {

// ---- components ----
struct Position : mdr_vec3 {};
struct Rotation : mdr_quat {};
struct Box : mdr_box {};
struct HitCount { int32_t n; };
struct Candidate : mdr_candidate {};
struct SortKey { uint32_t key; };
struct Health { int32_t hp; };
struct Action { float remainingTime; };
struct Mana { float mp; };
struct Quiver { int32_t numArrows; };
struct CleanupEntity : mdr_cleanup {};
struct GameState { uint32_t tick; static constexpr int kSingletonArch = 3; };

// ---- archetypes: kId = registration order within the app ----
namespace tg { struct Sphere { static constexpr int kId = 0; }; struct Agent { static constexpr int kId = 1; }; }
namespace fan { struct Obj { static constexpr int kId = 0; }; struct CandidateTmp { static constexpr int kId = 1; }; }
namespace fvs {
struct Dragon { static constexpr int kId = 0; };
struct Knight { static constexpr int kId = 1; };
struct CleanupTracker { static constexpr int kId = 2; };
}

// func ids of the dispatch table (the reference assigns these by scanning PTX)
enum Func {
    kTgDrift = 1, kTgClamp,
    kFanMove, kFanBox, kFanOverlap, kFanConsume,
    kFvsAction, kFvsCaster, kFvsArcher, kFvsMark, kFvsDestroy, kFvsRespawn, kFvsTick,
};

// =========================== config 1: simple_taskgraph =====================
__device__ __forceinline__ void tgDrift(Context &, Position &p) { mdr_tg_drift(&p); }
__device__ __forceinline__ void tgClamp(Context &, Position &p) { mdr_tg_clamp(&p); }   // simple.cpp:22-35

// =========================== config 2: fan-out ==============================
__device__ __forceinline__ void fanMove(Context &, Position &p) { mdr_fan_move(&p); }
__device__ __forceinline__ void fanBox(Context &, Position &p, Rotation &q, Box &b) { mdr_fan_box(&p, &q, &b); }
// simple_jobs simple.cpp:97-117: every later row is tested; hits become temporaries
__device__ __forceinline__ void fanOverlap(Context &ctx, Entity &e, Box &box)
{
    const int32_t row = ctx.getLoc(e).row;
    const int32_t n = ctx.numRows<fan::Obj>();
    const Entity *ents = ctx.column<Entity>(fan::Obj::kId, 0);
    const Box *boxes = ctx.column<Box>(fan::Obj::kId, 3);
    const Box mine = box;
    for (int32_t j = row + 1; j < n; j++) {
        const Box other = boxes[j];
        if (!mdr_box_overlap(&mine, &other)) continue;
        const Loc l = ctx.makeTemporary<fan::CandidateTmp>();
        if (l.archetype == kNoArch) return;
        Candidate c;
        c.a.gen = e.gen; c.a.id = e.id;
        c.b.gen = ents[j].gen; c.b.id = ents[j].id;
        ctx.getDirect<Candidate>(1, l) = c;
        ctx.getDirect<SortKey>(2, l).key = mdr_fan_key(row, j);
    }
}
__device__ __forceinline__ void fanConsume(Context &ctx, Candidate &c)
{
    const Loc la = ctx.getLoc(Entity { c.a.gen, c.a.id });
    const Loc lb = ctx.getLoc(Entity { c.b.gen, c.b.id });
    atomicAdd(&ctx.getDirect<HitCount>(4, la).n, 1);
    atomicAdd(&ctx.getDirect<HitCount>(4, lb).n, 1);
}

// =========================== config 4: fantasy_vs ===========================
__device__ __forceinline__ uint32_t fvsTickNow(Context &ctx) { return ctx.getSingleton<GameState>().tick; }
__device__ __forceinline__ void fvsAction(Context &ctx, Entity &e, Position &p, Action &a)   // fvs.cpp:111-152
{
    mdr_fvs_action_select(ctx.world_id, fvsTickNow(ctx), e.id, &p, &a.remainingTime);
}
__device__ __forceinline__ void fvsCaster(Context &ctx, Entity &e, Action &a, Mana &m)   // fvs.cpp:154-196
{
    mdr_vec3 target;
    if (!mdr_fvs_caster(ctx.world_id, fvsTickNow(ctx), e.id, &a.remainingTime, &m.mp, &target)) return;
    // fvs.cpp:180: ctx.parallelFor(healthQuery, blast) from inside the system
    ctx.submitNested(__float_as_uint(target.x), __float_as_uint(target.y), __float_as_uint(target.z), 0u);
}
// The nested job: every (Position, Health) row, Dragon table then Knight table, all lanes.
__device__ __forceinline__ void fvsBlast(Context &ctx, uint32_t tx, uint32_t ty, uint32_t tz, uint32_t)
{
    mdr_vec3 target;
    target.x = __uint_as_float(tx); target.y = __uint_as_float(ty); target.z = __uint_as_float(tz);
    for (int arch = 0; arch < 2; arch++) {
        const Position *pos = ctx.column<Position>(arch, 1);
        Health *hp = ctx.column<Health>(arch, 2);
        // four independent loads in flight per lane before the first test (the rows are independent; one load,
        // one test, one atomic at a time makes every 32-row slice wait a full memory latency)
        const int32_t n = ctx.numRows(arch);
        for (int32_t base = laneId(); base < n; base += 128) {
            mdr_vec3 p[4];
#pragma unroll
            for (int k = 0; k < 4; k++) if (base + 32 * k < n) p[k] = pos[base + 32 * k];
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (base + 32 * k < n && mdr_fvs_in_blast(&target, &p[k])) atomicSub(&hp[base + 32 * k].hp, MDR_FVS_BLAST_DAMAGE);
        }
    }
}
__device__ __forceinline__ void fvsArcher(Context &ctx, Entity &e, Action &a, Quiver &q)   // fvs.cpp:198-223
{
    const int32_t target = mdr_fvs_archer(ctx.world_id, fvsTickNow(ctx), e.id, &a.remainingTime, &q.numArrows,
                                          ctx.numRows<fvs::Dragon>());
    if (target < 0) return;
    atomicSub(&ctx.column<Health>(fvs::Dragon::kId, 2)[target].hp, MDR_FVS_ARROW_DAMAGE);
}
__device__ __forceinline__ void fvsMark(Context &ctx, Entity &e, Health &h)   // fvs.cpp:225-231
{
    if (h.hp > 0) return;
    const Loc at = ctx.getLoc(e);
    const Loc l = ctx.makeTemporary<fvs::CleanupTracker>();
    if (l.archetype == kNoArch) return;
    CleanupEntity c;
    c.e.gen = e.gen; c.e.id = e.id; c.kind = (int32_t)at.archetype;
    ctx.getDirect<CleanupEntity>(1, l) = c;
    ctx.getDirect<SortKey>(2, l).key = mdr_fvs_cleanup_key(at.archetype, at.row);
}
__device__ __forceinline__ void fvsDestroy(Context &ctx, CleanupEntity &c)   // fvs.cpp:233-236
{
    ctx.destroyEntityNow(Entity { c.e.gen, c.e.id });
}
__device__ __forceinline__ void fvsRespawn(Context &ctx, CleanupEntity &c)
{
    mdr_vec3 p; int32_t hp, arrows; float mp;
    mdr_fvs_spawn(ctx.world_id, fvsTickNow(ctx), c.e.id, c.kind, &p, &hp, &mp, &arrows);
    Position pos; pos.x = p.x; pos.y = p.y; pos.z = p.z;
    if (c.kind == MDR_FVS_DRAGON) ctx.makeEntityNow<fvs::Dragon>(pos, Health { hp }, Action { 0.f }, Mana { mp });
    else ctx.makeEntityNow<fvs::Knight>(pos, Health { hp }, Action { 0.f }, Quiver { arrows });
}
__device__ __forceinline__ void fvsTick(Context &, GameState &g) { g.tick += 1; }

// The generated `switch (func_id)` of the reference's megakernel
// (src/mw/cuda_exec.cpp:553-564), written out.
__device__ __forceinline__ void dispatchParallelFor(Context &ctx, const NodeDev &nd)
{
    switch (nd.func) {
    case kTgDrift: parallelFor<tgDrift, Position>(ctx, nd); break;
    case kTgClamp: parallelFor<tgClamp, Position>(ctx, nd); break;
    case kFanMove: parallelFor<fanMove, Position>(ctx, nd); break;
    case kFanBox: parallelFor<fanBox, Position, Rotation, Box>(ctx, nd); break;
    case kFanOverlap: parallelFor<fanOverlap, Entity, Box>(ctx, nd); break;
    case kFanConsume: parallelFor<fanConsume, Candidate>(ctx, nd); break;
    case kFvsAction: parallelFor<fvsAction, Entity, Position, Action>(ctx, nd); break;
    case kFvsCaster: parallelForNested<fvsCaster, fvsBlast, Entity, Action, Mana>(ctx, nd); break;
    case kFvsArcher: parallelFor<fvsArcher, Entity, Action, Quiver>(ctx, nd); break;
    case kFvsMark: parallelFor<fvsMark, Entity, Health>(ctx, nd); break;
    case kFvsDestroy: parallelFor<fvsDestroy, CleanupEntity>(ctx, nd); break;
    case kFvsRespawn: parallelFor<fvsRespawn, CleanupEntity>(ctx, nd); break;
    case kFvsTick: parallelFor<fvsTick, GameState>(ctx, nd); break;
    default: break;
    }
}

}  // namespace apps
}  // namespace ecs
}  // namespace mdr
