// TEST INFRASTRUCTURE ONLY -- drives the UNMODIFIED reference ECS state layer
// and task graph (/root/reference/src/core + src/common) behind oracle/ecs_api.h.
// No reference source is copied: oracle/Makefile compiles the reference .cpp
// files where they lie and links them with this file into
// oracle/_ref/libmadrona_ecs_ref.so.
//
// One single-world StateManager / StateCache / Context / TaskGraph per world,
// the arrangement of the reference's GPU backend (one StateManager per world,
// src/mw/device/include/madrona/mw_gpu_entry.hpp:29-31) and of
// include/madrona/mw_cpu.inl:5-68. Systems are written against the reference's
// own API (Context, ParallelForNode, ClearTmpNode, makeEntityNow, ...); their
// arithmetic comes from include/madrona_b200_apps.h.
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <atomic>
#include <array>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>

#include <madrona/taskgraph.hpp>
#include <madrona/hashmap.hpp>
#include <madrona/custom_context.hpp>
#include <madrona/components.hpp>

#include "worker_init.hpp"   // -I/root/reference/src/core

#include "ecs_api.h"

using namespace madrona;

namespace {

// ---- components (layouts of include/madrona_b200_apps.h) ----
// Type ids are process-global in the reference (src/common/type_tracker.cpp:126-168) while every
// StateManager counts from its own next id, so two apps in one process must not share a
// component type: shared layouts are tagged per app.
template <int App> struct PositionT : mdr_vec3 {};
template <int App> struct RotationT : mdr_quat {};
template <int App> struct SortKeyT { uint32_t key; };
struct Box : mdr_box {};
struct HitCount { int32_t n; };
struct Candidate : mdr_candidate {};
struct Health { int32_t hp; };
struct Action { float remainingTime; };
struct Mana { float mp; };
struct Quiver { int32_t numArrows; };
struct CleanupEntity : mdr_cleanup {};
struct GameState { uint32_t tick; };
struct CompA { uint32_t v; };
struct CompB { float v[3]; };
struct alignas(16) CompC { float v[4]; };
struct ScriptSingleton { int32_t v; };
struct Seed { uint32_t spawned; uint32_t checksum; };
struct Life { int32_t ticks; };
struct Buddy { Entity e; };
struct DeadRef { Entity e; Entity buddy; };
struct ChurnTick { uint32_t tick; };

// ---- archetypes ----
struct Sphere : Archetype<PositionT<1>, RotationT<1>> {};
struct Agent : Archetype<PositionT<1>, RotationT<1>> {};
struct Obj : Archetype<PositionT<2>, RotationT<2>, Box, HitCount> {};
struct CandidateTmp : Archetype<Candidate, SortKeyT<2>> {};
struct Dragon : Archetype<PositionT<3>, Health, Action, Mana> {};
struct Knight : Archetype<PositionT<3>, Health, Action, Quiver> {};
struct CleanupTracker : Archetype<CleanupEntity, SortKeyT<3>> {};
struct ArchX : Archetype<CompA> {};
struct ArchY : Archetype<CompA, CompB> {};
struct ArchZ : Archetype<CompA, CompB, CompC> {};
struct ArchT : Archetype<CompA> {};
struct Spawner : Archetype<Seed> {};
struct Bullet : Archetype<Life, Buddy> {};
struct DeadTmp : Archetype<DeadRef, SortKeyT<4>> {};

struct Sim;
class Engine : public CustomContext<Engine, Sim> {
    using CustomContext::CustomContext;
};
struct Sim : WorldBase {
    Sim(Engine &ctx) : WorldBase(ctx) {}
    uint32_t world = 0;
    StateManager *sm = nullptr;
};

struct World {
    StateManager *sm = nullptr;
    StateCache *cache = nullptr;
    Sim *sim = nullptr;
    Engine *ctx = nullptr;
    TaskGraph *graph = nullptr;
    std::vector<Entity> handles;   // SCRIPT app: creation order
};

struct Export { uint32_t arch, col, bytes; };

struct Handle {
    int app = 0;
    std::vector<World> worlds;
    std::vector<uint32_t> arch_ids;   // app-local archetype index -> reference archetype id
    std::vector<Export> exports;
};

template <typename A> uint32_t archId() { return TypeTracker::typeID<A>(); }

Table &tableOf(StateManager *sm, uint32_t arch) { return sm->archetype_stores_[arch]->tblStorage.tbl; }
template <typename A> Table &tableOf(Engine &ctx) { return tableOf(ctx.data().sm, archId<A>()); }

// =========================== config 1 =======================================
inline void tgDrift(Engine &, PositionT<1> &p) { mdr_tg_drift(&p); }
inline void tgClamp(Engine &, PositionT<1> &p) { mdr_tg_clamp(&p); }

// =========================== config 2 =======================================
inline void fanMove(Engine &, PositionT<2> &p) { mdr_fan_move(&p); }
inline void fanBox(Engine &, PositionT<2> &p, RotationT<2> &q, Box &b) { mdr_fan_box(&p, &q, &b); }
inline void fanOverlap(Engine &ctx, Entity &e, Box &box)
{
    Table &t = tableOf<Obj>(ctx);
    const int32_t row = ctx.getLoc(e).row;
    const int32_t n = (int32_t)t.numRows();
    const Entity *ents = (const Entity *)t.data(0);
    const Box *boxes = (const Box *)t.data(3);
    for (int32_t j = row + 1; j < n; j++) {
        if (!mdr_box_overlap(&box, &boxes[j])) continue;
        Loc l = ctx.makeTemporary<CandidateTmp>();
        Candidate &c = ctx.getDirect<Candidate>(1, l);
        c.a.gen = e.gen; c.a.id = e.id;
        c.b.gen = ents[j].gen; c.b.id = ents[j].id;
        ctx.getDirect<SortKeyT<2>>(2, l).key = mdr_fan_key(row, j);
    }
}
// The reference CPU iteration emits candidates already in key order; the sort
// node of the GPU graph (orphan src/mw/device/sort_archetype.cpp) is a no-op here.
inline void fanConsume(Engine &ctx, Candidate &c)
{
    Loc la = ctx.getLoc(Entity { c.a.gen, c.a.id });
    Loc lb = ctx.getLoc(Entity { c.b.gen, c.b.id });
    ctx.getDirect<HitCount>(4, la).n += 1;
    ctx.getDirect<HitCount>(4, lb).n += 1;
}

// =========================== config 4 =======================================
inline uint32_t fvsTick(Engine &ctx) { return ctx.getSingleton<GameState>().tick; }
inline void fvsAction(Engine &ctx, Entity &e, PositionT<3> &p, Action &a)
{
    mdr_fvs_action_select(ctx.data().world, fvsTick(ctx), e.id, &p, &a.remainingTime);
}
inline void fvsCaster(Engine &ctx, Entity &e, Action &a, Mana &m)
{
    mdr_vec3 target;
    if (!mdr_fvs_caster(ctx.data().world, fvsTick(ctx), e.id, &a.remainingTime, &m.mp, &target)) return;
    // fvs.cpp:180-190: every (Position, Health) row, Dragon table then Knight table
    for (uint32_t arch : { archId<Dragon>(), archId<Knight>() }) {
        Table &t = tableOf(ctx.data().sm, arch);
        const mdr_vec3 *pos = (const mdr_vec3 *)t.data(1);
        Health *hp = (Health *)t.data(2);
        for (int32_t i = 0, n = (int32_t)t.numRows(); i < n; i++)
            if (mdr_fvs_in_blast(&target, &pos[i])) hp[i].hp -= MDR_FVS_BLAST_DAMAGE;
    }
}
inline void fvsArcher(Engine &ctx, Entity &e, Action &a, Quiver &q)
{
    Table &dragons = tableOf<Dragon>(ctx);
    int32_t target = mdr_fvs_archer(ctx.data().world, fvsTick(ctx), e.id, &a.remainingTime, &q.numArrows,
                                    (int32_t)dragons.numRows());
    if (target < 0) return;
    ((Health *)dragons.data(2))[target].hp -= MDR_FVS_ARROW_DAMAGE;
}
inline void fvsMark(Engine &ctx, Entity &e, Health &h)   // fvs.cpp:225-231
{
    if (h.hp > 0) return;
    Loc at = ctx.getLoc(e);
    Loc l = ctx.makeTemporary<CleanupTracker>();
    CleanupEntity &c = ctx.getDirect<CleanupEntity>(1, l);
    c.e.gen = e.gen; c.e.id = e.id;
    c.kind = at.archetype == archId<Dragon>() ? MDR_FVS_DRAGON : MDR_FVS_KNIGHT;
    ctx.getDirect<SortKeyT<3>>(2, l).key =
        mdr_fvs_cleanup_key(at.archetype == archId<Dragon>() ? MDR_FVS_DRAGON : MDR_FVS_KNIGHT, at.row);
}
inline void fvsDestroy(Engine &ctx, CleanupEntity &c)   // fvs.cpp:233-236
{
    ctx.destroyEntityNow(Entity { c.e.gen, c.e.id });
}
inline void fvsRespawn(Engine &ctx, CleanupEntity &c)
{
    mdr_vec3 p; int32_t hp, arrows; float mp;
    mdr_fvs_spawn(ctx.data().world, fvsTick(ctx), c.e.id, c.kind, &p, &hp, &mp, &arrows);
    PositionT<3> pos; pos.x = p.x; pos.y = p.y; pos.z = p.z;
    if (c.kind == MDR_FVS_DRAGON)
        ctx.makeEntityNow<Dragon>(pos, Health { hp }, Action { 0.f }, Mana { mp });
    else
        ctx.makeEntityNow<Knight>(pos, Health { hp }, Action { 0.f }, Quiver { arrows });
}
inline void fvsTickSystem(Engine &, GameState &g) { g.tick += 1; }

// =========================== churn ==========================================
// Any number of makeEntityNow / destroyEntityNow per invocation, the returned Entity used at once
// (getUnsafe), tmpAlloc scratch, stale handles: the world of examples/churn/churn.cu, statement for statement.
inline void churnSpawn(Engine &ctx, Entity &e, Seed &s)
{
    const uint32_t world = ctx.data().world, tick = ctx.getSingleton<ChurnTick>().tick;
    const int32_t pairs = mdr_churn_pairs(world, tick, e.id);
    uint32_t *scratch = (uint32_t *)ctx.tmpAlloc(sizeof(uint32_t) * (2 * (uint64_t)pairs + 1));
    for (int32_t k = 0; k < pairs; k++) {
        Entity a = ctx.makeEntityNow<Bullet>();
        Entity b = ctx.makeEntityNow<Bullet>();
        ctx.getUnsafe<Life>(a).ticks = mdr_churn_life(world, tick, e.id, 2 * k);
        ctx.getUnsafe<Life>(b).ticks = mdr_churn_life(world, tick, e.id, 2 * k + 1);
        ctx.getUnsafe<Buddy>(a).e = b;
        ctx.getUnsafe<Buddy>(b).e = a;
        scratch[2 * k] = (uint32_t)a.id;
        scratch[2 * k + 1] = (uint32_t)b.id;
    }
    uint32_t sum = 0;
    for (int32_t k = 0; k < 2 * pairs; k++) sum = sum * 31u + scratch[k];
    s.spawned += 2u * (uint32_t)pairs;
    s.checksum ^= sum;
}
inline void churnAge(Engine &, Life &l) { l.ticks -= 1; }
inline void churnMark(Engine &ctx, Entity &e, Life &l, Buddy &b)
{
    if (l.ticks > 0) return;
    Loc at = ctx.getLoc(e);
    Loc t = ctx.makeTemporary<DeadTmp>();
    ctx.getUnsafe<DeadRef>(t) = DeadRef { e, b.e };
    ctx.getUnsafe<SortKeyT<4>>(t).key = (uint32_t)at.row;
}
inline void churnReap(Engine &ctx, DeadRef &d)
{
    ctx.destroyEntityNow(d.e);
    ctx.destroyEntityNow(d.buddy);     // often stale already: the reference returns on an invalid Loc
}
inline void churnTickSystem(Engine &, ChurnTick &t) { t.tick += 1; }

template <auto Fn, typename... Cs> using PF = ParallelForNode<Engine, Fn, Cs...>;

void setupWorld(Handle *h, World &w, uint32_t world_idx, const int32_t *params, int32_t nparams)
{
    auto P = [&](int i, int32_t dflt) { return i < nparams ? params[i] : dflt; };
    w.sm = new StateManager();
    w.cache = new StateCache();
    ECSRegistry reg(w.sm, nullptr);
    std::vector<uint32_t> ids;
    switch (h->app) {
    case MDR_APP_SCRIPT:
        reg.registerComponent<CompA>();
        reg.registerComponent<CompB>();
        reg.registerComponent<CompC>();
        reg.registerArchetype<ArchX>();
        reg.registerArchetype<ArchY>();
        reg.registerArchetype<ArchZ>();
        reg.registerArchetype<ArchT>();
        ids = { archId<ArchX>(), archId<ArchY>(), archId<ArchZ>(), archId<ArchT>() };
        if (P(0, 1)) {
            reg.registerSingleton<ScriptSingleton>();
            ids.push_back(archId<StateManager::SingletonArchetype<ScriptSingleton>>());
        }
        break;
    case MDR_APP_TASKGRAPH:
        reg.registerComponent<PositionT<1>>();
        reg.registerComponent<RotationT<1>>();
        reg.registerArchetype<Sphere>();
        reg.registerArchetype<Agent>();
        ids = { archId<Sphere>(), archId<Agent>() };
        h->exports = { { archId<Agent>(), 1, 12 }, { archId<Agent>(), 2, 16 } };   // simple.cpp:46-47
        break;
    case MDR_APP_FANOUT:
        reg.registerComponent<PositionT<2>>();
        reg.registerComponent<RotationT<2>>();
        reg.registerComponent<Box>();
        reg.registerComponent<HitCount>();
        reg.registerComponent<Candidate>();
        reg.registerComponent<SortKeyT<2>>();
        reg.registerArchetype<Obj>();
        reg.registerArchetype<CandidateTmp>();
        ids = { archId<Obj>(), archId<CandidateTmp>() };
        h->exports = { { archId<Obj>(), 4, 4 } };
        break;
    case MDR_APP_FVS:
        reg.registerComponent<PositionT<3>>();
        reg.registerComponent<Health>();
        reg.registerComponent<Action>();
        reg.registerComponent<Mana>();
        reg.registerComponent<Quiver>();
        reg.registerComponent<CleanupEntity>();
        reg.registerComponent<SortKeyT<3>>();
        reg.registerArchetype<Dragon>();
        reg.registerArchetype<Knight>();
        reg.registerArchetype<CleanupTracker>();
        reg.registerSingleton<GameState>();
        ids = { archId<Dragon>(), archId<Knight>(), archId<CleanupTracker>(),
                archId<StateManager::SingletonArchetype<GameState>>() };
        h->exports = { { archId<Dragon>(), 2, 4 }, { archId<Knight>(), 1, 12 } };
        break;
    case MDR_APP_CHURN:
        reg.registerComponent<Seed>();
        reg.registerComponent<Life>();
        reg.registerComponent<Buddy>();
        reg.registerComponent<DeadRef>();
        reg.registerComponent<SortKeyT<4>>();
        reg.registerArchetype<Spawner>();
        reg.registerArchetype<Bullet>();
        reg.registerArchetype<DeadTmp>();
        reg.registerSingleton<ChurnTick>();
        ids = { archId<Spawner>(), archId<Bullet>(), archId<DeadTmp>(),
                archId<StateManager::SingletonArchetype<ChurnTick>>() };
        h->exports = { { archId<Spawner>(), 1, 8 }, { archId<Bullet>(), 1, 4 } };
        break;
    default:
        abort();
    }
    h->arch_ids = ids;

    WorkerInit wi { w.sm, w.cache };
    w.sim = (Sim *)malloc(sizeof(Sim));
    w.ctx = new Engine(w.sim, wi);
    new (w.sim) Sim(*w.ctx);
    w.sim->world = world_idx;
    w.sim->sm = w.sm;

    // World constructor: entities in creation order; column values are
    // written afterwards through eorc_write_column.
    Engine &ctx = *w.ctx;
    TaskGraph::Builder b(ctx);
    switch (h->app) {
    case MDR_APP_SCRIPT:
        if (P(0, 1)) ctx.getSingleton<ScriptSingleton>().v = 0;
        break;
    case MDR_APP_TASKGRAPH: {
        for (int i = 0, n = P(0, 100); i < n; i++) ctx.makeEntityNow<Sphere>();
        ctx.makeEntityNow<Agent>();
        auto n0 = b.addToGraph<PF<tgDrift, PositionT<1>>>({});
        b.addToGraph<PF<tgClamp, PositionT<1>>>({ n0 });
    } break;
    case MDR_APP_FANOUT: {
        for (int i = 0, n = P(0, 100); i < n; i++) ctx.makeEntityNow<Obj>();
        auto n0 = b.addToGraph<ClearTmpNode<CandidateTmp>>({});
        auto n1 = b.addToGraph<PF<fanMove, PositionT<2>>>({ n0 });
        auto n2 = b.addToGraph<PF<fanBox, PositionT<2>, RotationT<2>, Box>>({ n1 });
        auto n3 = b.addToGraph<PF<fanOverlap, Entity, Box>>({ n2 });
        b.addToGraph<PF<fanConsume, Candidate>>({ n3 });
    } break;
    case MDR_APP_FVS: {
        ctx.getSingleton<GameState>().tick = 0;
        for (int i = 0, n = P(0, 50); i < n; i++) ctx.makeEntityNow<Dragon>();
        for (int i = 0, n = P(1, 200); i < n; i++) ctx.makeEntityNow<Knight>();
        auto n0 = b.addToGraph<PF<fvsAction, Entity, PositionT<3>, Action>>({});
        auto n1 = b.addToGraph<PF<fvsCaster, Entity, Action, Mana>>({ n0 });
        auto n2 = b.addToGraph<PF<fvsArcher, Entity, Action, Quiver>>({ n1 });
        auto n3 = b.addToGraph<PF<fvsMark, Entity, Health>>({ n2 });
        auto n4 = b.addToGraph<PF<fvsDestroy, CleanupEntity>>({ n3 });
        auto n5 = b.addToGraph<PF<fvsRespawn, CleanupEntity>>({ n4 });
        auto n6 = b.addToGraph<ClearTmpNode<CleanupTracker>>({ n5 });
        b.addToGraph<PF<fvsTickSystem, GameState>>({ n6 });
    } break;
    case MDR_APP_CHURN: {
        ctx.getSingleton<ChurnTick>().tick = 0;
        for (int i = 0, n = P(0, 24); i < n; i++) {
            Entity s = ctx.makeEntityNow<Spawner>();
            ctx.getUnsafe<Seed>(s) = Seed { 0u, 0u };
        }
        auto n0 = b.addToGraph<PF<churnSpawn, Entity, Seed>>({});
        auto n1 = b.addToGraph<PF<churnAge, Life>>({ n0 });
        auto n2 = b.addToGraph<PF<churnMark, Entity, Life, Buddy>>({ n1 });
        auto n3 = b.addToGraph<PF<churnReap, DeadRef>>({ n2 });
        auto n4 = b.addToGraph<ClearTmpNode<DeadTmp>>({ n3 });
        auto n5 = b.addToGraph<ResetTmpAllocNode>({ n4 });
        b.addToGraph<PF<churnTickSystem, ChurnTick>>({ n5 });
    } break;
    }
    // Builder::build() dereferences the first staged node: no graph for the node-less script app.
    if (h->app != MDR_APP_SCRIPT) w.graph = new TaskGraph(b.build());
}

}  // namespace

extern "C" {

const char *eorc_kind(void) { return "reference"; }

void *eorc_create(int32_t app, int32_t num_worlds, int32_t first_world, const int32_t *params, int32_t nparams)
{
    if (app < 0 || app >= MDR_NUM_APPS || num_worlds <= 0) return nullptr;
    Handle *h = new Handle;
    h->app = app;
    h->worlds.resize(num_worlds);
    for (int i = 0; i < num_worlds; i++) setupWorld(h, h->worlds[i], (uint32_t)(first_world + i), params, nparams);
    return h;
}

void eorc_destroy(void *hp) { delete (Handle *)hp; }   // per-world reference state is leaked, as the reference does

void eorc_step(void *hp, int32_t num_steps, int32_t num_threads)
{
    Handle *h = (Handle *)hp;
    const int W = (int)h->worlds.size();
    if (h->app == MDR_APP_SCRIPT) return;
    if (num_threads <= 1) {
        for (int s = 0; s < num_steps; s++)
            for (int i = 0; i < W; i++) h->worlds[i].graph->run(h->worlds[i].ctx);
        return;
    }
    // src/mw/cpu_exec.cpp:244-284: workers pull world indices from an atomic counter.
    std::atomic<int> next { 0 };
    std::vector<std::thread> pool;
    for (int t = 0; t < num_threads; t++) {
        pool.emplace_back([&]() {
            while (true) {
                int i = next.fetch_add(1, std::memory_order_relaxed);
                if (i >= W) break;
                for (int s = 0; s < num_steps; s++) h->worlds[i].graph->run(h->worlds[i].ctx);
            }
        });
    }
    for (auto &t : pool) t.join();
}

int32_t eorc_script(void *hp, int32_t world, const mdr_script_op *ops, int32_t nops)
{
    Handle *h = (Handle *)hp;
    if (h->app != MDR_APP_SCRIPT) return -1;
    World &w = h->worlds[world];
    Engine &ctx = *w.ctx;
    for (int i = 0; i < nops; i++) {
        const mdr_script_op &o = ops[i];
        switch (o.op) {
        case MDR_OP_CREATE: {
            float b[3], c[4];
            mdr_script_payload(o.b, b, c);
            Entity e;
            if (o.a == 0) e = ctx.makeEntityNow<ArchX>(CompA { (uint32_t)o.b });
            else if (o.a == 1) e = ctx.makeEntityNow<ArchY>(CompA { (uint32_t)o.b }, CompB { { b[0], b[1], b[2] } });
            else e = ctx.makeEntityNow<ArchZ>(CompA { (uint32_t)o.b }, CompB { { b[0], b[1], b[2] } },
                                              CompC { { c[0], c[1], c[2], c[3] } });
            w.handles.push_back(e);
        } break;
        case MDR_OP_DESTROY:
            if (o.a >= 0 && o.a < (int)w.handles.size()) ctx.destroyEntityNow(w.handles[o.a]);
            break;
        case MDR_OP_CLEAR:
            if (o.a == 0) ctx.clearArchetype<ArchX>();
            else if (o.a == 1) ctx.clearArchetype<ArchY>();
            else ctx.clearArchetype<ArchZ>();
            break;
        case MDR_OP_TMP: {
            Loc l = ctx.makeTemporary<ArchT>();
            ctx.getDirect<CompA>(1, l).v = (uint32_t)o.b;
        } break;
        case MDR_OP_CLEAR_TMP:
            ctx.clearTemporaries<ArchT>();
            break;
        default:
            return -1;
        }
    }
    return 0;
}

int32_t eorc_num_archetypes(void *hp) { return (int32_t)((Handle *)hp)->arch_ids.size(); }

int32_t eorc_num_cols(void *hp, int32_t arch)
{
    Handle *h = (Handle *)hp;
    return (int32_t)tableOf(h->worlds[0].sm, h->arch_ids[arch]).num_components_;
}

int32_t eorc_col_bytes(void *hp, int32_t arch, int32_t col)
{
    Handle *h = (Handle *)hp;
    return (int32_t)tableOf(h->worlds[0].sm, h->arch_ids[arch]).bytes_per_column_[col];
}

int32_t eorc_num_rows(void *hp, int32_t world, int32_t arch)
{
    Handle *h = (Handle *)hp;
    return (int32_t)tableOf(h->worlds[world].sm, h->arch_ids[arch]).numRows();
}

void eorc_read_column(void *hp, int32_t world, int32_t arch, int32_t col, void *out)
{
    Handle *h = (Handle *)hp;
    Table &t = tableOf(h->worlds[world].sm, h->arch_ids[arch]);
    memcpy(out, t.data(col), (size_t)t.numRows() * t.bytes_per_column_[col]);
}

void eorc_write_column(void *hp, int32_t world, int32_t arch, int32_t col, const void *in)
{
    Handle *h = (Handle *)hp;
    Table &t = tableOf(h->worlds[world].sm, h->arch_ids[arch]);
    memcpy(t.data(col), in, (size_t)t.numRows() * t.bytes_per_column_[col]);
}

int32_t eorc_id_count(void *hp, int32_t world)
{
    Handle *h = (Handle *)hp;
    return (int32_t)h->worlds[world].sm->entity_store_.map_.store_.numIDs;
}

void eorc_read_idmap(void *hp, int32_t world, int32_t *a, int32_t *b, uint32_t *gen, int32_t state[10])
{
    Handle *h = (Handle *)hp;
    World &w = h->worlds[world];
    auto &map = w.sm->entity_store_.map_;
    const int n = (int)map.store_.numIDs;
    // Loc.archetype of a live entity is reported as the app-local archetype index.
    std::map<uint32_t, int32_t> local;
    for (size_t i = 0; i < h->arch_ids.size(); i++) local[h->arch_ids[i]] = (int32_t)i;
    for (int i = 0; i < n; i++) {
        auto &node = map.store_[i];
        a[i] = node.freeNode.subNext;
        b[i] = node.freeNode.globalNext;
        gen[i] = node.gen.load_relaxed();
    }
    // Translate archetype ids of live nodes (those reachable from a table row).
    for (size_t ai = 0; ai < h->arch_ids.size(); ai++) {
        Table &t = tableOf(w.sm, h->arch_ids[ai]);
        // temporaries carry no ids: their Entity column is uninitialised
        const Entity *ents = (const Entity *)t.data(0);
        for (uint32_t r = 0; r < t.numRows(); r++) {
            Entity e = ents[r];
            if (e.id < 0 || e.id >= n) continue;
            auto &node = map.store_[e.id];
            if (node.gen.load_relaxed() != e.gen) continue;
            if (node.val.archetype == h->arch_ids[ai] && node.val.row == (int32_t)r) a[e.id] = (int32_t)ai;
        }
    }
    auto cacheOut = [](const EntityStore::Cache &c, int32_t *o) {
        o[0] = c.free_head_; o[1] = c.num_free_ids_; o[2] = c.overflow_head_; o[3] = c.num_overflow_ids_;
    };
    cacheOut(w.sm->init_state_cache_.entity_cache_, state);
    cacheOut(w.cache->entity_cache_, state + 4);
    state[8] = map.free_head_.load_relaxed().head;
    state[9] = n;
}

int64_t eorc_export(void *hp, int32_t slot, void *out, int64_t max_bytes)
{
    Handle *h = (Handle *)hp;
    if (slot < 0 || slot >= (int)h->exports.size()) return -1;
    const Export &ex = h->exports[slot];
    int64_t off = 0;
    for (World &w : h->worlds) {
        Table &t = tableOf(w.sm, ex.arch);
        int64_t nb = (int64_t)t.numRows() * ex.bytes;
        if (off + nb > max_bytes) return -1;
        memcpy((char *)out + off, t.data(ex.col), nb);
        off += nb;
    }
    return off;
}

// The reference's own StaticIntegerMap<128> (include/madrona/hashmap.hpp, src/common/hashmap.cpp:17-116):
// builds the map and hands back both storage arrays, for pinning csrc/static_map.hpp. The caller keeps to
// key sets the reference can hash (it FATALs otherwise).
void eorc_static_map_build(const uint32_t *keys, const uint32_t *values, uint32_t n,
                           uint32_t *out_keys /*[128]*/, uint32_t *out_values /*[128]*/)
{
    std::vector<IntegerMapPair> in(n);
    for (uint32_t i = 0; i < n; i++) in[i] = IntegerMapPair { keys[i], values[i] };
    StaticIntegerMap<128> m(in.data(), n);
    memcpy(out_keys, m.keys_.data(), 4 * 128);
    memcpy(out_values, m.values_.data(), 4 * 128);
}

}  // extern "C"
