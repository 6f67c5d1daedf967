// OUT-OF-TREE world: entity churn through the reference's call-and-return structural API.
//   spawnSystem   0, 2 or 4 makeEntityNow<Bullet>() per invocation, each returned Entity used at once
//                 (getUnsafe on the new row), Context::tmpAlloc scratch
//   ageSystem     Life.ticks -= 1
//   markSystem    dead bullets -> DeadTmp temporaries (makeTemporary), keyed by the bullet's row
//   sort          DeadTmp by key (temporaries made by 32 lanes land in any order; the reference makes
//                 them in row order, so this restores its order)
//   reapSystem    two destroyEntityNow per invocation, the second one often on a stale handle
//   ClearTmpNode<DeadTmp>, ResetTmpAllocNode, tickSystem (singleton)
// The same world, statement for statement, is in the reference-driven harness (oracle/ecs_ref_harness.cpp,
// MDR_APP_CHURN): tests/test_boundary.py compares ids, generations, row order, id-map free lists, exports.
#include <madrona/b200/mw_gpu.hpp>
#include <madrona_b200_apps.h>

using namespace madrona;

namespace Churn {

struct Seed { uint32_t spawned; uint32_t checksum; };
struct Life { int32_t ticks; };
struct Buddy { Entity e; };
struct DeadRef { Entity e; Entity buddy; };
struct SortKey { uint32_t key; };
struct Tick { uint32_t tick; };

struct Spawner : public Archetype<Seed> {};
struct Bullet : public Archetype<Life, Buddy> {};
struct DeadTmp : public Archetype<DeadRef, SortKey> {};

struct Config { int32_t numSpawners; int32_t maxBullets; };
struct Init { int32_t unused; };

class Engine;
struct Sim : public WorldBase {
    static void registerTypes(ECSRegistry &registry, const Config &cfg);
    static void setupTasks(TaskGraph::Builder &builder, const Config &cfg);
    __device__ Sim(Engine &ctx, const Config &cfg, const Init &init);
};
class Engine : public CustomContext<Engine, Sim> {
public:
    using CustomContext::CustomContext;
};

__device__ inline void spawnSystem(Engine &ctx, Entity &e, Seed &s)
{
    const uint32_t world = (uint32_t)ctx.worldID().idx, tick = ctx.getSingleton<Tick>().tick;
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
__device__ inline void ageSystem(Engine &, Life &l) { l.ticks -= 1; }
__device__ inline void markSystem(Engine &ctx, Entity &e, Life &l, Buddy &b)
{
    if (l.ticks > 0) return;
    Loc at = ctx.getLoc(e);
    Loc t = ctx.makeTemporary<DeadTmp>();
    ctx.getUnsafe<DeadRef>(t) = DeadRef { e, b.e };
    ctx.getUnsafe<SortKey>(t).key = (uint32_t)at.row;
}
__device__ inline void reapSystem(Engine &ctx, DeadRef &d)
{
    ctx.destroyEntityNow(d.e);
    ctx.destroyEntityNow(d.buddy);
}
__device__ inline void tickSystem(Engine &, Tick &t) { t.tick += 1; }

void Sim::registerTypes(ECSRegistry &registry, const Config &cfg)
{
    registry.registerComponent<Seed>();
    registry.registerComponent<Life>();
    registry.registerComponent<Buddy>();
    registry.registerComponent<DeadRef>();
    registry.registerComponent<SortKey>();
    registry.registerArchetype<Spawner>(cfg.numSpawners);
    registry.registerArchetype<Bullet>(cfg.maxBullets);
    registry.registerArchetype<DeadTmp>(cfg.maxBullets);
    registry.registerSingleton<Tick>();
    registry.exportColumn<Spawner, Seed>(0);
    registry.exportColumn<Bullet, Life>(1);
}

void Sim::setupTasks(TaskGraph::Builder &builder, const Config &)
{
    auto n0 = builder.addToGraph<ParallelForNode<Engine, spawnSystem, Entity, Seed>>({});
    auto n1 = builder.addToGraph<ParallelForNode<Engine, ageSystem, Life>>({ n0 });
    auto n2 = builder.addToGraph<ParallelForNode<Engine, markSystem, Entity, Life, Buddy>>({ n1 });
    auto ns = builder.addToGraph<SortArchetypeNode<DeadTmp, SortKey>>({ n2 }, false);
    auto n3 = builder.addToGraph<ParallelForNode<Engine, reapSystem, DeadRef>>({ ns });
    auto n4 = builder.addToGraph<ClearTmpNode<DeadTmp>>({ n3 });
    auto n5 = builder.addToGraph<ResetTmpAllocNode>({ n4 });
    builder.addToGraph<ParallelForNode<Engine, tickSystem, Tick>>({ n5 });
}

__device__ Sim::Sim(Engine &ctx, const Config &cfg, const Init &)
{
    ctx.getSingleton<Tick>().tick = 0;
    for (int i = 0; i < cfg.numSpawners; i++) {
        Entity s = ctx.makeEntityNow<Spawner>();
        ctx.getUnsafe<Seed>(s) = Seed { 0u, 0u };
    }
}

using Executor = TaskGraphExecutor<Engine, Sim, Config, Init>;

}  // namespace Churn

using namespace Churn;

extern "C" {

void *churn_create(int32_t num_worlds, int32_t num_spawners, int32_t max_bullets, int32_t first_world)
{
    std::vector<Init> inits(num_worlds, Init { 0 });
    Executor::Config cfg;
    cfg.numWorlds = (uint32_t)num_worlds;
    cfg.firstWorld = (uint32_t)first_world;
    cfg.tmpBytesPerWorld = 256u * (uint32_t)num_spawners;     // one 256-byte grain per spawner and tick
    return new Executor(cfg, Config { num_spawners, max_bullets }, inits.data());
}
void churn_run(void *h, int32_t steps) { for (int i = 0; i < steps; i++) ((Executor *)h)->run(); }
void *churn_ecs(void *h) { return ((Executor *)h)->handle(); }
void churn_destroy(void *h) { delete (Executor *)h; }

}
