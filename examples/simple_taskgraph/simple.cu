// OUT-OF-TREE world: examples/simple_taskgraph of the reference (simple.hpp / simple.cpp:22-60, mw_cpu.cpp)
// written against the reference's own surface -- WorldT::registerTypes / setupTasks, systems as non-type
// template arguments of ParallelForNode, a world constructor that makes entities through the Context --
// and compiled by nvcc against include/madrona/b200 + libmadrona_b200.so. Nothing of this file is known
// to the library: the systems below are linked into the megakernel instantiated in this translation unit.
// BASELINE.json configs[0] as SURVEY 8(d) restates it: 100 Sphere{Position, Rotation} + 1 Agent per world,
// a drift system so that clamping has work to do, clampSystem, Agent Position / Rotation exported.
#include <madrona/b200/mw_gpu.hpp>

using namespace madrona;

namespace SimpleTaskgraph {

struct Position { float x, y, z; };
struct Rotation { float w, x, y, z; };
struct Sphere : public Archetype<Position, Rotation> {};
struct Agent : public Archetype<Position, Rotation> {};

struct AABB { Position pMin, pMax; };
struct Config { int32_t numSpheres; };
struct Init { const Position *positions; };    // numSpheres + 1 starting positions of this world (device memory)

class Engine;
struct SimpleSim : public WorldBase {
    static void registerTypes(ECSRegistry &registry, const Config &cfg);
    static void setupTasks(TaskGraph::Builder &builder, const Config &cfg);
    __device__ SimpleSim(Engine &ctx, const Config &cfg, const Init &init);
    AABB worldBounds;
    Entity agent;
};
class Engine : public CustomContext<Engine, SimpleSim> {
public:
    using CustomContext::CustomContext;
};

__device__ inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (hi < v ? hi : v); }   // std::clamp

__device__ inline void driftSystem(Engine &, Position &position)
{
    const float x = position.x, y = position.y;
    position.x = x + 0.05f * y;
    position.y = y - 0.05f * x;
    position.z = position.z + 0.01f;
}

// simple.cpp:22-35
__device__ inline void clampSystem(Engine &ctx, Position &position)
{
    position.x = clampf(position.x, ctx.data().worldBounds.pMin.x, ctx.data().worldBounds.pMax.x);
    position.y = clampf(position.y, ctx.data().worldBounds.pMin.y, ctx.data().worldBounds.pMax.y);
    position.z = clampf(position.z, ctx.data().worldBounds.pMin.z, ctx.data().worldBounds.pMax.z);
}

void SimpleSim::registerTypes(ECSRegistry &registry, const Config &cfg)
{
    registry.registerComponent<Position>();
    registry.registerComponent<Rotation>();
    registry.registerArchetype<Sphere>(cfg.numSpheres);
    registry.registerArchetype<Agent>(1);
    registry.exportColumn<Agent, Position>(0);      // simple.cpp:46-47
    registry.exportColumn<Agent, Rotation>(1);
}

void SimpleSim::setupTasks(TaskGraph::Builder &builder, const Config &)
{
    auto drift_sys = builder.addToGraph<ParallelForNode<Engine, driftSystem, Position>>({});
    builder.addToGraph<ParallelForNode<Engine, clampSystem, Position>>({ drift_sys });
}

__device__ SimpleSim::SimpleSim(Engine &ctx, const Config &cfg, const Init &init)
{
    worldBounds = AABB { { -10.f, -10.f, 0.f }, { 10.f, 10.f, 10.f } };
    for (int i = 0; i < cfg.numSpheres; i++)
        ctx.makeEntityNow<Sphere>(init.positions[i], Rotation { 1.f, 0.f, 0.f, 0.f });
    agent = ctx.makeEntityNow<Agent>(init.positions[cfg.numSpheres], Rotation { 1.f, 0.f, 0.f, 0.f });
}

using Executor = TaskGraphExecutor<Engine, SimpleSim, Config, Init>;

}  // namespace SimpleTaskgraph

// ---- what mw_cpu.cpp / gpu.cpp of the example do, behind three C calls for the test harness ----
using namespace SimpleTaskgraph;
struct Demo { Executor *exec; int32_t num_worlds; Position *d_pos; };

extern "C" {

// positions: [num_worlds][num_spheres + 1][3] on the host
void *simple_create(int32_t num_worlds, int32_t num_spheres, int32_t first_world, const float *positions)
{
    Demo *d = new Demo;
    d->num_worlds = num_worlds;
    const size_t per = (size_t)num_spheres + 1;
    cudaMalloc((void **)&d->d_pos, sizeof(Position) * per * num_worlds);
    cudaMemcpy(d->d_pos, positions, sizeof(Position) * per * num_worlds, cudaMemcpyHostToDevice);
    std::vector<Init> inits(num_worlds);
    for (int w = 0; w < num_worlds; w++) inits[w].positions = d->d_pos + per * w;
    Executor::Config cfg;
    cfg.numWorlds = (uint32_t)num_worlds;
    cfg.numExportedBuffers = 2;
    cfg.firstWorld = (uint32_t)first_world;
    d->exec = new Executor(cfg, Config { num_spheres }, inits.data());
    return d;
}
void simple_run(void *h, int32_t steps) { for (int i = 0; i < steps; i++) ((Demo *)h)->exec->run(); }
void *simple_ecs(void *h) { return ((Demo *)h)->exec->handle(); }
void simple_destroy(void *h)
{
    Demo *d = (Demo *)h;
    delete d->exec;
    cudaFree(d->d_pos);
    delete d;
}

}
