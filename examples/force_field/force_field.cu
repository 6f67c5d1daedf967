// OUT-OF-TREE world with the rigid-body step as a node of its task graph (the reference wires
// RigidBodyPhysicsSystem::setup*Tasks into the user's builder, src/physics/physics.cpp:1142-1205, so a user
// system can sit on either side of it in the same run()):
//     forceSystem  (writes ExternalForce from Position -- an "action")
//  -> RigidBodyPhysicsSystem                        (the fused B200 step kernel on this archetype's columns)
//  -> obsSystem    (packs Position + Velocity into an exported observation)
// One archetype Body = the twelve physics components + Obs, 65 rows per world in collisions-64.
#include <madrona/b200/mw_gpu.hpp>

using namespace madrona;
using namespace madrona::base;
using namespace madrona::phys;

namespace ForceField {

struct Obs { float pos[3]; float vz; };
struct Body : public RigidBody<Obs> {};
struct Config { float stiffness; };
struct Init { int32_t numBodies; };

class Engine;
struct Sim : public WorldBase {
    static void registerTypes(ECSRegistry &registry, const Config &cfg);
    static void setupTasks(TaskGraph::Builder &builder, const Config &cfg);
    __device__ Sim(Engine &ctx, const Config &cfg, const Init &init);
    float stiffness;
    static inline TaskGraph::NodeID physicsNode { -1 };
    static inline int32_t rowsPerWorld = 65;
};
class Engine : public CustomContext<Engine, Sim> {
public:
    using CustomContext::CustomContext;
};

// a spring towards the z axis on every dynamic body: one rounding per component
__device__ inline void forceSystem(Engine &ctx, Position &pos, ResponseType &resp, ExternalForce &force)
{
    if (resp != ResponseType::Dynamic) return;
    const float k = ctx.data().stiffness;
    force.x = k * pos.x;
    force.y = k * pos.y;
    force.z = 0.f;
}

__device__ inline void obsSystem(Engine &, Position &pos, Velocity &vel, Obs &obs)
{
    obs.pos[0] = pos.x; obs.pos[1] = pos.y; obs.pos[2] = pos.z;
    obs.vz = vel.linear[2];
}

void Sim::registerTypes(ECSRegistry &registry, const Config &)
{
    RigidBodyPhysicsSystem::registerTypes(registry);
    registry.registerComponent<Obs>();
    registry.registerArchetype<Body>(rowsPerWorld);
    registry.exportColumn<Body, Obs>(0);
}

void Sim::setupTasks(TaskGraph::Builder &builder, const Config &)
{
    auto force_sys = builder.addToGraph<ParallelForNode<Engine, forceSystem, Position, ResponseType, ExternalForce>>({});
    physicsNode = RigidBodyPhysicsSystem::setupTasks(builder, { force_sys });
    builder.addToGraph<ParallelForNode<Engine, obsSystem, Position, Velocity, Obs>>({ physicsNode });
}

__device__ Sim::Sim(Engine &ctx, const Config &cfg, const Init &init)
{
    stiffness = cfg.stiffness;
    for (int i = 0; i < init.numBodies; i++) ctx.makeEntityNow<Body>();     // components written by the host below
}

using Executor = TaskGraphExecutor<Engine, Sim, Config, Init>;

}  // namespace ForceField

using namespace ForceField;

extern "C" {

void *ff_create(int32_t num_worlds, int32_t num_bodies, float stiffness, const mdr_phys_config *pcfg,
                const mdr_object *objs, int32_t num_objs, const float *verts, const uint32_t *poly_words)
{
    Sim::rowsPerWorld = num_bodies;
    std::vector<Init> inits(num_worlds, Init { num_bodies });
    Executor::Config cfg;
    cfg.numWorlds = (uint32_t)num_worlds;
    cfg.numExportedBuffers = 1;
    cfg.gpuID = pcfg->gpu_id;
    Executor *exec = new Executor(cfg, Config { stiffness }, inits.data());
    exec->initPhysics<Body>(Sim::physicsNode, *pcfg, objs, num_objs, verts, poly_words);
    return exec;
}
void ff_run(void *h, int32_t steps) { for (int i = 0; i < steps; i++) ((Executor *)h)->run(); }
void *ff_ecs(void *h) { return ((Executor *)h)->handle(); }
void *ff_exported(void *h, int32_t slot) { return ((Executor *)h)->getExported(slot); }
void ff_destroy(void *h) { delete (Executor *)h; }

}
