// RigidBodyPhysicsSystem of the reference (include/madrona/physics.hpp:150-171,402-464, src/physics/
// physics.cpp:1055-1205) for a world on the B200 task graph: the twelve components every physics
// archetype starts with (phys::Cols order), registerTypes, and setup*Tasks. The broadphase, substep
// and cleanup node chains of the reference are ONE node here (the fused per-world step kernel on the
// archetype's own columns): user nodes sit before and after it in the same run().
#pragma once
#include "taskgraph.hpp"

namespace madrona {
namespace base {
struct Position { float x, y, z; };
struct Rotation { float w, x, y, z; };
struct Scale { float x, y, z; };
struct ObjectID { int32_t idx; };
}  // namespace base

namespace phys {
struct Velocity { float linear[3], angular[3]; };
struct ExternalForce { float x, y, z; };
struct ExternalTorque { float x, y, z; };
enum class ResponseType : uint32_t { Dynamic = 0, Kinematic = 1, Static = 2 };
struct LeafID { int32_t id; };
namespace solver {
struct SubstepPrevState { float prevPosition[3]; float prevRotation[4]; };
struct PreSolvePositional { float x[3]; float q[4]; };
struct PreSolveVelocity { float v[3]; float omega[3]; };
}  // namespace solver

// struct Body : madrona::phys::RigidBody<MyObservation> {};   (user components after the twelve)
template <typename... UserComponentTs>
using RigidBody = Archetype<base::Position, base::Rotation, base::Scale, Velocity, base::ObjectID, ResponseType,
                            solver::SubstepPrevState, solver::PreSolvePositional, solver::PreSolveVelocity,
                            ExternalForce, ExternalTorque, LeafID, UserComponentTs...>;

struct RigidBodyPhysicsSystem {
    static void registerTypes(ECSRegistry &registry)
    {
        registry.registerComponent<base::Position>();
        registry.registerComponent<base::Rotation>();
        registry.registerComponent<base::Scale>();
        registry.registerComponent<Velocity>();
        registry.registerComponent<base::ObjectID>();
        registry.registerComponent<ResponseType>();
        registry.registerComponent<solver::SubstepPrevState>();
        registry.registerComponent<solver::PreSolvePositional>();
        registry.registerComponent<solver::PreSolveVelocity>();
        registry.registerComponent<ExternalForce>();
        registry.registerComponent<ExternalTorque>();
        registry.registerComponent<LeafID>();
    }
    // setupBroadphaseTasks + setupSubstepTasks + setupCleanupTasks in one call / one node
    static TaskGraph::NodeID setupTasks(TaskGraph::Builder &builder, std::initializer_list<TaskGraph::NodeID> deps)
    {
        const b200::Deps d(deps);
        return { b200::req(mdr_ecs_add_physics(builder.handle(), d.ids, d.n), "RigidBodyPhysicsSystem::setupTasks") };
    }
};
}  // namespace phys
}  // namespace madrona
