// TaskGraph::Builder and the node types of the reference (include/madrona/taskgraph.hpp:15-135,
// taskgraph.inl:58-101; src/mw/device/sort_archetype.cpp:1281-1392 for the sort node) over the C ABI.
//   builder.addToGraph<ParallelForNode<Engine, clampSystem, Position>>({deps})
// compiles as in the reference: the system is a non-type template argument, the node's row loop is a
// device function of THIS translation unit whose address goes into the graph (mdr_ecs_add_parallel_for_fn),
// and the megakernel that calls it is instantiated here as well (mw_gpu.hpp).
#pragma once
#include <initializer_list>

#include "context.hpp"
#include "registry.hpp"

namespace madrona {

class TaskGraph {
public:
    struct NodeID { int32_t id; };

    class Builder {
    public:
        explicit Builder(mdr_ecs *ecs) : ecs_(ecs) {}
        template <typename NodeT, typename... Args>
        NodeID addToGraph(std::initializer_list<NodeID> dependencies, Args &&...args)
        {
            return NodeT::addToGraph(*this, dependencies, std::forward<Args>(args)...);
        }
        mdr_ecs *handle() const { return ecs_; }

    private:
        mdr_ecs *ecs_;
    };
};

namespace b200 {
struct Deps {
    int32_t ids[64];
    int32_t n = 0;
    Deps(std::initializer_list<TaskGraph::NodeID> d) { for (auto x : d) if (n < 64) ids[n++] = x.id; }
};
template <typename NodeT> __device__ mdr::ecs::NodeFn g_node_entry = &NodeT::entry;
template <typename NodeT> inline const void *nodeEntryAddress()
{
    mdr::ecs::NodeFn fn = nullptr;
    reqCuda(cudaMemcpyFromSymbol(&fn, g_node_entry<NodeT>, sizeof(fn)), "cudaMemcpyFromSymbol(node entry)");
    return (const void *)fn;
}
}  // namespace b200

// ParallelForNode<ContextT, Fn, Components...>: Fn(ContextT &, Components &...) for every row of every
// archetype that has the components (include/madrona/taskgraph.inl:63-71 -> state.inl:358-396).
template <typename ContextT, auto Fn, typename... ComponentTs>
class ParallelForNode {
public:
    static __device__ void call(mdr::ecs::Context &low, ComponentTs &...cs)
    {
        ContextT ctx(low);
        Fn(ctx, cs...);
    }
    static __device__ void entry(mdr::ecs::Context &low, const mdr::ecs::NodeDev &nd)
    {
        mdr::ecs::parallelForImmediate<&ParallelForNode::call, ComponentTs...>(low, nd);
    }
    static TaskGraph::NodeID addToGraph(TaskGraph::Builder &builder, std::initializer_list<TaskGraph::NodeID> dependencies)
    {
        const b200::Deps d(dependencies);
        const int32_t comps[] = { b200::TypeInfo<ComponentTs>::id... };
        return { b200::req(mdr_ecs_add_parallel_for_fn(builder.handle(), b200::nodeEntryAddress<ParallelForNode>(), comps,
                                                       (int32_t)sizeof...(ComponentTs), d.ids, d.n), "ParallelForNode") };
    }
};

template <typename ArchetypeT> class ClearTmpNode {
public:
    static TaskGraph::NodeID addToGraph(TaskGraph::Builder &builder, std::initializer_list<TaskGraph::NodeID> dependencies)
    {
        const b200::Deps d(dependencies);
        return { b200::req(mdr_ecs_add_clear_tmp(builder.handle(), b200::TypeInfo<ArchetypeT>::id, d.ids, d.n), "ClearTmpNode") };
    }
};

class ResetTmpAllocNode {
public:
    static TaskGraph::NodeID addToGraph(TaskGraph::Builder &builder, std::initializer_list<TaskGraph::NodeID> dependencies)
    {
        const b200::Deps d(dependencies);
        return { b200::req(mdr_ecs_add_reset_tmp_alloc(builder.handle(), d.ids, d.n), "ResetTmpAllocNode") };
    }
};

// SortArchetypeNode<ArchetypeT, KeyComponentT>: stable sort of every world's rows by a u32 key component
template <typename ArchetypeT, typename KeyT> class SortArchetypeNode {
public:
    static TaskGraph::NodeID addToGraph(TaskGraph::Builder &builder, std::initializer_list<TaskGraph::NodeID> dependencies,
                                        bool remap_entities = true)
    {
        const b200::Deps d(dependencies);
        return { b200::req(mdr_ecs_add_sort(builder.handle(), b200::TypeInfo<ArchetypeT>::id, b200::TypeInfo<KeyT>::id,
                                            remap_entities ? 1 : 0, d.ids, d.n), "SortArchetypeNode") };
    }
};

}  // namespace madrona
