// The executor a world is handed to: TaskGraphExecutor<ContextT, WorldT, ConfigT, InitT> of the
// reference (include/madrona/mw_cpu.hpp:50-81, mw_cpu.inl:5-68; MWCudaExecutor include/madrona/
// mw_gpu.hpp:55-76), on one B200. The user's translation unit includes this header, defines
//   struct MyWorld : madrona::WorldBase {
//       static void registerTypes(madrona::ECSRegistry &, const Config &);
//       static void setupTasks(madrona::TaskGraph::Builder &, const Config &);
//       __device__ MyWorld(Engine &ctx, const Config &, const Init &);
//   };
// and is compiled with nvcc for sm_100a: the task-graph megakernel and the world-constructor kernel are
// instantiated HERE, next to the user's systems (the reference JIT-links them at run time,
// src/mw/cuda_exec.cpp:553-818, entry kernels src/mw/device/include/madrona/mw_gpu_entry.hpp:12-101),
// and linked against libmadrona_b200.so, which owns tables, id maps, exports and the run graph.
#pragma once
#include <new>
#include <vector>

#include "physics.hpp"
#include "taskgraph.hpp"

namespace madrona {
namespace b200 {

// initWorlds (mw_gpu_entry.hpp:38-62): one warp per world, lane 0 constructs the world in place.
template <typename ContextT, typename WorldT, typename ConfigT, typename InitT>
__global__ void __launch_bounds__(32 * mdr::ecs::kWarpsPerBlock)
initWorldsKernel(const mdr::ecs::EcsDev *__restrict__ E, const ConfigT *cfg, const InitT *user_inits)
{
    using namespace mdr::ecs;
    __shared__ WarpShared shared[kWarpsPerBlock];
    WarpShared *ws = &shared[threadIdx.x >> 5];
    const int warps = gridDim.x * kWarpsPerBlock;
    mdr::ecs::Context low;
    for (int w = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < E->num_worlds; w += warps) {
        loadWorld(low, E, ws, w);
        if (laneId() == 0) {
            low.serial = 1u;
            ContextT ctx(low);
            new (low.worldData()) WorldT(ctx, *cfg, user_inits[w]);
        }
        __syncwarp();
        storeWorld(low);
    }
}

template <int kMinBlocks> inline int32_t launchGraphKernel(const void *dev, int32_t first, int32_t last, int32_t num_worlds,
                                                           void *stream, void *)
{
    auto kernel = mdr::ecs::ecsGraphKernel<kMinBlocks, mdr::ecs::NoBuiltin>;
    static int blocks_cap = 0;
    if (blocks_cap == 0) {
        int dev_id = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev_id);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev_id);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 32 * mdr::ecs::kWarpsPerBlock, 0);
        blocks_cap = (per_sm > 0 ? per_sm : 1) * (sms > 0 ? sms : 1);
    }
    const int need = (num_worlds + mdr::ecs::kWarpsPerBlock - 1) / mdr::ecs::kWarpsPerBlock;
    kernel<<<need < blocks_cap ? need : blocks_cap, 32 * mdr::ecs::kWarpsPerBlock, 0, (cudaStream_t)stream>>>(
        (const mdr::ecs::EcsDev *)dev, first, last);
    return cudaGetLastError() == cudaSuccess ? MDR_OK : MDR_ESTATE;
}

}  // namespace b200

template <typename ContextT, typename WorldT, typename ConfigT, typename InitT>
class TaskGraphExecutor {
public:
    struct Config {
        uint32_t numWorlds;
        uint32_t numExportedBuffers = 0;        // informational: slots are sized by exportColumn calls
        int32_t gpuID = 0;
        uint32_t firstWorld = 0;                // global index of world 0 (worlds sharded over GPUs)
        uint32_t defaultRowsPerWorld = 256;     // row capacity of archetypes registered without one
        uint32_t tmpBytesPerWorld = 0;          // Context::tmpAlloc arena
    };

    TaskGraphExecutor(const Config &cfg, const ConfigT &user_cfg, const InitT *user_inits)
    {
        mdr_ecs_config ec {};
        ec.num_worlds = (int32_t)cfg.numWorlds;
        ec.first_world = (int32_t)cfg.firstWorld;
        ec.gpu_id = cfg.gpuID;
        b200::req(mdr_ecs_create(&ec, &ecs_), "mdr_ecs_create");
        b200::reqCuda(cudaSetDevice(cfg.gpuID), "cudaSetDevice");
        ECSRegistry registry(ecs_, (int32_t)cfg.defaultRowsPerWorld);
        WorldT::registerTypes(registry, user_cfg);
        TaskGraph::Builder builder(ecs_);
        WorldT::setupTasks(builder, user_cfg);
        b200::req(mdr_ecs_set_world_data_bytes(ecs_, (int32_t)sizeof(WorldT)), "world data");
        b200::req(mdr_ecs_set_tmp_bytes(ecs_, (int32_t)cfg.tmpBytesPerWorld), "tmp arena");
        b200::req(mdr_ecs_set_step_launcher(ecs_, &b200::launchGraphKernel<4>, nullptr), "step launcher");
        b200::req(mdr_ecs_build(ecs_), "mdr_ecs_build");

        // world constructors on the device (user inits are copied and may be freed by the caller,
        // src/mw/cuda_exec.cpp:1149-1157,1299-1300)
        cudaStream_t stream = (cudaStream_t)mdr_ecs_stream(ecs_);
        ConfigT *d_cfg = nullptr;
        InitT *d_inits = nullptr;
        b200::reqCuda(cudaMalloc((void **)&d_cfg, sizeof(ConfigT)), "cudaMalloc");
        b200::reqCuda(cudaMalloc((void **)&d_inits, sizeof(InitT) * cfg.numWorlds), "cudaMalloc");
        b200::reqCuda(cudaMemcpyAsync(d_cfg, &user_cfg, sizeof(ConfigT), cudaMemcpyHostToDevice, stream), "copy config");
        b200::reqCuda(cudaMemcpyAsync(d_inits, user_inits, sizeof(InitT) * cfg.numWorlds, cudaMemcpyHostToDevice, stream),
                      "copy inits");
        const int blocks = ((int)cfg.numWorlds + mdr::ecs::kWarpsPerBlock - 1) / mdr::ecs::kWarpsPerBlock;
        b200::initWorldsKernel<ContextT, WorldT, ConfigT, InitT>
            <<<blocks < 1184 ? blocks : 1184, 32 * mdr::ecs::kWarpsPerBlock, 0, stream>>>(
                (const mdr::ecs::EcsDev *)mdr_ecs_device_descriptor(ecs_), d_cfg, d_inits);
        b200::reqCuda(cudaGetLastError(), "initWorlds");
        b200::reqCuda(cudaStreamSynchronize(stream), "initWorlds");
        cudaFree(d_cfg);
        cudaFree(d_inits);
        b200::req(mdr_ecs_tables_written(ecs_), "tables written");
        b200::req(mdr_ecs_sync(ecs_), "world constructors");
    }
    ~TaskGraphExecutor()
    {
        for (mdr_phys *p : physics_) mdr_phys_destroy(p);
        mdr_ecs_destroy(ecs_);
    }
    TaskGraphExecutor(const TaskGraphExecutor &) = delete;

    // RigidBodyPhysicsSystem::init (src/physics/physics.cpp:1012-1036) for archetype ArchetypeT whose
    // physics node is `node` (returned by RigidBodyPhysicsSystem::setupTasks in setupTasks()).
    template <typename ArchetypeT>
    void initPhysics(TaskGraph::NodeID node, const mdr_phys_config &cfg, const mdr_object *objs, int32_t num_objs,
                     const float *verts, const uint32_t *poly_words)
    {
        mdr_phys *p = nullptr;
        b200::req(mdr_phys_create_on_ecs(&cfg, objs, num_objs, verts, poly_words, ecs_, b200::TypeInfo<ArchetypeT>::id, &p),
                  "mdr_phys_create_on_ecs");
        b200::req(mdr_ecs_attach_physics(ecs_, node.id, p), "mdr_ecs_attach_physics");
        b200::req(mdr_phys_flag_rebuild(p), "bvh.rebuildOnUpdate");
        b200::req(mdr_phys_sync(p), "mdr_phys_sync");
        physics_.push_back(p);
    }

    inline void run() { b200::req(mdr_ecs_step(ecs_, 1), "run"); }
    inline void *getExported(CountT slot) const { return mdr_ecs_get_exported(ecs_, slot); }
    mdr_ecs *handle() const { return ecs_; }

private:
    mdr_ecs *ecs_ = nullptr;
    std::vector<mdr_phys *> physics_;
};

}  // namespace madrona
