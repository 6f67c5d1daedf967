// ECS state layer + task-graph executor: host half (registration, query
// resolution, graph lowering, table allocation) and the kernels
//   ecsStepKernel        persistent megakernel: one warp walks the task graph of one world
//   ecsScriptKernel      immediate-mode create / destroy / clear (world constructors, tests)
//   ecsExportOffsets / ecsExportCopyOut / ecsExportCopyIn   exported-column staging
// behind the mdr_ecs_* C ABI of include/madrona_b200.h. sm_100a only.
//
// Reference interfaces restated: ECSRegistry (include/madrona/state.hpp:109-134,
// state.inl:17-215), StateManager::makeQuery (src/core/state.cpp:270-361),
// TaskGraph::Builder (src/core/taskgraph.cpp:11-122), export copy kernels
// (src/mw/device/consts.cpp:92-273), MWCudaExecutor::run (src/mw/cuda_exec.cpp:1777-1794).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/madrona_b200.h"
#include "ecs_apps.cuh"

using namespace mdr::ecs;
using namespace mdr;       // static_map.hpp

// physics.cu: one rigid-body step of `p` enqueued on `stream` (capturable), the physics node of an ECS graph
int32_t mdrPhysEnqueueStep(mdr_phys *p, cudaStream_t stream);

#define REQ_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (expr);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            fprintf(stderr, "madrona_b200: CUDA error %s at %s:%d: %s\n",                \
                    cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_));   \
            abort();                                                                     \
        }                                                                                \
    } while (0)

namespace {

// ------------------------------------------------------------------ kernels --
// The megakernel itself is mdr::ecs::ecsGraphKernel (include/madrona/b200/ecs_device.cuh); this is its
// instantiation for the built-in worlds, whose systems are named by func id (csrc/ecs_apps.cuh). A world
// whose systems live in another translation unit brings its own instantiation (mdr_ecs_set_step_launcher).
struct AppsBuiltin {
    __device__ __forceinline__ void operator()(Context &ctx, const NodeDev &nd) const { apps::dispatchParallelFor(ctx, nd); }
};
template <int kMinBlocks>
__global__ void __launch_bounds__(32 * kWarpsPerBlock, kMinBlocks)
ecsStepKernel(const EcsDev *__restrict__ E, int first, int last, int steps)
{
    __shared__ WarpShared shared[kWarpsPerBlock];
    __shared__ BlockDescriptors descriptors;
    WarpShared *ws = &shared[threadIdx.x >> 5];
    const int warps = gridDim.x * kWarpsPerBlock;
    Context ctx;
    const NodeDev *nodes = loadDescriptors(&descriptors, ctx, E, first, last);
    for (int w = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < E->num_worlds; w += warps) {
        loadWorld(ctx, E, ws, w);
        // `steps` walks of the graph back to back (mdr_ecs_step_async(e, n > 1)): no world reads another world,
        // so a warp may run its world ahead; row counts and id state stay in shared memory between the walks
        for (int k = 0; k < steps; k++) {
            if (k) beginStep(ctx);
            runNodes(ctx, nodes, last - first, AppsBuiltin());
        }
        storeWorld(ctx);
    }
}

// ops of world w are ops[offsets[w] .. offsets[w + 1]) or, with offsets == null,
// the same ops[0 .. num_shared) for every world.
__global__ void __launch_bounds__(32 * kWarpsPerBlock)
ecsScriptKernel(const EcsDev *__restrict__ E, const mdr_script_op *__restrict__ ops,
                const int32_t *__restrict__ offsets, int32_t num_shared)
{
    __shared__ WarpShared shared[kWarpsPerBlock];
    WarpShared *ws = &shared[threadIdx.x >> 5];
    const int warps = gridDim.x * kWarpsPerBlock;
    const int lane = laneId();
    const bool script_app = E->app == MDR_APP_SCRIPT;
    Context ctx;
    for (int w = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5); w < E->num_worlds; w += warps) {
        loadWorld(ctx, E, ws, w);
        const int32_t begin = offsets ? offsets[w] : 0, end = offsets ? offsets[w + 1] : num_shared;
        Entity *handles = E->handles ? E->handles + (size_t)w * E->max_handles : nullptr;
        int32_t nh = handles ? E->num_handles[w] : 0;
        int32_t pos = begin;
        while (pos < end) {
            const int32_t i = pos + lane;
            mdr_script_op op { 0, 0, 0 };
            if (i < end) op = ops[i];
            const unsigned special = __ballot_sync(kFull, op.op >= MDR_OP_CLEAR);
            int len = special ? __ffs(special) - 1 : 32;
            if (len == 0) {
                const int32_t kind = __shfl_sync(kFull, op.op, 0), a = __shfl_sync(kFull, op.a, 0),
                              b = __shfl_sync(kFull, op.b, 0);
                if (kind == MDR_OP_CLEAR) {
                    if (a >= 0 && a < E->num_arch) clearArchetype(ctx, a, false);
                } else if (kind == MDR_OP_CLEAR_TMP) {
                    clearArchetype(ctx, 3, true);
                } else if (kind == MDR_OP_TMP) {
                    if (lane == 0) {
                        const Loc l = ctx.makeTemporary(3);
                        if (l.archetype != kNoArch) ctx.column<uint32_t>(3, 1)[l.row] = (uint32_t)b;
                    }
                    __syncwarp();
                }
                pos += 1;
                continue;
            }
            bool in = lane < len && i < end;
            bool is_create = in && op.op == MDR_OP_CREATE && op.a >= 0 && op.a < E->num_arch;
            int before = __popc(__ballot_sync(kFull, is_create) & lowMask(lane));
            // a destroy naming a handle that an earlier op of this slice creates ends the slice
            const unsigned split = __ballot_sync(kFull, in && op.op == MDR_OP_DESTROY && op.a >= nh && op.a < nh + before);
            if (split) {
                len = __ffs(split) - 1;
                in = lane < len;
                is_create = is_create && in;
            }
            const bool is_destroy = in && op.op == MDR_OP_DESTROY && op.a >= 0 && op.a < nh;
            if (is_destroy) ctx.destroyEntityNow(handles[op.a]);
            if (is_create) {
                float pb[3], pc[4];
                mdr_script_payload(op.b, pb, pc);
                if (script_app) {
                    struct { uint32_t v; float b[3]; float c[4]; } row { (uint32_t)op.b, { pb[0], pb[1], pb[2] },
                                                                      { pc[0], pc[1], pc[2], pc[3] } };
                    ctx.makeEntityNowIn(op.a, row);
                } else {
                    struct { uint32_t z[kMaxPayloadWords]; } row {};
                    ctx.makeEntityNowIn(op.a, row);
                }
            }
            __syncwarp();
            applyPending(ctx);
            const unsigned creates = __ballot_sync(kFull, is_create);
            if (is_create && handles) {
                const int32_t h = nh + before;
                if (h < E->max_handles) handles[h] = ctx.created;
                else atomicOr(E->error, kErrHandleOverflow);
            }
            nh += __popc(creates);
            if (nh > E->max_handles && handles) nh = E->max_handles;
            pos += len < end - pos ? len : end - pos;
            __syncwarp();
        }
        if (handles && lane == 0) E->num_handles[w] = nh;
        storeWorld(ctx);
    }
}

// StateManager::copyOutExportedColumns (src/core/state.cpp:516-563) /
// madronaMWGPUExportBlockSums + ExportCopyOut (src/mw/device/consts.cpp:137-273):
// offsets[w] = rows of the export's archetype in worlds < w. One block per slot.
template <int kPer>
__global__ void __launch_bounds__(1024) ecsExportOffsets(const EcsDev *__restrict__ E)
{
    __shared__ int64_t warp_sums[32];
    __shared__ int64_t carry;
    const ExportDev &x = E->exports[blockIdx.x];
    if (x.fixed) return;
    const int32_t *counts = E->arch[x.arch].row_count;
    const int W = E->num_worlds, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    // kPer consecutive worlds per thread: with 8, 8192 worlds are one pass of the block (three barriers in
    // all); up to 1024 worlds one world per thread is the shorter chain (launchExportOffsets picks).
    for (int base = 0; base < W; base += 1024 * kPer) {
        const int w0 = base + (int)threadIdx.x * kPer;
        int32_t v[kPer];
        if (kPer == 8 && w0 + kPer <= W && (((uintptr_t)(counts + w0)) & 15u) == 0) {
            const int4 a = *(const int4 *)(counts + w0), b = *(const int4 *)(counts + w0 + 4);
            const int32_t t[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
#pragma unroll
            for (int k = 0; k < kPer; k++) v[k] = t[k];
        } else {
#pragma unroll
            for (int k = 0; k < kPer; k++) v[k] = w0 + k < W ? counts[w0 + k] : 0;
        }
        int64_t mine = 0;
#pragma unroll
        for (int k = 0; k < kPer; k++) mine += v[k];
        int64_t incl = mine;
        for (int o = 1; o < 32; o <<= 1) {
            const int64_t t = __shfl_up_sync(kFull, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) warp_sums[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int64_t s = warp_sums[lane], si = s;
            for (int o = 1; o < 32; o <<= 1) {
                const int64_t t = __shfl_up_sync(kFull, si, o);
                if (lane >= o) si += t;
            }
            warp_sums[lane] = si - s;
        }
        __syncthreads();
        const int64_t c = carry;
        int64_t run = c + warp_sums[wid] + incl - mine;
#pragma unroll
        for (int k = 0; k < kPer; k++) {
            if (w0 + k < W) x.offsets[w0 + k] = run;
            run += v[k];
        }
        __syncthreads();
        if (threadIdx.x == 1023) carry = run;
        __syncthreads();
    }
    if (threadIdx.x == 0) x.offsets[W] = carry;
}

static void launchExportOffsets(const EcsDev *dev, int num_exports, int num_worlds, cudaStream_t stream)
{
    if (num_worlds <= 1024) ecsExportOffsets<1><<<num_exports, 1024, 0, stream>>>(dev);
    else ecsExportOffsets<8><<<num_exports, 1024, 0, stream>>>(dev);
}

// One warp per world; every non-fixed export: tile rows -> contiguous buffer
// (to_buffer) or back (the reference's copy-in of actions before the step).
// A world's share is a few KB at an offset that is only word-aligned: words,
// four loads in flight per lane before the first store.
__device__ __forceinline__ void copyWordsWarp(uint32_t *__restrict__ dst, const uint32_t *__restrict__ src, int total, int lane)
{
    int i = lane;
    for (; i + 96 < total; i += 128) {
        const uint32_t a = src[i], b = src[i + 32], c = src[i + 64], d = src[i + 96];
        dst[i] = a; dst[i + 32] = b; dst[i + 64] = c; dst[i + 96] = d;
    }
    for (; i < total; i += 32) dst[i] = src[i];
}

__global__ void __launch_bounds__(256) ecsExportCopy(const EcsDev *__restrict__ E, int to_buffer)
{
    const int w = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= E->num_worlds) return;
    for (int s = 0; s < E->num_exports; s++) {
        const ExportDev &x = E->exports[s];
        if (x.fixed) continue;
        const ArchDev &A = E->arch[x.arch];
        const int64_t off = x.offsets[w];
        const int32_t n = (int32_t)(x.offsets[w + 1] - off);
        const int words = x.bytes >> 2;
        uint32_t *tile = (uint32_t *)(A.col[x.col] + (size_t)w * A.stride * x.bytes);
        uint32_t *buf = (uint32_t *)x.buffer + off * words;
        const int total = n * words;
        if (to_buffer) copyWordsWarp(buf, tile, total, lane);
        else copyWordsWarp(tile, buf, total, lane);
    }
}

// ---- the whole step in ONE launch (every world resident at once: num_worlds <= grid x 8 warps) ----------
// export copy-in -> the nodes of the graph -> GRID BARRIER -> export offsets -> export copy-out, one warp per
// world. What the four launches of the run graph do (MWCudaExecutor's graph, src/mw/cuda_exec.cpp:1519-1690),
// without the three launch boundaries: at 1024 worlds a step is a few microseconds of work and the boundaries
// are most of it. The barrier is the only cross-world point of the step: the offsets are a prefix sum over the
// row counts of ALL worlds. It is a counter barrier over co-resident blocks (the grid is never larger than the
// resident capacity of the device: mdr_ecs_build sizes it).
__device__ __forceinline__ void gridBarrier(unsigned *bar /*[2]: arrivals, generation*/)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned gen = *(volatile unsigned *)(bar + 1);
        if (atomicAdd(bar, 1u) == gridDim.x - 1) {
            *(volatile unsigned *)bar = 0u;
            __threadfence();
            atomicAdd(bar + 1, 1u);
        } else {
            while (*(volatile unsigned *)(bar + 1) == gen) __nanosleep(100);   // leave the issue slots to the blocks still at work
        }
        __threadfence();
    }
    __syncthreads();
}

template <int kMinBlocks>
__global__ void __launch_bounds__(32 * kWarpsPerBlock, kMinBlocks)
ecsFusedStepKernel(const EcsDev *__restrict__ E, int first, int last, unsigned *bar, int steps, int align)
{
    __shared__ WarpShared shared[kWarpsPerBlock];
    __shared__ BlockDescriptors descriptors;
    __shared__ long long s_warp_sum[kWarpsPerBlock];
    __shared__ long long s_base;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int W = E->num_worlds;
    const int w = blockIdx.x * kWarpsPerBlock + warp;
    const bool active = w < W;
    Context ctx;
    const NodeDev *nodes = loadDescriptors(&descriptors, ctx, E, first, last);
    // 1. copy-in at the offsets the previous step (or refreshExports) left
    if (active) {
        for (int s = 0; s < E->num_exports; s++) {
            const ExportDev &x = E->exports[s];
            if (x.fixed) continue;
            const ArchDev &A = E->arch[x.arch];
            const int64_t off = x.offsets[w];
            const int total = (int)(x.offsets[w + 1] - off) * (x.bytes >> 2);
            copyWordsWarp((uint32_t *)(A.col[x.col] + (size_t)w * A.stride * x.bytes), (const uint32_t *)x.buffer + off * (x.bytes >> 2), total, lane);
        }
        __syncwarp();
        // 2. TaskGraph::run, `steps` times (a batch: the staged exports are observable only at its end, so
        //    copy-in happens once before and offsets + copy-out once after; nothing else crosses worlds)
        loadWorld(ctx, E, &shared[warp], w);
        const bool aligned = align && (W % kWarpsPerBlock) == 0;       // no idle warp in any block
        for (int k = 0; k < steps; k++) {
            if (k) beginStep(ctx);
            runNodes(ctx, nodes, last - first, AppsBuiltin(), aligned);
        }
        storeWorld(ctx);
    }
    // 3. every world's row counts are final
    gridBarrier(bar);
    // 4 + 5. offsets of this block's worlds (prefix over the worlds before them), then copy-out
    const int w0 = blockIdx.x * kWarpsPerBlock;
    for (int s = 0; s < E->num_exports; s++) {
        const ExportDev &x = E->exports[s];
        if (x.fixed) continue;
        const ArchDev &A = E->arch[x.arch];
        const int32_t *counts = A.row_count;
        long long part = 0;
        for (int i = threadIdx.x; i < w0 && i < W; i += blockDim.x) part += __ldcg(counts + i);
        for (int o = 16; o; o >>= 1) part += __shfl_xor_sync(kFull, part, o);
        if (lane == 0) s_warp_sum[warp] = part;
        __syncthreads();
        if (threadIdx.x == 0) {
            long long t = 0;
            for (int k = 0; k < kWarpsPerBlock; k++) t += s_warp_sum[k];
            s_base = t;
        }
        __syncthreads();
        long long off = s_base;
        for (int k = 0; k < warp; k++) if (w0 + k < W) off += __ldcg(counts + w0 + k);
        const int32_t n = active ? __ldcg(counts + w) : 0;
        if (active) {
            if (lane == 0) {
                x.offsets[w] = off;
                if (w == W - 1) x.offsets[W] = off + n;
            }
            const int words = x.bytes >> 2;
            copyWordsWarp((uint32_t *)x.buffer + off * words, (const uint32_t *)(A.col[x.col] + (size_t)w * A.stride * x.bytes), n * words, lane);
        }
        __syncthreads();
    }
}

// Device-side column lookups through the uploaded maps (what Context::get<T>(Loc) does per call).
__global__ void ecsColumnLookupKernel(const EcsDev *__restrict__ E, int arch, const int32_t *comps, int n, int32_t *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = comps[i] < 0 ? -1 : E->col_maps[arch].lookup((uint32_t)comps[i]);
}

// Replicates the id nodes every world starts with (singleton entities).
__global__ void ecsBroadcastNodes(IdNode *dst, int32_t id_cap, const IdNode *src, int32_t n, int32_t num_worlds)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)num_worlds * n) return;
    const int32_t w = (int32_t)(i / n), k = (int32_t)(i - (int64_t)w * n);
    dst[(size_t)w * id_cap + k] = src[k];
}

// ---------------------------------------------------------------- host side --
struct HostComponent { int32_t bytes, align; };
struct HostArchetype { std::vector<int32_t> comps; int32_t max_rows, flags; bool singleton; };
struct HostNode { int32_t kind, func, arch, key_comp, remap; std::vector<int32_t> comps, deps; void *fn = nullptr; };
struct HostExport { int32_t arch, comp; };

}  // namespace

struct mdr_ecs {
    mdr_ecs_config cfg;
    std::vector<HostComponent> comps;
    std::vector<HostArchetype> archs;
    std::vector<HostNode> nodes;
    std::vector<HostExport> exports;    // by slot
    bool built = false;
    EcsDev host {};                     // host copy of the device descriptor
    std::vector<ColumnMapDev> col_maps; // per archetype, built at mdr_ecs_build
    EcsDev *dev = nullptr;
    std::vector<void *> allocs;
    cudaStream_t stream = nullptr;
    cudaGraphExec_t graph = nullptr;
    bool exports_dirty = true;
    int blocks = 0;
    void (*step_kernel)(const EcsDev *, int, int, int) = nullptr;
    void (*fused_kernel)(const EcsDev *, int, int, unsigned *, int, int) = nullptr;
    int align_nodes = -1;           // the warps of a block enter each node together (fused kernel): -1 = by policy,
                                    // MDR_ECS_ALIGN_NODES=0 / 1 force it   // the whole step in one launch, when it applies
    bool all_resident = false;      // every world fits the fused kernel's resident grid (and MDR_ECS_NO_FUSED is not set)
    bool batch_ok = true;           // n > 1 steps of one call run inside one launch (MDR_ECS_NO_BATCH=1: n graph launches)
    int last_call_launches = 0;     // kernels launched by the last mdr_ecs_step_async call
    volatile int32_t *h_error = nullptr;   // host view of EcsDev::error (mapped page-locked memory)
    int64_t *h_totals = nullptr;    // page-locked: row totals of the staged exports, read back by mdr_ecs_step_read
    unsigned *grid_bar = nullptr;
    bool fused_ok = false, fused_in_graph = false;
    mdr_ecs_launch_fn launcher = nullptr;   // megakernel instantiated in the user's translation unit
    void *launcher_user = nullptr;
    int32_t tmp_bytes = 0, world_data_bytes = 0;
    std::vector<mdr_phys *> physics;        // by node index (kNodePhysics), attached after build
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double timed_ms = 0;
    int timed_steps = 0, pending_steps = 0;
    bool timing_open = false;

    bool alloc_failed = false;      // a table did not fit the device: mdr_ecs_build frees everything, MDR_ECAPACITY
    template <typename T> T *dalloc(size_t n, bool zero = true)
    {
        void *p = nullptr;
        const size_t bytes = (n ? n : 1) * sizeof(T);
        if (alloc_failed) return nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess) {
            cudaGetLastError();     // out of memory is not sticky
            alloc_failed = true;
            return nullptr;
        }
        if (zero) REQ_CUDA(cudaMemset(p, 0, bytes));
        allocs.push_back(p);
        return (T *)p;
    }
};

namespace {

int colOf(const HostArchetype &a, int32_t comp)
{
    if (comp == 0) return 0;   // Entity is column 0 of every table (src/core/state.cpp:383-385)
    for (size_t i = 0; i < a.comps.size(); i++)
        if (a.comps[i] == comp) return (int)i + 1;
    return -1;
}

// The archetype's column map: Entity (component 0) -> column 0, user components -> 1.. in registration order
// (the lookupInputs of StateManager::registerArchetype, src/core/state.cpp:383-404).
bool buildArchColumnMap(ColumnMapDev &m, const HostArchetype &a)
{
    std::vector<uint32_t> keys { 0u }, vals { 0u };
    for (size_t i = 0; i < a.comps.size(); i++) { keys.push_back((uint32_t)a.comps[i]); vals.push_back((uint32_t)i + 1); }
    memset(&m, 0, sizeof(m));
    return buildColumnMap(m, keys.data(), vals.data(), (uint32_t)keys.size());
}

bool depsOk(const mdr_ecs *e, const int32_t *deps, int32_t ndeps)
{
    // Builder::build (src/core/taskgraph.cpp:39-109) emits nodes in insertion order
    // and spins if a dependency is not yet emitted: only earlier nodes are valid.
    if (ndeps < 0 || (ndeps > 0 && !deps)) return false;
    for (int i = 0; i < ndeps; i++)
        if (deps[i] < 0 || deps[i] >= (int32_t)e->nodes.size()) return false;
    return true;
}

void closeTiming(mdr_ecs *e)
{
    if (!e->timing_open) return;
    float ms = 0.f;
    REQ_CUDA(cudaEventSynchronize(e->ev1));
    REQ_CUDA(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    e->timed_ms += ms;
    e->timed_steps += e->pending_steps;
    e->pending_steps = 0;
    e->timing_open = false;
}

void refreshExports(mdr_ecs *e)
{
    if (e->host.num_exports > 0) {
        launchExportOffsets(e->dev, e->host.num_exports, e->host.num_worlds, e->stream);
        ecsExportCopy<<<(e->host.num_worlds + 7) / 8, 256, 0, e->stream>>>(e->dev, 1);
        REQ_CUDA(cudaGetLastError());
    }
    e->exports_dirty = false;
}

void runScript(mdr_ecs *e, const mdr_script_op *d_ops, const int32_t *d_offsets, int32_t num_shared)
{
    ecsScriptKernel<<<e->blocks, 32 * kWarpsPerBlock, 0, e->stream>>>(e->dev, d_ops, d_offsets, num_shared);
    REQ_CUDA(cudaGetLastError());
    e->exports_dirty = true;
}

int32_t checkError(mdr_ecs *e)
{
    REQ_CUDA(cudaStreamSynchronize(e->stream));      // callers have waited already: returns at once
    const int32_t err = e->h_error ? *e->h_error : 0;
    if (err) {
        *e->h_error = 0;
        return MDR_EOVERFLOW;
    }
    return MDR_OK;
}

}  // namespace

extern "C" {

int32_t mdr_ecs_create(const mdr_ecs_config *cfg, mdr_ecs **out)
{
    if (!cfg || !out || cfg->num_worlds <= 0 || cfg->first_world < 0) return MDR_EINVAL;
    if (cfg->gpu_id < 0 || cfg->gpu_id >= mdr_device_count()) return MDR_ENODEVICE;
    REQ_CUDA(cudaSetDevice(cfg->gpu_id));
    cudaDeviceProp prop;
    REQ_CUDA(cudaGetDeviceProperties(&prop, cfg->gpu_id));
    if (prop.major != 10) {
        fprintf(stderr, "madrona_b200: device %d is sm_%d%d; this library is built for sm_100a only\n",
                cfg->gpu_id, prop.major, prop.minor);
        return MDR_ENODEVICE;
    }
    mdr_ecs *e = new mdr_ecs;
    e->cfg = *cfg;
    e->comps.push_back({ 8, 4 });   // Entity is component 0 (StateManager ctor, src/core/state.cpp:148-162)
    REQ_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    REQ_CUDA(cudaEventCreate(&e->ev0));
    REQ_CUDA(cudaEventCreate(&e->ev1));
    *out = e;
    return MDR_OK;
}

void mdr_ecs_destroy(mdr_ecs *e)
{
    if (!e) return;
    cudaSetDevice(e->cfg.gpu_id);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->graph) cudaGraphExecDestroy(e->graph);
    if (e->stream) cudaStreamDestroy(e->stream);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    for (void *a : e->allocs) cudaFree(a);
    if (e->h_totals) cudaFreeHost(e->h_totals);
    if (e->h_error) cudaFreeHost((void *)e->h_error);
    delete e;
}

int32_t mdr_ecs_register_component(mdr_ecs *e, int32_t bytes, int32_t align)
{
    if (!e || bytes <= 0 || bytes % 4 != 0 || align <= 0 || align > 16) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    e->comps.push_back({ bytes, align });
    return (int32_t)e->comps.size() - 1;
}

static int32_t addArchetype(mdr_ecs *e, const int32_t *comps, int32_t n, int32_t max_rows, int32_t flags, bool singleton)
{
    if (!e || n <= 0 || n + 1 > kMaxCols || !comps || max_rows <= 0) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    if ((int)e->archs.size() >= kMaxArch) return MDR_ECAPACITY;
    HostArchetype a;
    for (int i = 0; i < n; i++) {
        if (comps[i] <= 0 || comps[i] >= (int32_t)e->comps.size()) return MDR_EINVAL;
        a.comps.push_back(comps[i]);
    }
    a.max_rows = max_rows; a.flags = flags; a.singleton = singleton;
    e->archs.push_back(a);
    return (int32_t)e->archs.size() - 1;
}

int32_t mdr_ecs_register_archetype(mdr_ecs *e, const int32_t *comps, int32_t n, int32_t max_rows, int32_t flags)
{
    return addArchetype(e, comps, n, max_rows, flags, false);
}

int32_t mdr_ecs_register_singleton(mdr_ecs *e, int32_t bytes, int32_t align)
{
    int32_t c = mdr_ecs_register_component(e, bytes, align);
    if (c < 0) return c;
    return addArchetype(e, &c, 1, 1, MDR_ARCH_FIXED, true);
}

int32_t mdr_ecs_export_column(mdr_ecs *e, int32_t arch, int32_t comp, int32_t slot)
{
    if (!e || arch < 0 || arch >= (int32_t)e->archs.size() || slot < 0 || slot >= kMaxExports) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    if (colOf(e->archs[arch], comp) < 0) return MDR_EINVAL;
    if ((int)e->exports.size() <= slot) e->exports.resize(slot + 1, HostExport { -1, -1 });
    e->exports[slot] = { arch, comp };
    return MDR_OK;
}

static int32_t addNode(mdr_ecs *e, HostNode n, const int32_t *deps, int32_t ndeps)
{
    if (!e) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    if (!depsOk(e, deps, ndeps)) return MDR_EINVAL;
    if ((int)e->nodes.size() >= kMaxNodes) return MDR_ECAPACITY;
    n.deps.assign(deps, deps + ndeps);
    e->nodes.push_back(n);
    return (int32_t)e->nodes.size() - 1;
}

int32_t mdr_ecs_add_parallel_for(mdr_ecs *e, int32_t func, const int32_t *comps, int32_t n, const int32_t *deps,
                                 int32_t ndeps)
{
    if (!e || !comps || n <= 0 || n > kMaxQueryComps || func <= 0) return MDR_EINVAL;
    HostNode node { kNodeParallelFor, func, -1, -1, 0, {}, {} };
    for (int i = 0; i < n; i++) {
        if (comps[i] < 0 || comps[i] >= (int32_t)e->comps.size()) return MDR_EINVAL;
        node.comps.push_back(comps[i]);
    }
    return addNode(e, node, deps, ndeps);
}

int32_t mdr_ecs_add_parallel_for_fn(mdr_ecs *e, const void *device_fn, const int32_t *comps, int32_t n,
                                    const int32_t *deps, int32_t ndeps)
{
    if (!e || !comps || n <= 0 || n > kMaxQueryComps || !device_fn) return MDR_EINVAL;
    HostNode node { kNodeParallelFor, 0, -1, -1, 0, {}, {} };
    node.fn = (void *)device_fn;
    for (int i = 0; i < n; i++) {
        if (comps[i] < 0 || comps[i] >= (int32_t)e->comps.size()) return MDR_EINVAL;
        node.comps.push_back(comps[i]);
    }
    return addNode(e, node, deps, ndeps);
}

int32_t mdr_ecs_add_physics(mdr_ecs *e, const int32_t *deps, int32_t ndeps)
{
    return addNode(e, HostNode { kNodePhysics, 0, -1, -1, 0, {}, {} }, deps, ndeps);
}

int32_t mdr_ecs_attach_physics(mdr_ecs *e, int32_t node, mdr_phys *p)
{
    if (!e || !p) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (node < 0 || node >= (int32_t)e->nodes.size() || e->nodes[node].kind != kNodePhysics) return MDR_EINVAL;
    e->physics.resize(e->nodes.size(), nullptr);
    e->physics[node] = p;
    if (e->graph) { cudaGraphExecDestroy(e->graph); e->graph = nullptr; }
    return MDR_OK;
}

// RigidBodyPhysicsSystem on an archetype of this ECS: the archetype must start with the twelve physics
// components in phys::Cols order after Entity (include/madrona/physics.hpp:449-464), byte sizes checked.
int32_t mdr_phys_create_on_ecs(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                               const float *verts, const uint32_t *poly_words, mdr_ecs *e, int32_t arch,
                               mdr_phys **out)
{
    if (!cfg || !e || !out) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (arch < 0 || arch >= e->host.num_arch) return MDR_EINVAL;
    const ArchDev &A = e->host.arch[arch];
    if (A.num_cols < MDR_NUM_PHYS_COLS || cfg->num_worlds != e->host.num_worlds || cfg->gpu_id != e->cfg.gpu_id ||
        cfg->max_bodies_per_world != A.max_rows)
        return MDR_EINVAL;
    void *cols[MDR_NUM_PHYS_COLS];
    for (int c = 0; c < MDR_NUM_PHYS_COLS; c++) {
        if (A.col_bytes[c] != mdr_phys_column_bytes(c)) return MDR_EINVAL;
        cols[c] = A.col[c];
    }
    return mdr_phys_create_on_columns(cfg, objs, num_objs, verts, poly_words, A.stride, cols, A.row_count, out);
}

int32_t mdr_ecs_set_step_launcher(mdr_ecs *e, mdr_ecs_launch_fn fn, void *user)
{
    if (!e) return MDR_EINVAL;
    e->launcher = fn;
    e->launcher_user = user;
    if (e->graph) { cudaGraphExecDestroy(e->graph); e->graph = nullptr; }
    return MDR_OK;
}

int32_t mdr_ecs_set_tmp_bytes(mdr_ecs *e, int32_t bytes_per_world)
{
    if (!e || bytes_per_world < 0) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    e->tmp_bytes = (bytes_per_world + 255) / 256 * 256;
    return MDR_OK;
}

int32_t mdr_ecs_set_world_data_bytes(mdr_ecs *e, int32_t bytes_per_world)
{
    if (!e || bytes_per_world < 0) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    e->world_data_bytes = (bytes_per_world + 15) / 16 * 16;
    return MDR_OK;
}

const void *mdr_ecs_device_descriptor(mdr_ecs *e) { return e && e->built ? e->dev : nullptr; }
int32_t mdr_ecs_num_worlds(mdr_ecs *e) { return e ? e->cfg.num_worlds : MDR_EINVAL; }

int32_t mdr_ecs_tables_written(mdr_ecs *e)
{
    if (!e) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    e->exports_dirty = true;
    return MDR_OK;
}

int32_t mdr_ecs_add_clear_tmp(mdr_ecs *e, int32_t arch, const int32_t *deps, int32_t ndeps)
{
    if (!e || arch < 0 || arch >= (int32_t)e->archs.size()) return MDR_EINVAL;
    return addNode(e, HostNode { kNodeClearTmp, 0, arch, -1, 0, {}, {} }, deps, ndeps);
}

int32_t mdr_ecs_add_reset_tmp_alloc(mdr_ecs *e, const int32_t *deps, int32_t ndeps)
{
    return addNode(e, HostNode { kNodeResetTmpAlloc, 0, -1, -1, 0, {}, {} }, deps, ndeps);
}

int32_t mdr_ecs_add_sort(mdr_ecs *e, int32_t arch, int32_t key_comp, int32_t remap_entities, const int32_t *deps,
                         int32_t ndeps)
{
    if (!e || arch < 0 || arch >= (int32_t)e->archs.size()) return MDR_EINVAL;
    const int col = colOf(e->archs[arch], key_comp);
    if (col <= 0 || e->comps[key_comp].bytes != 4) return MDR_EINVAL;
    return addNode(e, HostNode { kNodeSort, 0, arch, key_comp, remap_entities, {}, {} }, deps, ndeps);
}

static int32_t buildImpl(mdr_ecs *e);

// A build that fails (capacity of the descriptor, or tables that do not fit the device) leaves nothing behind:
// every allocation is released and the registry can be built again (with fewer worlds, say).
int32_t mdr_ecs_build(mdr_ecs *e)
{
    if (!e) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    if (e->archs.empty()) return MDR_EINVAL;
    int32_t rc = buildImpl(e);
    if (rc == MDR_OK && e->alloc_failed) rc = MDR_ECAPACITY;
    if (rc != MDR_OK) {
        if (e->alloc_failed)
            fprintf(stderr, "madrona_b200: %d worlds of this registry do not fit the device memory\n", e->cfg.num_worlds);
        for (void *a : e->allocs) cudaFree(a);
        e->allocs.clear();
        if (e->h_error) { cudaFreeHost((void *)e->h_error); e->h_error = nullptr; }
        e->alloc_failed = false;
        e->built = false;
        e->dev = nullptr;
        memset(&e->host, 0, sizeof(e->host));
    }
    return rc;
}

static int32_t buildImpl(mdr_ecs *e)
{
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    EcsDev &H = e->host;
    memset(&H, 0, sizeof(H));
    const int W = e->cfg.num_worlds;
    H.num_worlds = W;
    H.first_world = e->cfg.first_world;
    H.num_arch = (int32_t)e->archs.size();
    H.app = e->cfg.app;
    int64_t total_rows = 0;
    int num_singletons = 0;
    e->col_maps.resize(H.num_arch);
    for (int a = 0; a < H.num_arch; a++)
        if (!buildArchColumnMap(e->col_maps[a], e->archs[a])) return MDR_ECAPACITY;   // the reference FATALs
    H.col_maps = e->dalloc<ColumnMapDev>((size_t)H.num_arch, false);
    if (e->alloc_failed) return MDR_ECAPACITY;
    REQ_CUDA(cudaMemcpy(H.col_maps, e->col_maps.data(), sizeof(ColumnMapDev) * (size_t)H.num_arch, cudaMemcpyHostToDevice));
    for (int a = 0; a < H.num_arch; a++) {
        const HostArchetype &ha = e->archs[a];
        ArchDev &A = H.arch[a];
        A.num_cols = (int32_t)ha.comps.size() + 1;
        A.max_rows = ha.max_rows;
        A.stride = (ha.max_rows + 15) / 16 * 16;
        A.flags = ha.flags;
        A.row_count = e->dalloc<int32_t>(W);
        for (int c = 0; c < A.num_cols; c++) {
            A.col_bytes[c] = c == 0 ? 8 : e->comps[ha.comps[c - 1]].bytes;
            A.col[c] = e->dalloc<char>((size_t)W * A.stride * A.col_bytes[c]);
        }
        total_rows += ha.max_rows;
        num_singletons += ha.singleton;
    }
    H.id_cap = e->cfg.ids_per_world > 0 ? (e->cfg.ids_per_world + 63) / 64 * 64
                                        : (int32_t)((total_rows + 63) / 64 * 64 + 64 * (4 + (num_singletons ? 1 : 0)));
    H.id_nodes = e->dalloc<IdNode>((size_t)W * H.id_cap);
    H.id_state = e->dalloc<IdState>(W);
    {   // the error word lives in page-locked host memory mapped into the device: checkError reads it after a
        // stream wait without a blocking 4-byte copy per step
        void *h = nullptr, *d = nullptr;
        if (cudaHostAlloc(&h, 64, cudaHostAllocMapped) != cudaSuccess || cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
            cudaGetLastError();
            if (h) cudaFreeHost(h);
            e->alloc_failed = true;
        } else {
            memset(h, 0, 64);
            e->h_error = (volatile int32_t *)h;
        }
        H.error = (int32_t *)d;
    }
    H.tmp_bytes = e->tmp_bytes;
    if (H.tmp_bytes > 0) {
        H.tmp_arena = e->dalloc<char>((size_t)W * H.tmp_bytes, false);
        H.tmp_used = e->dalloc<int32_t>(W);
    }
    H.world_data_bytes = e->world_data_bytes;
    if (H.world_data_bytes > 0) H.world_data = e->dalloc<char>((size_t)W * H.world_data_bytes);
    H.max_handles = e->cfg.max_script_handles;
    if (H.max_handles > 0) {
        H.handles = e->dalloc<Entity>((size_t)W * H.max_handles);
        H.num_handles = e->dalloc<int32_t>(W);
    }

    if (e->alloc_failed) return MDR_ECAPACITY;

    // Initial id state of every world. registerSingleton (include/madrona/state.inl:
    // 171-187) makes one entity per singleton through the StateManager's own init cache.
    std::vector<IdNode> nodes0(H.id_cap, IdNode { 0, 0, 0u });
    IdState st0;
    st0.init_cache = IdCache { kSentinel, 0, kSentinel, 0 };
    st0.cache = st0.init_cache;
    st0.global_head = kSentinel;
    st0.num_ids = 0;
    int32_t herr = 0;
    for (int a = 0; a < H.num_arch; a++) {
        if (!e->archs[a].singleton) continue;
        Entity ent = idAcquire(nodes0.data(), st0, st0.init_cache, H.id_cap, &herr);
        if (ent.id < 0) return MDR_ECAPACITY;
        nodes0[ent.id].a = a;
        nodes0[ent.id].b = 0;
        std::vector<Entity> col((size_t)W * H.arch[a].stride, Entity { 0u, 0 });
        std::vector<int32_t> ones(W, 1);
        for (int w = 0; w < W; w++) col[(size_t)w * H.arch[a].stride] = ent;
        REQ_CUDA(cudaMemcpy(H.arch[a].col[0], col.data(), col.size() * 8, cudaMemcpyHostToDevice));
        REQ_CUDA(cudaMemcpy(H.arch[a].row_count, ones.data(), 4 * (size_t)W, cudaMemcpyHostToDevice));
    }
    if (st0.num_ids > 0) {
        IdNode *d0 = e->dalloc<IdNode>(st0.num_ids, false);
        if (e->alloc_failed) return MDR_ECAPACITY;
        REQ_CUDA(cudaMemcpy(d0, nodes0.data(), sizeof(IdNode) * (size_t)st0.num_ids, cudaMemcpyHostToDevice));
        const int64_t total = (int64_t)W * st0.num_ids;
        ecsBroadcastNodes<<<(unsigned)((total + 255) / 256), 256>>>(H.id_nodes, H.id_cap, d0, st0.num_ids, W);
        REQ_CUDA(cudaGetLastError());
        REQ_CUDA(cudaDeviceSynchronize());
    }
    {
        std::vector<IdState> states(W, st0);
        REQ_CUDA(cudaMemcpy(H.id_state, states.data(), sizeof(IdState) * (size_t)W, cudaMemcpyHostToDevice));
    }

    // Lower the graph: insertion order is the sorted order (see depsOk); resolve every
    // query the way StateManager::makeQuery does (archetypes in id order).
    H.num_nodes = (int32_t)e->nodes.size();
    for (int ni = 0; ni < H.num_nodes; ni++) {
        const HostNode &hn = e->nodes[ni];
        NodeDev &N = H.node[ni];
        N.kind = hn.kind; N.func = hn.func; N.arch = hn.arch; N.remap = hn.remap;
        N.fn = (NodeFn)hn.fn;
        if (hn.kind == kNodeSort) {
            N.key_col = colOf(e->archs[hn.arch], hn.key_comp);
            SortScratch &S = H.sort[hn.arch];
            if (!S.key[0]) {
                const ArchDev &A = H.arch[hn.arch];
                const size_t R = (size_t)W * A.stride;
                int mb = 0;
                for (int c = 0; c < A.num_cols; c++) mb = A.col_bytes[c] > mb ? A.col_bytes[c] : mb;
                S.staging_bytes = (mb + 15) / 16 * 16;
                S.key[0] = e->dalloc<uint32_t>(R, false); S.key[1] = e->dalloc<uint32_t>(R, false);
                S.idx[0] = e->dalloc<int32_t>(R, false); S.idx[1] = e->dalloc<int32_t>(R, false);
                S.staging = e->dalloc<char>(R * S.staging_bytes, false);
            }
        }
        if (hn.kind != kNodeParallelFor) continue;
        for (int a = 0; a < H.num_arch; a++) {
            // StateManager::makeQuery (src/core/state.cpp:305-356): columnLookup.exists, then columnLookup[]
            const ColumnMapDev &cm = e->col_maps[a];
            bool has = true;
            for (int32_t c : hn.comps) has = has && cm.exists((uint32_t)c);
            if (!has) continue;
            if (N.num_match >= kMaxMatch) return MDR_ECAPACITY;
            MatchDev &M = N.match[N.num_match++];
            M.arch = a;
            for (size_t k = 0; k < hn.comps.size(); k++) M.col[k] = (int32_t)cm.at((uint32_t)hn.comps[k]);
        }
    }
    H.num_exports = (int32_t)e->exports.size();
    for (int s = 0; s < H.num_exports; s++) {
        const HostExport &hx = e->exports[s];
        if (hx.arch < 0) return MDR_EINVAL;
        ExportDev &X = H.exports[s];
        const ArchDev &A = H.arch[hx.arch];
        X.arch = hx.arch; X.col = colOf(e->archs[hx.arch], hx.comp); X.bytes = A.col_bytes[X.col];
        X.fixed = (A.flags & MDR_ARCH_FIXED) ? 1 : 0;
        if (X.fixed) {
            X.buffer = A.col[X.col];   // fixed-size archetypes export in place (src/core/state.cpp:481-483)
        } else {
            X.buffer = e->dalloc<char>((size_t)W * A.stride * X.bytes);
            X.offsets = e->dalloc<int64_t>((size_t)W + 1);
        }
    }
    e->dev = e->dalloc<EcsDev>(1, false);
    if (e->alloc_failed) return MDR_ECAPACITY;
    REQ_CUDA(cudaMemcpy(e->dev, &H, sizeof(H), cudaMemcpyHostToDevice));

    // Register budget of the megakernel: more resident warps hide the latency of the
    // dependent table accesses (the step is latency-bound, profiles/r01e_summary.md).
    // Measured (profiles/r01e_summary.md): 4 blocks per SM (62 registers, no spills) is the
    // fastest per warp; 6 or 8 (40 / 32 registers, spills) only pay when they let every world
    // be resident at once instead of leaving a partial second round.
    cudaDeviceProp prop0;
    REQ_CUDA(cudaGetDeviceProperties(&prop0, e->cfg.gpu_id));
    const char *mb = getenv("MDR_ECS_MIN_BLOCKS");
    int want = mb && *mb ? atoi(mb) : 0;
    if (want == 0) {
        auto rounds = [&](int v) {
            const long long cap = (long long)prop0.multiProcessorCount * v * kWarpsPerBlock;
            return (W + cap - 1) / cap;
        };
        want = 4;
        if (rounds(6) < rounds(want)) want = 6;
        if (rounds(8) < rounds(want)) want = 8;
    }
    e->step_kernel = want == 6 ? ecsStepKernel<6> : want == 8 ? ecsStepKernel<8> : ecsStepKernel<4>;
    e->fused_kernel = want == 6 ? ecsFusedStepKernel<6> : want == 8 ? ecsFusedStepKernel<8> : ecsFusedStepKernel<4>;
    e->grid_bar = e->dalloc<unsigned>(2);
    if (e->alloc_failed) return MDR_ECAPACITY;
    int per_sm = 0;
    REQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, e->step_kernel, 32 * kWarpsPerBlock, 0));
    cudaDeviceProp prop;
    REQ_CUDA(cudaGetDeviceProperties(&prop, e->cfg.gpu_id));
    long long grid = (long long)(per_sm > 0 ? per_sm : 1) * prop.multiProcessorCount;
    const long long need = ((long long)W + kWarpsPerBlock - 1) / kWarpsPerBlock;
    e->blocks = (int)(grid < need ? grid : need);
    // the one-launch step needs every world resident at once under the fused kernel's own occupancy
    int fused_per_sm = 0;
    REQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fused_per_sm, e->fused_kernel, 32 * kWarpsPerBlock, 0));
    // Measured (profiles/r02_summary.md section 6): level with the four-launch graph from 4096 worlds on, 7 % slower
    // at 1024 (the barrier is two more dependent round trips in a step that is all latency): on by default from 2048
    // worlds, MDR_ECS_FUSED=1 / MDR_ECS_NO_FUSED=1 force it either way.
    const char *nf = getenv("MDR_ECS_NO_FUSED"), *ff = getenv("MDR_ECS_FUSED");
    e->all_resident = need <= (long long)fused_per_sm * prop.multiProcessorCount && !(nf && *nf == '1');
    e->fused_ok = !(nf && *nf == '1') && (W >= 2048 || (ff && *ff == '1')) && e->all_resident;
    const char *an = getenv("MDR_ECS_ALIGN_NODES");
    e->align_nodes = an && *an ? (*an == '1') : -1;
    const char *nb = getenv("MDR_ECS_NO_BATCH");
    e->batch_ok = !(nb && *nb == '1');
    e->built = true;
    e->exports_dirty = true;
    return MDR_OK;
}

int32_t mdr_ecs_make_entities(mdr_ecs *e, int32_t arch, int32_t count)
{
    if (!e || !e->built) return e ? MDR_ESTATE : MDR_EINVAL;
    if (arch < 0 || arch >= e->host.num_arch || count < 0) return MDR_EINVAL;
    if (count == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    std::vector<mdr_script_op> ops(count, mdr_script_op { MDR_OP_CREATE, arch, 0 });
    mdr_script_op *d = nullptr;
    REQ_CUDA(cudaMalloc(&d, sizeof(mdr_script_op) * ops.size()));
    REQ_CUDA(cudaMemcpyAsync(d, ops.data(), sizeof(mdr_script_op) * ops.size(), cudaMemcpyHostToDevice, e->stream));
    runScript(e, d, nullptr, count);
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    REQ_CUDA(cudaFree(d));
    return checkError(e);
}

int32_t mdr_ecs_script(mdr_ecs *e, const mdr_script_op *ops, const int32_t *offsets)
{
    if (!e || !ops || !offsets) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    const int W = e->host.num_worlds;
    const int32_t total = offsets[W];
    if (total <= 0) return MDR_OK;
    for (int32_t i = 0; i < total; i++) {
        // MDR_OP_TMP / MDR_OP_CLEAR_TMP address the script app's temporary archetype (3)
        if ((ops[i].op == MDR_OP_TMP || ops[i].op == MDR_OP_CLEAR_TMP) && (e->cfg.app != MDR_APP_SCRIPT || e->host.num_arch <= 3))
            return MDR_ESTATE;
        if (ops[i].op < MDR_OP_CREATE || ops[i].op > MDR_OP_CLEAR_TMP) return MDR_EINVAL;
    }
    mdr_script_op *d = nullptr;
    int32_t *doff = nullptr;
    REQ_CUDA(cudaMalloc(&d, sizeof(mdr_script_op) * (size_t)total));
    REQ_CUDA(cudaMalloc(&doff, 4 * ((size_t)W + 1)));
    REQ_CUDA(cudaMemcpyAsync(d, ops, sizeof(mdr_script_op) * (size_t)total, cudaMemcpyHostToDevice, e->stream));
    REQ_CUDA(cudaMemcpyAsync(doff, offsets, 4 * ((size_t)W + 1), cudaMemcpyHostToDevice, e->stream));
    runScript(e, d, doff, 0);
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    REQ_CUDA(cudaFree(d));
    REQ_CUDA(cudaFree(doff));
    return checkError(e);
}

static int32_t columnCopy(mdr_ecs *e, int32_t arch, int32_t col, int32_t rows, void *host, bool to_device)
{
    if (!e || !host) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (arch < 0 || arch >= e->host.num_arch) return MDR_EINVAL;
    const ArchDev &A = e->host.arch[arch];
    if (col < 0 || col >= A.num_cols || rows < 0 || rows > A.stride) return MDR_EINVAL;
    if (rows == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    const size_t nb = A.col_bytes[col];
    if (to_device) {
        REQ_CUDA(cudaMemcpy2DAsync(A.col[col], A.stride * nb, host, rows * nb, rows * nb, e->host.num_worlds,
                                   cudaMemcpyHostToDevice, e->stream));
        e->exports_dirty = true;
    } else {
        REQ_CUDA(cudaMemcpy2DAsync(host, rows * nb, A.col[col], A.stride * nb, rows * nb, e->host.num_worlds,
                                   cudaMemcpyDeviceToHost, e->stream));
    }
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    return MDR_OK;
}

int32_t mdr_ecs_write_column(mdr_ecs *e, int32_t arch, int32_t col, int32_t rows, const void *host)
{
    return columnCopy(e, arch, col, rows, (void *)host, true);
}

int32_t mdr_ecs_read_column(mdr_ecs *e, int32_t arch, int32_t col, int32_t rows, void *host)
{
    return columnCopy(e, arch, col, rows, host, false);
}

int32_t mdr_ecs_row_counts(mdr_ecs *e, int32_t arch, int32_t *host)
{
    if (!e || !host) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (arch < 0 || arch >= e->host.num_arch) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    REQ_CUDA(cudaMemcpy(host, e->host.arch[arch].row_count, 4 * (size_t)e->host.num_worlds, cudaMemcpyDeviceToHost));
    return MDR_OK;
}

int32_t mdr_ecs_num_components(mdr_ecs *e) { return e ? (int32_t)e->comps.size() : MDR_EINVAL; }
int32_t mdr_ecs_num_archetypes(mdr_ecs *e) { return e ? (int32_t)e->archs.size() : MDR_EINVAL; }
int32_t mdr_ecs_num_columns(mdr_ecs *e, int32_t arch)
{
    return e && arch >= 0 && arch < (int32_t)e->archs.size() ? (int32_t)e->archs[arch].comps.size() + 1 : MDR_EINVAL;
}
int32_t mdr_ecs_column_bytes(mdr_ecs *e, int32_t arch, int32_t col)
{
    if (!e || arch < 0 || arch >= (int32_t)e->archs.size()) return MDR_EINVAL;
    if (col == 0) return 8;
    if (col < 0 || col > (int32_t)e->archs[arch].comps.size()) return MDR_EINVAL;
    return e->comps[e->archs[arch].comps[col - 1]].bytes;
}
int32_t mdr_ecs_row_stride(mdr_ecs *e, int32_t arch)
{
    return e && e->built && arch >= 0 && arch < e->host.num_arch ? e->host.arch[arch].stride : MDR_EINVAL;
}
void *mdr_ecs_column_ptr(mdr_ecs *e, int32_t arch, int32_t col)
{
    if (!e || !e->built || arch < 0 || arch >= e->host.num_arch || col < 0 || col >= e->host.arch[arch].num_cols)
        return nullptr;
    return e->host.arch[arch].col[col];
}

int32_t mdr_ecs_id_capacity(mdr_ecs *e) { return e && e->built ? e->host.id_cap : MDR_EINVAL; }

int32_t mdr_ecs_read_idmap(mdr_ecs *e, int32_t world, int32_t *a, int32_t *b, uint32_t *gen, int32_t state[10])
{
    if (!e || !a || !b || !gen || !state) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (world < 0 || world >= e->host.num_worlds) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    IdState st;
    REQ_CUDA(cudaMemcpy(&st, e->host.id_state + world, sizeof(st), cudaMemcpyDeviceToHost));
    std::vector<IdNode> nodes(st.num_ids > 0 ? st.num_ids : 1);
    if (st.num_ids > 0)
        REQ_CUDA(cudaMemcpy(nodes.data(), e->host.id_nodes + (size_t)world * e->host.id_cap,
                            sizeof(IdNode) * (size_t)st.num_ids, cudaMemcpyDeviceToHost));
    for (int i = 0; i < st.num_ids; i++) { a[i] = nodes[i].a; b[i] = nodes[i].b; gen[i] = nodes[i].gen; }
    memcpy(state, &st, sizeof(st));
    return st.num_ids;
}

int32_t mdr_ecs_step_async(mdr_ecs *e, int32_t num_steps)
{
    if (!e || num_steps < 0) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    if (num_steps == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    // never block the host here: a timed batch that is still running is extended, not waited for
    bool extend = false;
    if (e->timing_open) {
        if (cudaEventQuery(e->ev1) == cudaSuccess) closeTiming(e);
        else { cudaGetLastError(); extend = true; }
    }
    if (e->exports_dirty) refreshExports(e);
    const bool staged = [&] {
        for (int s = 0; s < e->host.num_exports; s++)
            if (!e->host.exports[s].fixed) return true;
        return false;
    }();
    // A batch of n > 1 steps without a host-visible point between them: the n walks of a world's graph run inside
    // ONE launch (the staged exports are copied in before and out after; the tables, row counts and id state of a
    // world stay with its warp). Same results as n single steps, bit for bit (tests/test_ecs_gpu.py). Not for graphs
    // with a physics node (a launch boundary) or a user-instantiated megakernel (its launcher takes one step).
    bool has_physics_node = false;
    for (int ni = 0; ni < e->host.num_nodes; ni++) has_physics_node = has_physics_node || e->host.node[ni].kind == kNodePhysics;
    if (num_steps > 1 && e->batch_ok && !e->launcher && !has_physics_node) {
        if (!extend) REQ_CUDA(cudaEventRecord(e->ev0, e->stream));
        const int W = e->host.num_worlds, N = e->host.num_nodes;
        if (staged && e->all_resident) {
            // Measured (profiles/r02_summary.md section 11): with 64 warps per SM each at its own node, a long graph stalls on
            // instruction fetch (fantasy_vs, 13 nodes: `no_instruction` 13.5 cycles per issue); a block barrier per node
            // keeps the 8 warps of a block on the same code: +12 % for fantasy_vs batches, a loss for short graphs
            // (barriers with nothing to gain) and for single steps.
            const int align = e->align_nodes >= 0 ? e->align_nodes : (N >= 8);
            e->fused_kernel<<<(W + kWarpsPerBlock - 1) / kWarpsPerBlock, 32 * kWarpsPerBlock, 0, e->stream>>>(e->dev, 0, N, e->grid_bar, num_steps, align);
        } else {
            if (staged) ecsExportCopy<<<(W + 7) / 8, 256, 0, e->stream>>>(e->dev, 0);
            e->step_kernel<<<e->blocks, 32 * kWarpsPerBlock, 0, e->stream>>>(e->dev, 0, N, num_steps);
            if (staged) {
                launchExportOffsets(e->dev, e->host.num_exports, W, e->stream);
                ecsExportCopy<<<(W + 7) / 8, 256, 0, e->stream>>>(e->dev, 1);
            }
        }
        REQ_CUDA(cudaGetLastError());
        REQ_CUDA(cudaEventRecord(e->ev1, e->stream));
        e->timing_open = true;
        e->pending_steps = (extend ? e->pending_steps : 0) + num_steps;
        e->last_call_launches = staged && !e->all_resident ? 4 : 1;
        return MDR_OK;
    }
    if (!e->graph) {
        // MWCudaExecutor's run graph (src/mw/cuda_exec.cpp:1519-1690):
        // exportCopyIn -> megakernel -> exportBlockSums -> exportCopyOut
        cudaGraph_t g = nullptr;
        REQ_CUDA(cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal));
        bool has_physics = false;
        for (int ni = 0; ni < e->host.num_nodes; ni++) has_physics = has_physics || e->host.node[ni].kind == kNodePhysics;
        const bool fused = staged && e->fused_ok && !e->launcher && !has_physics;
        e->fused_in_graph = fused;
        if (fused) {
            const int blocks = (e->host.num_worlds + kWarpsPerBlock - 1) / kWarpsPerBlock;
            e->fused_kernel<<<blocks, 32 * kWarpsPerBlock, 0, e->stream>>>(e->dev, 0, e->host.num_nodes, e->grid_bar, 1, e->align_nodes > 0);
        }
        if (staged && !fused) ecsExportCopy<<<(e->host.num_worlds + 7) / 8, 256, 0, e->stream>>>(e->dev, 0);
        // The megakernel walks maximal runs of in-kernel nodes; a physics node is a launch of the
        // rigid-body step kernel on the same tables between two such runs (the reference's generated
        // megakernel takes the same (start_node, end_node) pair, src/mw/cuda_exec.cpp:728-745).
        int32_t rc_launch = MDR_OK;
        auto segment = [&](int first, int last) {
            if (first >= last) return;
            if (e->launcher) {
                const int32_t r = e->launcher(e->dev, first, last, e->host.num_worlds, (void *)e->stream, e->launcher_user);
                if (r != MDR_OK) rc_launch = r;
            } else {
                e->step_kernel<<<e->blocks, 32 * kWarpsPerBlock, 0, e->stream>>>(e->dev, first, last, 1);
            }
        };
        int first = fused ? e->host.num_nodes : 0;
        for (int ni = 0; ni < e->host.num_nodes && !fused; ni++) {
            if (e->host.node[ni].kind != kNodePhysics) continue;
            segment(first, ni);
            first = ni + 1;
            mdr_phys *ph = ni < (int)e->physics.size() ? e->physics[ni] : nullptr;
            if (!ph || mdrPhysEnqueueStep(ph, e->stream) != MDR_OK) rc_launch = MDR_ESTATE;
        }
        segment(first, e->host.num_nodes);
        if (staged && !fused) {
            launchExportOffsets(e->dev, e->host.num_exports, e->host.num_worlds, e->stream);
            ecsExportCopy<<<(e->host.num_worlds + 7) / 8, 256, 0, e->stream>>>(e->dev, 1);
        }
        REQ_CUDA(cudaStreamEndCapture(e->stream, &g));
        if (rc_launch != MDR_OK) { cudaGraphDestroy(g); return rc_launch; }
        REQ_CUDA(cudaGraphInstantiate(&e->graph, g, 0));
        REQ_CUDA(cudaGraphDestroy(g));
    }
    if (!extend) REQ_CUDA(cudaEventRecord(e->ev0, e->stream));
    for (int s = 0; s < num_steps; s++) REQ_CUDA(cudaGraphLaunch(e->graph, e->stream));
    REQ_CUDA(cudaEventRecord(e->ev1, e->stream));
    e->timing_open = true;
    e->pending_steps = (extend ? e->pending_steps : 0) + num_steps;
    e->last_call_launches = num_steps * mdr_ecs_launches_per_step(e);
    return MDR_OK;
}

int32_t mdr_ecs_last_launches(mdr_ecs *e) { return e ? e->last_call_launches : MDR_EINVAL; }

int32_t mdr_ecs_sync(mdr_ecs *e)
{
    if (!e) return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    closeTiming(e);
    return checkError(e);
}

int32_t mdr_ecs_step(mdr_ecs *e, int32_t num_steps)
{
    int32_t rc = mdr_ecs_step_async(e, num_steps);
    if (rc != MDR_OK) return rc;
    return mdr_ecs_sync(e);
}

void *mdr_ecs_stream(mdr_ecs *e) { return e ? (void *)e->stream : nullptr; }
void *mdr_ecs_done_event(mdr_ecs *e) { return e ? (void *)e->ev1 : nullptr; }

float mdr_ecs_step_ms(mdr_ecs *e, int32_t *steps)
{
    if (!e) return -1.f;
    closeTiming(e);
    const int n = e->timed_steps;
    const double ms = e->timed_ms;
    e->timed_steps = 0;
    e->timed_ms = 0;
    if (steps) *steps = n;
    return n > 0 ? (float)(ms / n) : -1.f;
}

#ifdef MDR_ECS_NODE_CLOCKS
// profiling builds only: reads and clears the per-node clock sums
int32_t mdr_ecs_node_clocks(mdr_ecs *e, unsigned long long *out, int32_t n)
{
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    REQ_CUDA(cudaMemcpyFromSymbol(out, g_node_clocks, sizeof(unsigned long long) * (size_t)n));
    static unsigned long long zeros[kMaxNodes + 2] = {};
    REQ_CUDA(cudaMemcpyToSymbol(g_node_clocks, zeros, sizeof(zeros)));
    return MDR_OK;
}
#endif

int32_t mdr_ecs_launches_per_step(mdr_ecs *e)
{
    if (!e || !e->built) return MDR_EINVAL;
    if (e->graph && e->fused_in_graph) return 1;
    int n = 0, run = 0;
    for (int ni = 0; ni < e->host.num_nodes; ni++) {
        if (e->host.node[ni].kind == kNodePhysics) { n += 2 + (run > 0); run = 0; }   // BVH rebuild check + step kernel
        else run++;
    }
    n += run > 0 || e->host.num_nodes == 0;
    for (int s = 0; s < e->host.num_exports; s++)
        if (!e->host.exports[s].fixed) return n + 3;
    return n;
}

int32_t mdr_static_map_build(const uint32_t *keys, const uint32_t *values, int32_t n,
                             uint32_t *key_storage, uint32_t *value_storage)
{
    if (!keys || !values || !key_storage || !value_storage || n < 0) return MDR_EINVAL;
    if ((uint32_t)n > kMapMaxKeys) return MDR_ECAPACITY;
    for (int i = 0; i < n; i++) if (keys[i] == kMapInvalidKey) return MDR_EINVAL;
    ColumnMapDev m;
    memset(&m, 0, sizeof(m));
    if (!buildColumnMap(m, keys, values, (uint32_t)n)) return MDR_ECAPACITY;
    memcpy(key_storage, m.keys, sizeof(m.keys));
    memcpy(value_storage, m.values, sizeof(m.values));
    return MDR_OK;
}

int32_t mdr_static_map_lookup(const uint32_t *key_storage, const uint32_t *value_storage, uint32_t key)
{
    if (!key_storage || !value_storage) return MDR_EINVAL;
    const uint32_t idx = staticMapIndex(key, key_storage[kMapShiftIdx], key_storage[kMapConstIdx]);
    if (idx >= kMapSlots || key_storage[idx] != key) return -1;
    return (int32_t)value_storage[idx];
}

int32_t mdr_ecs_column_map(mdr_ecs *e, int32_t arch, uint32_t *key_storage, uint32_t *value_storage)
{
    if (!e || !e->built || arch < 0 || arch >= (int)e->col_maps.size() || !key_storage || !value_storage) return MDR_EINVAL;
    memcpy(key_storage, e->col_maps[arch].keys, sizeof(uint32_t) * kMapSlots);
    memcpy(value_storage, e->col_maps[arch].values, sizeof(uint32_t) * kMapSlots);
    return MDR_OK;
}

int32_t mdr_ecs_lookup_columns(mdr_ecs *e, int32_t arch, const int32_t *comp_ids, int32_t n, int32_t *out_cols)
{
    if (!e || !e->built || arch < 0 || arch >= e->host.num_arch || !comp_ids || !out_cols || n < 0) return MDR_EINVAL;
    if (n == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    int32_t *d = nullptr;
    REQ_CUDA(cudaMalloc(&d, 8 * (size_t)n));
    REQ_CUDA(cudaMemcpy(d, comp_ids, 4 * (size_t)n, cudaMemcpyHostToDevice));
    ecsColumnLookupKernel<<<(n + 127) / 128, 128, 0, e->stream>>>(e->dev, arch, d, n, d + n);
    REQ_CUDA(cudaGetLastError());
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    REQ_CUDA(cudaMemcpy(out_cols, d + n, 4 * (size_t)n, cudaMemcpyDeviceToHost));
    REQ_CUDA(cudaFree(d));
    return MDR_OK;
}

void *mdr_ecs_get_exported(mdr_ecs *e, int32_t slot)
{
    if (!e || !e->built || slot < 0 || slot >= e->host.num_exports) return nullptr;
    return e->host.exports[slot].buffer;
}

int64_t mdr_ecs_export_rows(mdr_ecs *e, int32_t slot)
{
    if (!e || !e->built || slot < 0 || slot >= e->host.num_exports) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    const ExportDev &X = e->host.exports[slot];
    if (X.fixed) return (int64_t)e->host.num_worlds * e->host.arch[X.arch].stride;
    if (e->exports_dirty) refreshExports(e);
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    int64_t total = 0;
    REQ_CUDA(cudaMemcpy(&total, X.offsets + e->host.num_worlds, 8, cudaMemcpyDeviceToHost));
    return total;
}

int64_t mdr_ecs_read_exported(mdr_ecs *e, int32_t slot, void *host, int64_t max_bytes)
{
    if (!e || !host) return MDR_EINVAL;
    const int64_t rows = mdr_ecs_export_rows(e, slot);
    if (rows < 0) return rows;
    const int64_t bytes = rows * e->host.exports[slot].bytes;
    if (bytes > max_bytes) return MDR_ECAPACITY;
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    if (bytes) REQ_CUDA(cudaMemcpy(host, e->host.exports[slot].buffer, (size_t)bytes, cudaMemcpyDeviceToHost));
    return bytes;
}

// One step and its exported columns in host memory with two host synchronisations instead of one per call and
// column (mdr_ecs_step + mdr_ecs_export_rows + mdr_ecs_read_exported per slot: six blocking round trips for two
// slots, ~100 us -- more than a 1024-world step takes on the device). The step, the row totals of the staged
// exports (8 bytes each, to page-locked memory) and the fixed-size columns are enqueued together; after the first
// wait the staged columns are copied at their now known sizes.
int32_t mdr_ecs_step_read(mdr_ecs *e, int32_t num_slots, const int32_t *slots, void *const *host,
                          const int64_t *max_bytes, int64_t *out_bytes)
{
    if (!e || num_slots < 0 || num_slots > kMaxExports || (num_slots && (!slots || !host || !max_bytes || !out_bytes)))
        return MDR_EINVAL;
    if (!e->built) return MDR_ESTATE;
    for (int k = 0; k < num_slots; k++)
        if (slots[k] < 0 || slots[k] >= e->host.num_exports || !host[k]) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(e->cfg.gpu_id));
    if (!e->h_totals) REQ_CUDA(cudaHostAlloc((void **)&e->h_totals, sizeof(int64_t) * kMaxExports, cudaHostAllocDefault));
    int32_t rc = mdr_ecs_step_async(e, 1);
    if (rc != MDR_OK) return rc;
    const int W = e->host.num_worlds;
    bool staged = false;
    for (int k = 0; k < num_slots; k++) {
        const ExportDev &X = e->host.exports[slots[k]];
        if (X.fixed) {
            out_bytes[k] = (int64_t)W * e->host.arch[X.arch].stride * X.bytes;
            if (out_bytes[k] > max_bytes[k]) rc = MDR_ECAPACITY;
            else if (out_bytes[k]) REQ_CUDA(cudaMemcpyAsync(host[k], X.buffer, (size_t)out_bytes[k], cudaMemcpyDeviceToHost, e->stream));
        } else {
            staged = true;
            REQ_CUDA(cudaMemcpyAsync(e->h_totals + k, X.offsets + W, 8, cudaMemcpyDeviceToHost, e->stream));
        }
    }
    REQ_CUDA(cudaStreamSynchronize(e->stream));
    if (staged) {
        for (int k = 0; k < num_slots; k++) {
            const ExportDev &X = e->host.exports[slots[k]];
            if (X.fixed) continue;
            out_bytes[k] = e->h_totals[k] * X.bytes;
            if (out_bytes[k] > max_bytes[k]) rc = MDR_ECAPACITY;
            else if (out_bytes[k]) REQ_CUDA(cudaMemcpyAsync(host[k], X.buffer, (size_t)out_bytes[k], cudaMemcpyDeviceToHost, e->stream));
        }
        REQ_CUDA(cudaStreamSynchronize(e->stream));
    }
    closeTiming(e);
    const int32_t err = checkError(e);
    return rc != MDR_OK ? rc : err;
}

int32_t mdr_ecs_export_bytes_per_row(mdr_ecs *e, int32_t slot)
{
    if (!e || !e->built || slot < 0 || slot >= e->host.num_exports) return MDR_EINVAL;
    return e->host.exports[slot].bytes;
}

// ---- the built-in worlds: what a user's registerTypes / setupTasks / world
// constructor do, expressed through the C ABI above --------------------------------
int32_t mdr_app_setup(mdr_ecs *e, int32_t app, const int32_t *params, int32_t nparams)
{
    if (!e || app < 0 || app >= MDR_APP_CHURN) return MDR_EINVAL;
    if (e->built) return MDR_ESTATE;
    auto P = [&](int i, int32_t dflt) { return params && i < nparams ? params[i] : dflt; };
    using namespace mdr::ecs::apps;
    e->cfg.app = app;
    int32_t rc = MDR_OK;
    auto ok = [&](int32_t v) { if (v < 0 && rc == MDR_OK) rc = v; return v; };
    switch (app) {
    case MDR_APP_SCRIPT: {
        const int32_t cap = P(1, 4096);
        const int32_t A = ok(mdr_ecs_register_component(e, 4, 4)), B = ok(mdr_ecs_register_component(e, 12, 4)),
                      C = ok(mdr_ecs_register_component(e, 16, 16));
        const int32_t x[] = { A }, y[] = { A, B }, z[] = { A, B, C };
        ok(mdr_ecs_register_archetype(e, x, 1, cap, 0));
        ok(mdr_ecs_register_archetype(e, y, 2, cap, 0));
        ok(mdr_ecs_register_archetype(e, z, 3, cap, 0));
        ok(mdr_ecs_register_archetype(e, x, 1, cap, 0));
        if (P(0, 1)) ok(mdr_ecs_register_singleton(e, 4, 4));
        if (e->cfg.max_script_handles <= 0) e->cfg.max_script_handles = P(2, 1 << 16);
        if (rc == MDR_OK) rc = mdr_ecs_build(e);
    } break;
    case MDR_APP_TASKGRAPH: {
        const int32_t n = P(0, 100);
        const int32_t pos = ok(mdr_ecs_register_component(e, 12, 4)), rot = ok(mdr_ecs_register_component(e, 16, 4));
        const int32_t pr[] = { pos, rot };
        const int32_t sphere = ok(mdr_ecs_register_archetype(e, pr, 2, n, 0));
        const int32_t agent = ok(mdr_ecs_register_archetype(e, pr, 2, 1, 0));
        ok(mdr_ecs_export_column(e, agent, pos, 0));    // simple.cpp:46-47
        ok(mdr_ecs_export_column(e, agent, rot, 1));
        const int32_t q[] = { pos };
        const int32_t n0 = ok(mdr_ecs_add_parallel_for(e, kTgDrift, q, 1, nullptr, 0));
        ok(mdr_ecs_add_parallel_for(e, kTgClamp, q, 1, &n0, 1));
        if (rc == MDR_OK) rc = mdr_ecs_build(e);
        if (rc == MDR_OK) rc = mdr_ecs_make_entities(e, sphere, n);
        if (rc == MDR_OK) rc = mdr_ecs_make_entities(e, agent, 1);
    } break;
    case MDR_APP_FANOUT: {
        const int32_t n = P(0, 100), max_cand = P(1, n * (n - 1) / 2 < 4096 ? n * (n - 1) / 2 : 4096);
        const int32_t pos = ok(mdr_ecs_register_component(e, 12, 4)), rot = ok(mdr_ecs_register_component(e, 16, 4)),
                      box = ok(mdr_ecs_register_component(e, 24, 4)), hit = ok(mdr_ecs_register_component(e, 4, 4)),
                      cand = ok(mdr_ecs_register_component(e, 16, 4)), key = ok(mdr_ecs_register_component(e, 4, 4));
        const int32_t oc[] = { pos, rot, box, hit }, cc[] = { cand, key };
        const int32_t obj = ok(mdr_ecs_register_archetype(e, oc, 4, n, 0));
        const int32_t tmp = ok(mdr_ecs_register_archetype(e, cc, 2, max_cand > 0 ? max_cand : 1, 0));
        ok(mdr_ecs_export_column(e, obj, hit, 0));
        const int32_t q1[] = { pos }, q2[] = { pos, rot, box }, q3[] = { 0, box }, q4[] = { cand };
        int32_t d = ok(mdr_ecs_add_clear_tmp(e, tmp, nullptr, 0));
        d = ok(mdr_ecs_add_parallel_for(e, kFanMove, q1, 1, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFanBox, q2, 3, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFanOverlap, q3, 2, &d, 1));
        d = ok(mdr_ecs_add_sort(e, tmp, key, 0, &d, 1));
        ok(mdr_ecs_add_parallel_for(e, kFanConsume, q4, 1, &d, 1));
        if (rc == MDR_OK) rc = mdr_ecs_build(e);
        if (rc == MDR_OK) rc = mdr_ecs_make_entities(e, obj, n);
    } break;
    case MDR_APP_FVS: {
        const int32_t nd = P(0, 50), nk = P(1, 200);
        const int32_t pos = ok(mdr_ecs_register_component(e, 12, 4)), hp = ok(mdr_ecs_register_component(e, 4, 4)),
                      act = ok(mdr_ecs_register_component(e, 4, 4)), mana = ok(mdr_ecs_register_component(e, 4, 4)),
                      quiv = ok(mdr_ecs_register_component(e, 4, 4)), cl = ok(mdr_ecs_register_component(e, 12, 4)),
                      key = ok(mdr_ecs_register_component(e, 4, 4));
        const int32_t dc[] = { pos, hp, act, mana }, kc[] = { pos, hp, act, quiv }, tc[] = { cl, key };
        // respawning keeps the totals constant; leave head-room for a custom graph
        const int32_t dragon = ok(mdr_ecs_register_archetype(e, dc, 4, nd + 16, 0));
        const int32_t knight = ok(mdr_ecs_register_archetype(e, kc, 4, nk + 16, 0));
        const int32_t tracker = ok(mdr_ecs_register_archetype(e, tc, 2, nd + nk, 0));
        const int32_t game = ok(mdr_ecs_register_singleton(e, 4, 4));
        ok(mdr_ecs_export_column(e, dragon, hp, 0));
        ok(mdr_ecs_export_column(e, knight, pos, 1));
        const int32_t q0[] = { 0, pos, act }, q1[] = { 0, act, mana }, q2[] = { 0, act, quiv }, q3[] = { 0, hp },
                      q4[] = { cl }, q5[] = { (int32_t)e->comps.size() - 1 };
        int32_t d = ok(mdr_ecs_add_parallel_for(e, kFvsAction, q0, 3, nullptr, 0));
        d = ok(mdr_ecs_add_parallel_for(e, kFvsCaster, q1, 3, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFvsArcher, q2, 3, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFvsMark, q3, 2, &d, 1));
        d = ok(mdr_ecs_add_sort(e, tracker, key, 0, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFvsDestroy, q4, 1, &d, 1));
        d = ok(mdr_ecs_add_parallel_for(e, kFvsRespawn, q4, 1, &d, 1));
        d = ok(mdr_ecs_add_clear_tmp(e, tracker, &d, 1));
        ok(mdr_ecs_add_parallel_for(e, kFvsTick, q5, 1, &d, 1));
        (void)game;
        if (rc == MDR_OK) rc = mdr_ecs_build(e);
        if (rc == MDR_OK) rc = mdr_ecs_make_entities(e, dragon, nd);
        if (rc == MDR_OK) rc = mdr_ecs_make_entities(e, knight, nk);
    } break;
    }
    return rc;
}

}  // extern "C"
