// CUDA executor of the rigid-body world step: persistent kernel (one block per
// world at a time, worlds strided over a grid sized to the resident-block
// capacity of the GPU), the rare BVH rebuild kernel, and the mdr_phys_* C ABI.
// sm_100a only; compiled with -fmad=false (see phys_math.cuh).
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <vector>

#include "phys_kernel.cuh"
#include "phys_lidar.cuh"
#include "phys_feeder.cuh"
#include "phys_assets.hpp"
#include "obj_hull.hpp"
#include "phys_cols.hpp"

using namespace mdr;

#define REQ_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (expr);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            fprintf(stderr, "madrona_b200: CUDA error %s at %s:%d: %s\n",                \
                    cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_));   \
            abort();                                                                     \
        }                                                                                \
    } while (0)

// ------------------------------------------------------------------ kernels --
// physStepKernel lives in phys_kernel.cuh and is instantiated by phys_inst*.cu (parallel builds).
// The reference's top-down build reduced to traversal order + slot-box reset (phys_step.cuh).
// One WARP per flagged world (an episode reset names few worlds, the others return at once): the lanes compute the
// current leaf boxes and reset the slot boxes row-parallel; the first lane then runs the top-down splits, which are
// sequential by nature (each split partitions what the previous one left).
constexpr int kRebuildWarps = 8;
__global__ void __launch_bounds__(32 * kRebuildWarps) bvhRebuildKernel(const __grid_constant__ StepParams p)
{
    const int w = blockIdx.x * kRebuildWarps + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (w >= p.t.num_worlds || !p.t.rebuild_flag[w]) return;
    const PhysTables &t = p.t;
    const size_t r0 = (size_t)w * t.stride;
    const int n = t.row_count[w];
    for (int i = lane; i < n; i += 32) {
        const uint32_t oid = (uint32_t)t.obj[r0 + i];
        const ObjDev *o = p.a.objs + (oid < (uint32_t)p.a.num_objs ? oid : 0u);      // flagged by the step kernel
        Box wb = applyTRS(objAABB(o), ld3(t.pos + 3 * r0, i), ld4(t.rot + 4 * r0, i), ld3(t.scale + 3 * r0, i));
        const Box leaf = expandWithMotion(wb, ldVelLin(t.vel + 6 * r0, i), p.vel_expand, p.accel_expand);
        stBox(t.leaf_aabb + 6 * r0, i, leaf);
        stBox(t.slot_box + 6 * r0, i, leaf);
        t.sorted[r0 + i] = (uint16_t)i;
    }
    __syncwarp();
    if (lane == 0) {
        p.t.rebuild_flag[w] = 0;
        bvhRankOrder(t.leaf_aabb + 6 * r0, t.order + r0, t.sorted + r0, n);
    }
}

static void launchRebuild(const StepParams &params, int num_worlds, cudaStream_t stream)
{
    bvhRebuildKernel<<<(num_worlds + kRebuildWarps - 1) / kRebuildWarps, 32 * kRebuildWarps, 0, stream>>>(params);
}

// -------------------------------------------------------------- host state --
// Compiled register budgets, gathered from the instantiation units (ascending threads, then blocks).
static std::vector<KernelVariant> gatherVariants()
{
    std::vector<KernelVariant> all;
    const KernelVariant *(*parts[])(int *) = { mdrStepVariants0, mdrStepVariants1, mdrStepVariants2,
                                                mdrStepVariants3, mdrStepVariants4 };
    for (auto part : parts) {
        int n = 0;
        const KernelVariant *v = part(&n);
        all.insert(all.end(), v, v + n);
    }
    return all;
}
static const std::vector<KernelVariant> kVariants = gatherVariants();

struct mdr_phys {
    mdr_phys_config cfg;
    HostAssets assets;
    StepParams params;
    std::vector<void *> allocs;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // mdr_phys_step_io: copy-in / copy-out streams and per-chunk events
    cudaStream_t io_in = nullptr, io_out = nullptr;
    cudaStream_t lane_stream[kMaxLanes] = {};      // [0] = stream; the world ranges of step_io rotate over io_lanes of them
    int io_lanes = 3;
    std::vector<cudaEvent_t> io_events;
    // mdr_phys_trace_rays: device staging of the host-side ray buffers
    char *ray_buf = nullptr;
    size_t ray_buf_bytes = 0;
    bool timing_open = false;
    int pending_launches = 0;                // steps (or world ranges of step_io) between ev0 and ev1
    int last_call_launches = 0;              // step kernels the last mdr_phys_step_async call launched
    bool io_trace_on = false;                // MDR_PHYS_IO_TRACE=1
    std::vector<cudaEvent_t> io_trace;
    int io_trace_ranges = 0;
    bool batch_ok = true;                    // n > 1 steps of one call run as one launch (MDR_PHYS_NO_BATCH=1: n launches)
    uint32_t batch_step_base = 0;            // steps done inside batches so far: the epoch of StepParams::group_done
    double timed_ms = 0.0;
    int timed_launches = 0;
    bool external_tables = false;            // columns and row counts belong to an ECS archetype
    StepKernel kernel = nullptr;             // the variant in use: kernel_plain, or kernel_joints once joints are set
    StepKernel kernel_plain = nullptr, kernel_joints = nullptr;
    bool fixed_layout = false;               // the plain kernel is a compile-time-layout instantiation
    bool big_hulls = false;                  // an object has more than 32 vertices / faces: the two-slot SAT variant
    int threads = 0, blocks = 0, blocks_per_sm = 0, overflow_per_slot = 0;
    int threads_per_world = 0, worlds_per_block = 0, lockstep = 1;
    size_t smem_bytes = 0, slot_bytes = 0;
    bool maybe_rebuild = true;
    int32_t *d_err = nullptr;                // device view of h_err
    volatile int32_t *h_err = nullptr;       // the step's error word: page-locked host memory mapped into the device, so
                                             // that mdr_phys_sync reads it without a blocking 4-byte copy every step
    ProbeDev d_probe;   // device buffers of the probe
    int probe_pairs_cap = 0;
    JointDev *d_joints = nullptr;       // mdr_phys_set_joints
    int32_t *d_joint_count = nullptr;

    bool alloc_failed = false;   // a table did not fit the device: creation returns MDR_ECAPACITY
    template <typename T> T *dalloc(size_t n, bool zero = true)
    {
        void *p = nullptr;
        size_t bytes = (n ? n : 1) * sizeof(T);
        if (alloc_failed) return nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess) {
            cudaGetLastError();          // out of memory is not sticky
            alloc_failed = true;
            return nullptr;
        }
        if (zero) REQ_CUDA(cudaMemset(p, 0, bytes));
        allocs.push_back(p);
        return (T *)p;
    }
    template <typename T> T *upload(const std::vector<T> &v)
    {
        T *d = dalloc<T>(v.size(), false);
        if (d) REQ_CUDA(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        return d;
    }
};

// Called by the ASYNC entry points before they enqueue more work: the batch already timed is closed only
// if it has finished (cudaEventQuery); otherwise it stays open and the new launches extend it, so that
// back-to-back asynchronous calls never block the host (the executor's stream keeps them in order).
static bool timingStillRunning(mdr_phys *p)
{
    if (!p->timing_open) return false;
    if (cudaEventQuery(p->ev1) == cudaSuccess) return false;
    cudaGetLastError();      // cudaErrorNotReady is not an error
    return true;
}

static void closeTiming(mdr_phys *p)
{
    if (!p->timing_open) return;
    float ms = 0.f;
    REQ_CUDA(cudaEventSynchronize(p->ev1));
    REQ_CUDA(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
    p->timed_ms += ms;
    p->timed_launches += p->pending_launches;
    p->pending_launches = 0;
    p->timing_open = false;
}

// The device records one contact per PAIR SLOT; the reference's list is the
// slots that produced a contact, in order.
static void compactProbeContacts(ContactDev *contacts, int32_t *num_contacts, int substeps, int max_contacts)
{
    for (int s = 0; s < substeps; s++) {
        ContactDev *row = contacts + (size_t)s * max_contacts;
        int out = 0;
        for (int i = 0; i < max_contacts; i++) {
            if (row[i].num_points <= 0) continue;
            ContactDev c = row[i];
            row[out++] = c;
        }
        for (int i = out; i < max_contacts; i++) memset(&row[i], 0, sizeof(ContactDev));
        num_contacts[s] = out;
    }
}

extern "C" {

const char *mdr_version(void) { return "madrona_b200 0.1 (sm_100a)"; }

int32_t mdr_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int32_t mdr_phys_column_bytes(int32_t col)
{
    return col >= 0 && col < MDR_NUM_PHYS_COLS ? kColBytes[col] : 0;
}

}   // extern "C"

// `ext` non-null: the rigid-body columns are somebody else's tables (an ECS archetype,
// mdr_phys_create_on_columns); only the BVH singleton state and scratch are allocated here.
struct ExternalTables { int32_t stride; void *const *cols; int32_t *row_count; };

static int32_t createImpl(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                          const float *verts, const uint32_t *poly_words, const ExternalTables *ext, mdr_phys **out)
{
    if (!cfg || !out || !objs || num_objs <= 0 || cfg->num_worlds <= 0 ||
        cfg->max_bodies_per_world <= 0 || cfg->max_bodies_per_world > MDR_MAX_BODIES ||
        cfg->num_substeps <= 0 || cfg->max_contacts <= 0 || !(cfg->delta_t > 0.f))
        return MDR_EINVAL;
    if (cfg->gpu_id < 0 || cfg->gpu_id >= mdr_device_count()) return MDR_ENODEVICE;
    REQ_CUDA(cudaSetDevice(cfg->gpu_id));
    cudaDeviceProp prop;
    REQ_CUDA(cudaGetDeviceProperties(&prop, cfg->gpu_id));
    if (prop.major != 10) {
        fprintf(stderr, "madrona_b200: device %d is sm_%d%d; this library is built for sm_100a only\n",
                cfg->gpu_id, prop.major, prop.minor);
        return MDR_ENODEVICE;
    }

    mdr_phys *p = new mdr_phys;
    p->cfg = *cfg;
    if (!buildAssets(p->assets, objs, num_objs, verts, poly_words)) {
        fprintf(stderr, "madrona_b200: %s\n", p->assets.error.c_str());
        delete p;
        return MDR_EINVAL;
    }
    memset(&p->params, 0, sizeof(p->params));
    StepParams &sp = p->params;
    fillStepConstants(sp, *cfg);
    const int W = cfg->num_worlds;
    const int stride = ext ? ext->stride : (cfg->max_bodies_per_world + 3) / 4 * 4;   // every column tile 16 B aligned
    if (stride < cfg->max_bodies_per_world || stride % 4) { delete p; return MDR_EINVAL; }
    const size_t R = (size_t)W * stride;
    PhysTables &t = sp.t;
    t.stride = stride;
    t.num_worlds = W;
    for (int c = 0; c < MDR_NUM_PHYS_COLS; c++) *colSlot(t, c) = ext ? ext->cols[c] : (void *)p->dalloc<char>(R * kColBytes[c]);
    t.row_count = ext ? ext->row_count : p->dalloc<int32_t>(W);
    p->external_tables = ext != nullptr;
    t.leaf_aabb = p->dalloc<float>(R * 6);
    t.slot_box = p->dalloc<float>(R * 6);
    t.order = p->dalloc<uint16_t>(R);
    t.sorted = p->dalloc<uint16_t>(R);
    t.rebuild_flag = p->dalloc<uint8_t>(W);
    sp.a.objs = p->upload(p->assets.objs);
    sp.a.verts = p->upload(p->assets.verts);
    sp.a.planes = p->upload(p->assets.planes);
    sp.a.hedges = p->upload(p->assets.hedges);
    sp.a.edges = p->upload(p->assets.edges);
    sp.a.polys = p->upload(p->assets.polys);
    sp.a.num_objs = num_objs;
    sp.a.max_verts = p->assets.max_verts;
    sp.a.max_faces = p->assets.max_faces;
    sp.max_pairs = cfg->max_contacts;
    sp.smem_contacts = cfg->smem_contacts > 0 ? cfg->smem_contacts : 71;
    if (sp.smem_contacts > sp.max_pairs) sp.smem_contacts = sp.max_pairs;
    sp.sat_hint = p->dalloc<uint8_t>((size_t)W * ((sp.max_pairs + 3) / 4 * 4));
    sp.world_pairs = p->dalloc<int32_t>(W);
    sp.world_contacts = p->dalloc<int32_t>(W);
    {
        void *h = nullptr, *d = nullptr;
        if (cudaHostAlloc(&h, 64, cudaHostAllocMapped) != cudaSuccess || cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) {
            cudaGetLastError();
            if (h) cudaFreeHost(h);
            p->alloc_failed = true;
        } else {
            memset(h, 0, 64);
            p->h_err = (volatile int32_t *)h;
            p->d_err = (int32_t *)d;
        }
    }
    sp.error_flags = p->d_err;
    sp.work = p->dalloc<uint32_t>(2 * kMaxLanes);
    sp.group_done = p->dalloc<uint32_t>(W);
    sp.probe.world = -1;
#ifdef MDR_PHASE_CLOCKS
    sp.phase_clocks = p->dalloc<unsigned long long>(32);
#endif

    // Launch geometry. A block steps `worlds_per_block` worlds side by side with
    // `threads_per_world` threads each; the grid is the resident capacity of the GPU.
    auto envInt = [](const char *name, int dflt) {
        const char *e = getenv(name);
        return e && *e ? atoi(e) : dflt;
    };
    int tpw = envInt("MDR_THREADS_PER_WORLD", cfg->block_threads > 0 ? cfg->block_threads : 64);
    int wpb = envInt("MDR_WORLDS_PER_BLOCK", cfg->worlds_per_block > 0 ? cfg->worlds_per_block : 4);
    int want_min_blocks = envInt("MDR_MIN_BLOCKS", 0);
    p->lockstep = envInt("MDR_LOCKSTEP", 1);
    p->batch_ok = !envInt("MDR_PHYS_NO_BATCH", 0);
    p->io_lanes = envInt("MDR_PHYS_IO_LANES", 3);
    if (p->io_lanes < 1) p->io_lanes = 1;
    if (p->io_lanes > kMaxLanes) p->io_lanes = kMaxLanes;
    p->io_trace_on = envInt("MDR_PHYS_IO_TRACE", 0) != 0;
    if (!p->lockstep && wpb > 15) wpb = 15;     // one named barrier per world slot
    if (tpw < 32 || tpw % 32 || wpb < 1 || tpw * wpb > 1024) {
        fprintf(stderr, "madrona_b200: threads per world must be a multiple of 32 and threads x worlds per block <= 1024\n");
        mdr_phys_destroy(p);
        return MDR_EINVAL;
    }
    if (wpb > W) wpb = W;
    WorldView v;
    p->slot_bytes = carveWorld(v, nullptr, stride, sp.max_pairs, sp.smem_contacts);
    p->smem_bytes = p->slot_bytes * wpb;
    while (wpb > 1 && p->smem_bytes > (size_t)prop.sharedMemPerBlockOptin) {
        wpb--;
        p->smem_bytes = p->slot_bytes * wpb;
    }
    if (p->smem_bytes > (size_t)prop.sharedMemPerBlockOptin) {
        fprintf(stderr, "madrona_b200: one world needs %zu B of shared memory, the device offers %zu B per block\n",
                p->smem_bytes, (size_t)prop.sharedMemPerBlockOptin);
        mdr_phys_destroy(p);
        return MDR_ECAPACITY;
    }
    p->threads_per_world = tpw;
    p->worlds_per_block = wpb;
    p->threads = tpw * wpb;
    // How many such blocks fit one SM by shared memory decides the register budget.
    int fit = (int)((size_t)prop.sharedMemPerMultiprocessor / (p->smem_bytes + 1024));
    if (fit < 1) fit = 1;
    const KernelVariant *var = nullptr;
    for (const KernelVariant &kv : kVariants) {
        if (kv.threads < p->threads) continue;
        if (var && kv.threads > var->threads) break;
        if (!var) var = &kv;
        if (want_min_blocks ? kv.min_blocks == want_min_blocks : kv.min_blocks <= fit) var = &kv;
    }
    if (!var) { mdr_phys_destroy(p); return MDR_EINVAL; }
    p->big_hulls = p->assets.max_verts > 32 || p->assets.max_faces > 32;
    p->kernel = p->kernel_plain = p->big_hulls ? var->fn_big : var->fn;
    p->kernel_joints = p->big_hulls ? var->fn_big : var->fn_joints;
    // a variant compiled for exactly this shape (phys_kernel.cuh, FixedLayout); MDR_NO_FIXED_LAYOUT=1 keeps the runtime one
    if (!p->big_hulls && var->fn_fixed && !envInt("MDR_NO_FIXED_LAYOUT", 0) && p->lockstep && wpb <= kDenseSlots &&
        stride == var->fixed_np && sp.max_pairs == var->fixed_pairs && sp.smem_contacts == var->fixed_contacts &&
        tpw == var->fixed_tpw) {
        p->kernel = p->kernel_plain = var->fn_fixed;
        p->fixed_layout = true;
    }
    for (StepKernel k : { p->kernel_plain, p->kernel_joints }) {
        REQ_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_bytes));
        REQ_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    }
    int per_sm = 0;
    REQ_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, p->kernel, p->threads, p->smem_bytes));
    if (per_sm < 1) per_sm = 1;
    // Shared memory and L1 share 256 KB of an SM: ask for no more shared memory than the resident blocks need, the
    // rest is L1 for the object tables (vertices, planes, half-edges) and the row columns read in place. The hint is a
    // percentage of the largest carve-out and the driver rounds it UP to a supported size: rounding down here asks
    // for the smallest size that still holds the blocks (collisions-64: 2 x (93 504 + 5 792 + 1 024) B = 195.9 KB
    // fits the 196 KB carve-out exactly, 60 KB of L1 instead of 28: +1.8 % settled, +5 % falling).
    {
        int max_sm = 0, reserved = 0;
        REQ_CUDA(cudaDeviceGetAttribute(&max_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, cfg->gpu_id));
        REQ_CUDA(cudaDeviceGetAttribute(&reserved, cudaDevAttrReservedSharedMemoryPerBlock, cfg->gpu_id));
        for (StepKernel k : { p->kernel_plain, p->kernel_joints }) {
            cudaFuncAttributes fa;
            REQ_CUDA(cudaFuncGetAttributes(&fa, k));
            const long long need = (long long)per_sm * ((long long)p->smem_bytes + (long long)fa.sharedSizeBytes + reserved);
            // the carve-outs of compute capability 9.0 / 10.0; the smallest one that holds the resident blocks, asked for
            // as the percentage just below it (never one that the round-up could land under the need)
            static const int kCarveKB[] = { 0, 8, 16, 32, 64, 100, 132, 164, 196, 228 };
            long long pick = max_sm;
            for (int kb : kCarveKB) if ((long long)kb * 1024 >= need && (long long)kb * 1024 <= max_sm) { pick = (long long)kb * 1024; break; }
            int hint = max_sm > 0 ? (int)(pick * 100 / max_sm) : 100;
            if (hint > 100 || pick >= max_sm) hint = 100;
            hint = envInt("MDR_CARVEOUT", hint);
            REQ_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, hint));
        }
    }
    p->blocks_per_sm = per_sm;
    long long grid = (long long)per_sm * prop.multiProcessorCount;
    const long long need = ((long long)W + wpb - 1) / wpb;
    if (grid > need) grid = need;
    p->blocks = (int)grid;
    p->overflow_per_slot = sp.max_pairs - sp.smem_contacts;
    // one set per lane: the world ranges of mdr_phys_step_io_async run on a few streams, side by side at their ends
    sp.overflow = p->dalloc<SolverContact>((size_t)p->io_lanes * p->blocks * wpb * (p->overflow_per_slot + 1), false);
    if (p->alloc_failed) {
        fprintf(stderr, "madrona_b200: %d worlds x %d bodies (%d contact slots) do not fit the device memory\n",
                W, cfg->max_bodies_per_world, sp.max_pairs);
        mdr_phys_destroy(p);
        return MDR_ECAPACITY;
    }

    REQ_CUDA(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    REQ_CUDA(cudaEventCreate(&p->ev0));
    REQ_CUDA(cudaEventCreate(&p->ev1));
    memset(&p->d_probe, 0, sizeof(p->d_probe));
    *out = p;
    return MDR_OK;
}

// The physics node of an ECS task graph (mdr_ecs_add_physics): the BVH rebuild check and one step on
// the caller's stream, capturable into its run graph. Timing belongs to the caller.
int32_t mdrPhysEnqueueStep(mdr_phys *p, cudaStream_t stream)
{
    if (!p) return MDR_EINVAL;
    const int W = p->cfg.num_worlds;
    launchRebuild(p->params, W, stream);
    p->kernel<<<p->blocks, p->threads, p->smem_bytes, stream>>>(
        p->params, p->threads_per_world, (int)p->slot_bytes, p->overflow_per_slot, p->lockstep, 0, W, 1, 0u, 0, 0);
    return cudaGetLastError() == cudaSuccess ? MDR_OK : MDR_ESTATE;
}

extern "C" {

int32_t mdr_phys_create(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                        const float *verts, const uint32_t *poly_words, mdr_phys **out)
{
    return createImpl(cfg, objs, num_objs, verts, poly_words, nullptr, out);
}

int32_t mdr_phys_create_on_columns(const mdr_phys_config *cfg, const mdr_object *objs, int32_t num_objs,
                                   const float *verts, const uint32_t *poly_words, int32_t row_stride,
                                   void *const *columns, int32_t *row_counts, mdr_phys **out)
{
    if (!columns || !row_counts || row_stride <= 0) return MDR_EINVAL;
    for (int c = 0; c < MDR_NUM_PHYS_COLS; c++)
        if (!columns[c] || ((uintptr_t)columns[c] & 15u)) return MDR_EINVAL;
    const ExternalTables ext { row_stride, columns, row_counts };
    return createImpl(cfg, objs, num_objs, verts, poly_words, &ext, out);
}

void mdr_phys_destroy(mdr_phys *p)
{
    if (!p) return;
    cudaSetDevice(p->cfg.gpu_id);
    if (p->stream) { cudaStreamSynchronize(p->stream); cudaStreamDestroy(p->stream); }
    if (p->ray_buf) cudaFree(p->ray_buf);
    if (p->d_joints) cudaFree(p->d_joints);
    if (p->d_joint_count) cudaFree(p->d_joint_count);
    if (p->io_in) { cudaStreamSynchronize(p->io_in); cudaStreamDestroy(p->io_in); }
    if (p->io_out) { cudaStreamSynchronize(p->io_out); cudaStreamDestroy(p->io_out); }
    for (int l = 1; l < kMaxLanes; l++) if (p->lane_stream[l]) { cudaStreamSynchronize(p->lane_stream[l]); cudaStreamDestroy(p->lane_stream[l]); }
    for (cudaEvent_t e : p->io_events) cudaEventDestroy(e);
    for (cudaEvent_t e : p->io_trace) cudaEventDestroy(e);
    if (p->ev0) cudaEventDestroy(p->ev0);
    if (p->ev1) cudaEventDestroy(p->ev1);
    for (void *a : p->allocs) cudaFree(a);
    if (p->h_err) cudaFreeHost((void *)p->h_err);
    delete p;
}

// One world's row count (entity ids and leaf ids 0..rows-1, as mdr_phys_set_row_counts gives every world)
// and its BVH flagged for rebuild: the structural half of a per-world reset.
int32_t mdr_phys_set_world_rows(mdr_phys *p, int32_t world, int32_t rows)
{
    if (!p || world < 0 || world >= p->cfg.num_worlds || rows < 0 || rows > p->cfg.max_bodies_per_world) return MDR_EINVAL;
    if (p->external_tables) return MDR_ESTATE;      // the ECS owns entities and row counts
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    PhysTables &t = p->params.t;
    std::vector<int32_t> ent((size_t)t.stride * 2), leaf((size_t)t.stride);
    for (int i = 0; i < t.stride; i++) {
        ent[2 * i] = 0;
        ent[2 * i + 1] = leaf[i] = i < rows ? i : -1;
    }
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    REQ_CUDA(cudaMemcpy(t.row_count + world, &rows, 4, cudaMemcpyHostToDevice));
    REQ_CUDA(cudaMemcpy(t.entity + 2 * (size_t)world * t.stride, ent.data(), ent.size() * 4, cudaMemcpyHostToDevice));
    REQ_CUDA(cudaMemcpy(t.leaf_id + (size_t)world * t.stride, leaf.data(), leaf.size() * 4, cudaMemcpyHostToDevice));
    return mdr_phys_reset_worlds(p, &world, 1);
}

int32_t mdr_phys_set_row_counts(mdr_phys *p, const int32_t *rows)
{
    if (!p || !rows) return MDR_EINVAL;
    if (p->external_tables) return MDR_ESTATE;      // the ECS owns entities and row counts
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    PhysTables &t = p->params.t;
    const int W = t.num_worlds;
    std::vector<int32_t> ent((size_t)W * t.stride * 2), leaf((size_t)W * t.stride);
    for (int w = 0; w < W; w++) {
        if (rows[w] < 0 || rows[w] > p->cfg.max_bodies_per_world) return MDR_EINVAL;
        for (int i = 0; i < t.stride; i++) {
            size_t r = (size_t)w * t.stride + i;
            ent[2 * r] = 0;
            ent[2 * r + 1] = i < rows[w] ? i : -1;
            leaf[r] = i < rows[w] ? i : -1;
        }
    }
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    REQ_CUDA(cudaMemcpy(t.row_count, rows, 4 * (size_t)W, cudaMemcpyHostToDevice));
    REQ_CUDA(cudaMemcpy(t.entity, ent.data(), ent.size() * 4, cudaMemcpyHostToDevice));
    REQ_CUDA(cudaMemcpy(t.leaf_id, leaf.data(), leaf.size() * 4, cudaMemcpyHostToDevice));
    return mdr_phys_flag_rebuild(p);
}

static_assert(sizeof(JointDev) == sizeof(mdr_joint), "JointDev mirrors mdr_joint");

int32_t mdr_phys_set_joints(mdr_phys *p, int32_t max_joints, const int32_t *counts, const mdr_joint *joints)
{
    if (!p || max_joints < 0 || (max_joints > 0 && (!counts || !joints))) return MDR_EINVAL;
    const int W = p->params.t.num_worlds;
    for (int w = 0; w < W && max_joints > 0; w++) {
        if (counts[w] < 0 || counts[w] > max_joints) return MDR_EINVAL;
        for (int k = 0; k < counts[w]; k++) {
            const mdr_joint &j = joints[(size_t)w * max_joints + k];
            if (j.type < 0 || j.type > 1 || j.row1 < 0 || j.row2 < 0 ||
                j.row1 >= p->cfg.max_bodies_per_world || j.row2 >= p->cfg.max_bodies_per_world)
                return MDR_EINVAL;
        }
    }
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    if (max_joints > 0) {       // joint rows must be live rows of their world
        std::vector<int32_t> rows((size_t)W);
        REQ_CUDA(cudaMemcpy(rows.data(), p->params.t.row_count, 4 * (size_t)W, cudaMemcpyDeviceToHost));
        for (int w = 0; w < W; w++)
            for (int k = 0; k < counts[w]; k++) {
                const mdr_joint &j = joints[(size_t)w * max_joints + k];
                if (j.row1 >= rows[w] || j.row2 >= rows[w]) return MDR_EINVAL;
            }
    }
    if (p->d_joints) { REQ_CUDA(cudaFree(p->d_joints)); p->d_joints = nullptr; }
    if (p->d_joint_count) { REQ_CUDA(cudaFree(p->d_joint_count)); p->d_joint_count = nullptr; }
    if (max_joints > 0) {
        const size_t nj = (size_t)W * max_joints;
        REQ_CUDA(cudaMalloc((void **)&p->d_joints, nj * sizeof(JointDev)));
        REQ_CUDA(cudaMalloc((void **)&p->d_joint_count, (size_t)W * 4));
        REQ_CUDA(cudaMemcpy(p->d_joints, joints, nj * sizeof(JointDev), cudaMemcpyHostToDevice));
        REQ_CUDA(cudaMemcpy(p->d_joint_count, counts, (size_t)W * 4, cudaMemcpyHostToDevice));
    }
    p->params.joints = p->d_joints;
    p->params.joint_count = p->d_joint_count;
    p->params.joint_stride = max_joints;
    // Same launch bounds, hence the same residency: only the code path differs.
    p->kernel = max_joints > 0 ? p->kernel_joints : p->kernel_plain;
    return MDR_OK;
}

int32_t mdr_phys_write_column(mdr_phys *p, int32_t col, const void *src)
{
    if (!p || !src) return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t nb = kColBytes[col], n = p->cfg.max_bodies_per_world;
    if (col == MDR_COL_OBJECT_ID) {     // the step indexes the shared object table with it
        const int32_t *ids = (const int32_t *)src;
        for (size_t i = 0; i < n * (size_t)p->cfg.num_worlds; i++)
            if (ids[i] < 0 || ids[i] >= p->params.a.num_objs) return MDR_EINVAL;
    }
    REQ_CUDA(cudaMemcpy2DAsync(*s, p->params.t.stride * nb, src, n * nb, n * nb, p->cfg.num_worlds,
                               cudaMemcpyHostToDevice, p->stream));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    return MDR_OK;
}

int32_t mdr_phys_read_column(mdr_phys *p, int32_t col, void *dst)
{
    if (!p || !dst) return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t nb = kColBytes[col], n = p->cfg.max_bodies_per_world;
    REQ_CUDA(cudaMemcpy2DAsync(dst, n * nb, *s, p->params.t.stride * nb, n * nb, p->cfg.num_worlds,
                               cudaMemcpyDeviceToHost, p->stream));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    return MDR_OK;
}

int32_t mdr_phys_row_stride(mdr_phys *p) { return p ? p->params.t.stride : MDR_EINVAL; }

int32_t mdr_phys_write_column_raw_async(mdr_phys *p, int32_t col, const void *src)
{
    if (!p || !src) return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t bytes = (size_t)kColBytes[col] * p->params.t.stride * p->cfg.num_worlds;
    REQ_CUDA(cudaMemcpyAsync(*s, src, bytes, cudaMemcpyHostToDevice, p->stream));
    return MDR_OK;
}

int32_t mdr_phys_read_column_raw_async(mdr_phys *p, int32_t col, void *dst)
{
    if (!p || !dst) return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t bytes = (size_t)kColBytes[col] * p->params.t.stride * p->cfg.num_worlds;
    REQ_CUDA(cudaMemcpyAsync(dst, *s, bytes, cudaMemcpyDeviceToHost, p->stream));
    return MDR_OK;
}

void *mdr_host_alloc(uint64_t bytes)
{
    void *ptr = nullptr;
    if (cudaHostAlloc(&ptr, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return ptr;
}

void mdr_host_free(void *ptr) { if (ptr) cudaFreeHost(ptr); }

// The step's quaternion arithmetic (csrc/phys_math.cuh, the host instantiation of the same source) for the
// reference's known-answer tests (tests/math.cpp:26-51): op 0 angleAxis(a[0], a[1..3]), 1 a * b, 2 rotateVec(a, b[0..2]),
// 3 normalize(a), 4 inv(a).
int32_t mdr_math_quat(int32_t op, const float *a, const float *b, float *out)
{
    if (!a || !out || op < 0 || op > 4 || ((op == 1 || op == 2) && !b)) return MDR_EINVAL;
    Q4 r = mk4(0.f, 0.f, 0.f, 0.f);
    switch (op) {
    case 0: r = qangleAxis(a[0], mk3(a[1], a[2], a[3])); break;
    case 1: r = mk4(a[0], a[1], a[2], a[3]) * mk4(b[0], b[1], b[2], b[3]); break;
    case 2: { V3 v = rotate(mk4(a[0], a[1], a[2], a[3]), mk3(b[0], b[1], b[2])); r = mk4(0.f, v.x, v.y, v.z); } break;
    case 3: r = qnormalize(mk4(a[0], a[1], a[2], a[3])); break;
    case 4: r = qinv(mk4(a[0], a[1], a[2], a[3])); break;
    }
    out[0] = r.w; out[1] = r.x; out[2] = r.y; out[3] = r.z;
    return MDR_OK;
}

void *mdr_phys_done_event(mdr_phys *p) { return p ? (void *)p->ev1 : nullptr; }

int32_t mdr_stream_wait_event(void *stream, void *event)
{
    if (!event) return MDR_EINVAL;
    return cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0) == cudaSuccess ? MDR_OK : MDR_ESTATE;
}

void *mdr_phys_column_ptr(mdr_phys *p, int32_t col)
{
    if (!p) return nullptr;
    void **s = colSlot(p->params.t, col);
    return s ? *s : nullptr;
}

#ifdef MDR_PHASE_CLOCKS
// profiling builds only: reads and clears the per-phase clock64 sums (32 slots)
int32_t mdr_phys_phase_clocks(mdr_phys *p, unsigned long long *out)
{
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    REQ_CUDA(cudaMemcpy(out, p->params.phase_clocks, 32 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    REQ_CUDA(cudaMemset(p->params.phase_clocks, 0, 32 * sizeof(unsigned long long)));
    return MDR_OK;
}
#endif

int32_t mdr_phys_flag_rebuild(mdr_phys *p)
{
    if (!p) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaMemsetAsync(p->params.t.rebuild_flag, 1, p->cfg.num_worlds, p->stream));
    p->maybe_rebuild = true;
    return MDR_OK;
}

// RigidBodyPhysicsSystem::reset of the listed worlds only (physics.cpp:1038-1043: the context's own BVH is
// flagged, nobody else's): their slot boxes are reset and their traversal order rebuilt before the next
// step; every other world keeps its grown boxes and its pair order, as in the reference.
int32_t mdr_phys_reset_worlds(mdr_phys *p, const int32_t *worlds, int32_t n)
{
    if (!p || n < 0 || (n > 0 && !worlds)) return MDR_EINVAL;
    const int W = p->cfg.num_worlds;
    for (int i = 0; i < n; i++) if (worlds[i] < 0 || worlds[i] >= W) return MDR_EINVAL;
    if (n == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    // runs of consecutive worlds become one memset each (an episode reset names few worlds)
    std::vector<int32_t> sorted(worlds, worlds + n);
    std::sort(sorted.begin(), sorted.end());
    for (int i = 0; i < n;) {
        int j = i + 1;
        while (j < n && sorted[j] <= sorted[j - 1] + 1) j++;
        REQ_CUDA(cudaMemsetAsync(p->params.t.rebuild_flag + sorted[i], 1, (size_t)(sorted[j - 1] - sorted[i] + 1), p->stream));
        i = j;
    }
    p->maybe_rebuild = true;
    return MDR_OK;
}

// One world's rows of a column, host -> device ([rows][bytes], rows <= max_bodies_per_world): the
// per-world half of an episode reset (the reference writes components through ctx.getUnsafe in the
// world's own reset code).
int32_t mdr_phys_write_world_column(mdr_phys *p, int32_t col, int32_t world, int32_t rows, const void *src)
{
    if (!p || !src || world < 0 || world >= p->cfg.num_worlds || rows < 0 || rows > p->cfg.max_bodies_per_world)
        return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    if (col == MDR_COL_OBJECT_ID) {
        const int32_t *ids = (const int32_t *)src;
        for (int i = 0; i < rows; i++) if (ids[i] < 0 || ids[i] >= p->params.a.num_objs) return MDR_EINVAL;
    }
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t nb = kColBytes[col];
    REQ_CUDA(cudaMemcpyAsync((char *)*s + (size_t)world * p->params.t.stride * nb, src, rows * nb,
                             cudaMemcpyHostToDevice, p->stream));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    return MDR_OK;
}

int32_t mdr_phys_read_world_column(mdr_phys *p, int32_t col, int32_t world, int32_t rows, void *dst)
{
    if (!p || !dst || world < 0 || world >= p->cfg.num_worlds || rows < 0 || rows > p->cfg.max_bodies_per_world)
        return MDR_EINVAL;
    void **s = colSlot(p->params.t, col);
    if (!s) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t nb = kColBytes[col];
    REQ_CUDA(cudaMemcpyAsync(dst, (const char *)*s + (size_t)world * p->params.t.stride * nb, rows * nb,
                             cudaMemcpyDeviceToHost, p->stream));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    return MDR_OK;
}

int32_t mdr_phys_step_async(mdr_phys *p, int32_t num_steps)
{
    if (!p || num_steps < 0) return MDR_EINVAL;
    if (num_steps == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const bool extend = timingStillRunning(p);
    if (!extend) closeTiming(p);
    if (p->maybe_rebuild) {
        const int W = p->cfg.num_worlds;
        launchRebuild(p->params, W, p->stream);
        REQ_CUDA(cudaGetLastError());
        p->maybe_rebuild = false;
    }
    if (!extend) REQ_CUDA(cudaEventRecord(p->ev0, p->stream));
    // A call of n > 1 steps is a batch nobody observes in between: ONE launch walks (step, world group) tasks in
    // step order, a group's step waiting for its previous one (phys_kernel.cuh). MDR_PHYS_NO_BATCH=1: n launches.
    if (num_steps > 1 && p->batch_ok) {
        p->kernel<<<p->blocks, p->threads, p->smem_bytes, p->stream>>>(
            p->params, p->threads_per_world, (int)p->slot_bytes, p->overflow_per_slot, p->lockstep,
            0, p->cfg.num_worlds, num_steps, p->batch_step_base, 0, 0);
        p->batch_step_base += (uint32_t)num_steps;
        p->last_call_launches = 1;
    } else {
        for (int s = 0; s < num_steps; s++) {
            p->kernel<<<p->blocks, p->threads, p->smem_bytes, p->stream>>>(
                p->params, p->threads_per_world, (int)p->slot_bytes, p->overflow_per_slot, p->lockstep,
                0, p->cfg.num_worlds, 1, 0u, 0, 0);
        }
        p->last_call_launches = num_steps;
    }
    REQ_CUDA(cudaGetLastError());
    REQ_CUDA(cudaEventRecord(p->ev1, p->stream));
    p->timing_open = true;
    p->pending_launches = (extend ? p->pending_launches : 0) + num_steps;
    return MDR_OK;
}

// One step with host-side inputs and outputs, pipelined over world ranges: the
// copy-in of range k+1 and the copy-out of range k-1 run beside the step of
// range k (three streams). A range is a whole number of resident waves.
int32_t mdr_phys_step_io_async(mdr_phys *p, int32_t num_in, const int32_t *in_cols, const void *const *in_host,
                               int32_t num_out, const int32_t *out_cols, void *const *out_host, int32_t chunks)
{
    if (!p || num_in < 0 || num_out < 0 || (num_in && (!in_cols || !in_host)) || (num_out && (!out_cols || !out_host)))
        return MDR_EINVAL;
    for (int i = 0; i < num_in; i++) if (!colSlot(p->params.t, in_cols[i]) || !in_host[i]) return MDR_EINVAL;
    for (int i = 0; i < num_out; i++) if (!colSlot(p->params.t, out_cols[i]) || !out_host[i]) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    closeTiming(p);
    const int W = p->cfg.num_worlds;
    const int wave = p->blocks * p->worlds_per_block;
    const int waves = (W + wave - 1) / wave;
    if (chunks < 1) chunks = 1;
    if (chunks > waves) chunks = waves;
    // World ranges in whole resident waves. What the pipeline cannot hide is the copy-in of the FIRST range and the
    // copy-out of the LAST one, so with five ranges or more the ends are short (1 and 2 waves) and the middle carries
    // the rest: 14 waves in 7 ranges = 1 2 3 3 2 2 1 ... the exposed copies shrink from a quarter of the batch to 1/14.
    std::vector<int> range_waves;
    if (chunks >= 5 && waves >= chunks + 2) {
        const int mid = chunks - 4, rest = waves - 6;
        range_waves = { 1, 2 };
        for (int k = 0; k < mid; k++) range_waves.push_back(rest / mid + (k < rest % mid ? 1 : 0));
        range_waves.push_back(2);
        range_waves.push_back(1);
    } else {
        const int per_waves = (waves + chunks - 1) / chunks;
        for (int left = waves; left > 0; left -= per_waves) range_waves.push_back(left < per_waves ? left : per_waves);
    }
    const int ranges = (int)range_waves.size();
    std::vector<int> range_first(ranges + 1, 0);
    for (int r = 0; r < ranges; r++) {
        const long long nxt = (long long)range_first[r] + (long long)range_waves[r] * wave;
        range_first[r + 1] = nxt < W ? (int)nxt : W;
    }
    range_first[ranges] = W;
    if (!p->io_in) {
        REQ_CUDA(cudaStreamCreateWithFlags(&p->io_in, cudaStreamNonBlocking));
        REQ_CUDA(cudaStreamCreateWithFlags(&p->io_out, cudaStreamNonBlocking));
        p->lane_stream[0] = p->stream;
        for (int l = 1; l < p->io_lanes; l++) REQ_CUDA(cudaStreamCreateWithFlags(&p->lane_stream[l], cudaStreamNonBlocking));
    }
    while ((int)p->io_events.size() < 2 * ranges + 4 + kMaxLanes) {
        cudaEvent_t e;
        REQ_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        p->io_events.push_back(e);
    }
    // A flagged BVH rebuild reads poses of EVERY world: on such a step (a reset) all inputs go in first and
    // the rebuild follows them; the per-range overlap of copy-in and step is given up for that one step.
    const bool inputs_first = p->maybe_rebuild;
    REQ_CUDA(cudaEventRecord(p->ev0, p->stream));
    // MDR_PHYS_IO_TRACE=1: device timestamps of every range -- copy-in landed, kernel
    // done, copy-out done -- printed by the next mdr_phys_sync (profiles/io_probe.py)
    const bool trace = p->io_trace_on;
    if (trace) {
        while ((int)p->io_trace.size() < 3 * ranges + 1) {
            cudaEvent_t e;
            REQ_CUDA(cudaEventCreate(&e));
            p->io_trace.push_back(e);
        }
        p->io_trace_ranges = ranges;
        REQ_CUDA(cudaEventRecord(p->io_trace[0], p->stream));
    }
    // the side streams start after everything already queued on the executor's stream
    cudaEvent_t start = p->io_events[2 * ranges], done = p->io_events[2 * ranges + 1];
    REQ_CUDA(cudaEventRecord(start, p->stream));
    REQ_CUDA(cudaStreamWaitEvent(p->io_in, start, 0));
    REQ_CUDA(cudaStreamWaitEvent(p->io_out, start, 0));
    for (int l = 1; l < p->io_lanes; l++) REQ_CUDA(cudaStreamWaitEvent(p->lane_stream[l], start, 0));
    const size_t stride = p->params.t.stride;
    if (inputs_first) {
        for (int i = 0; i < num_in; i++) {
            const size_t bytes = (size_t)kColBytes[in_cols[i]] * stride * W;
            REQ_CUDA(cudaMemcpyAsync(*colSlot(p->params.t, in_cols[i]), in_host[i], bytes, cudaMemcpyHostToDevice, p->io_in));
        }
        REQ_CUDA(cudaEventRecord(p->io_events[0], p->io_in));
        REQ_CUDA(cudaStreamWaitEvent(p->stream, p->io_events[0], 0));
        launchRebuild(p->params, W, p->stream);
        REQ_CUDA(cudaGetLastError());
        p->maybe_rebuild = false;
        REQ_CUDA(cudaEventRecord(p->io_events[2 * ranges + 2], p->stream));
        for (int l = 1; l < p->io_lanes; l++) REQ_CUDA(cudaStreamWaitEvent(p->lane_stream[l], p->io_events[2 * ranges + 2], 0));
    }
    // The ranges rotate over io_lanes streams (the executor's own first): a launch ends with blocks finishing within
    // one group time of each other, and the next ranges' blocks (other worlds, nothing shared but the read-only
    // tables) move into the freed SMs instead of waiting for the stragglers -- six launch tails were a fifth of the
    // end-to-end step. Each lane has its own task counter and overflow contact slots (lane, block_base).
    for (int r = 0; r < ranges; r++) {
        const int first = range_first[r], end = range_first[r + 1];
        if (end <= first) continue;
        for (int i = 0; i < num_in && !inputs_first; i++) {
            const size_t row = (size_t)kColBytes[in_cols[i]] * stride;
            REQ_CUDA(cudaMemcpyAsync((char *)*colSlot(p->params.t, in_cols[i]) + row * first,
                                     (const char *)in_host[i] + row * first, row * (end - first),
                                     cudaMemcpyHostToDevice, p->io_in));
        }
        const int lane = r % p->io_lanes;
        cudaStream_t ks = p->lane_stream[lane];
        REQ_CUDA(cudaEventRecord(p->io_events[2 * r], p->io_in));
        REQ_CUDA(cudaStreamWaitEvent(ks, p->io_events[2 * r], 0));
        long long grid = ((long long)(end - first) + p->worlds_per_block - 1) / p->worlds_per_block;
        if (grid > p->blocks) grid = p->blocks;
        p->kernel<<<(int)grid, p->threads, p->smem_bytes, ks>>>(
            p->params, p->threads_per_world, (int)p->slot_bytes, p->overflow_per_slot, p->lockstep, first, end, 1, 0u,
            lane * p->blocks, lane);
        REQ_CUDA(cudaGetLastError());
        REQ_CUDA(cudaEventRecord(p->io_events[2 * r + 1], ks));
        if (trace) REQ_CUDA(cudaEventRecord(p->io_trace[1 + 3 * r], p->io_in));
        if (trace) REQ_CUDA(cudaEventRecord(p->io_trace[2 + 3 * r], ks));
        REQ_CUDA(cudaStreamWaitEvent(p->io_out, p->io_events[2 * r + 1], 0));
        for (int i = 0; i < num_out; i++) {
            const size_t row = (size_t)kColBytes[out_cols[i]] * stride;
            REQ_CUDA(cudaMemcpyAsync((char *)out_host[i] + row * first,
                                     (const char *)*colSlot(p->params.t, out_cols[i]) + row * first,
                                     row * (end - first), cudaMemcpyDeviceToHost, p->io_out));
        }
        if (trace) REQ_CUDA(cudaEventRecord(p->io_trace[3 + 3 * r], p->io_out));
    }
    // mdr_phys_sync (the executor's stream) also covers the last copy-out; the copy-in stream joins through the
    // kernels that waited for it, the second kernel stream through the copy-outs that waited for its kernels
    REQ_CUDA(cudaEventRecord(done, p->io_out));
    REQ_CUDA(cudaStreamWaitEvent(p->stream, done, 0));
    REQ_CUDA(cudaEventRecord(p->io_events[2 * ranges + 3], p->io_in));       // every side stream joins the executor's
    REQ_CUDA(cudaStreamWaitEvent(p->stream, p->io_events[2 * ranges + 3], 0));
    for (int l = 1; l < p->io_lanes; l++) {
        REQ_CUDA(cudaEventRecord(p->io_events[2 * ranges + 3 + l], p->lane_stream[l]));
        REQ_CUDA(cudaStreamWaitEvent(p->stream, p->io_events[2 * ranges + 3 + l], 0));
    }
    REQ_CUDA(cudaEventRecord(p->ev1, p->stream));
    p->timing_open = true;
    p->pending_launches = ranges;
    return MDR_OK;
}

int32_t mdr_phys_sync(mdr_phys *p)
{
    if (!p) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    closeTiming(p);
    if (p->io_trace_ranges > 0) {
        fprintf(stderr, "[mdr io trace] range: copy-in landed / kernel done / copy-out done (ms after the call)\n");
        for (int r = 0; r < p->io_trace_ranges; r++) {
            float t[3] = { 0.f, 0.f, 0.f };
            for (int k = 0; k < 3; k++) cudaEventElapsedTime(&t[k], p->io_trace[0], p->io_trace[1 + 3 * r + k]);
            fprintf(stderr, "[mdr io trace] %2d: %.3f / %.3f / %.3f\n", r, t[0], t[1], t[2]);
        }
        cudaGetLastError();
        p->io_trace_ranges = 0;
    }
    int32_t err = 0;
    err = p->h_err ? *p->h_err : 0;          // the stream is idle: every atomicOr of the step has landed
    if (err) *p->h_err = 0;
    if (err & MDR_ERR_STALL) return MDR_ESTATE;            // a batch task gave up waiting for its world's previous step
    if (err & MDR_ERR_BAD_OBJECT) return MDR_EINVAL;       // an ObjectID outside the object table reached the step
    if (err & MDR_ERR_PAIR_OVERFLOW) return MDR_EOVERFLOW;
    return MDR_OK;
}

int32_t mdr_phys_step(mdr_phys *p, int32_t num_steps)
{
    int32_t rc = mdr_phys_step_async(p, num_steps);
    if (rc != MDR_OK) return rc;
    return mdr_phys_sync(p);
}

void *mdr_phys_stream(mdr_phys *p) { return p ? (void *)p->stream : nullptr; }

int32_t mdr_phys_pack_instances(mdr_phys *p, const float *world_offset, float *out_transforms,
                                int32_t *out_object_ids)
{
    if (!p || !out_transforms || !out_object_ids) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    FeederJob job;
    memset(&job, 0, sizeof(job));
    for (int k = 0; k < 3; k++) job.world_offset[k] = world_offset ? world_offset[k] : 0.f;
    job.transforms = out_transforms;
    job.object_ids = out_object_ids;
    const size_t total = (size_t)p->cfg.num_worlds * p->params.t.stride;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    physFeedInstancesKernel<<<blocks, 256, 0, p->stream>>>(p->params.t, job);
    REQ_CUDA(cudaGetLastError());
    return MDR_OK;
}

int32_t mdr_phys_pack_views(mdr_phys *p, int32_t views_per_world, const int32_t *view_rows,
                            const float *camera_offset, const float *y_scale, const float *world_offset,
                            float *out_views)
{
    if (!p || views_per_world < 0 || (views_per_world && (!view_rows || !camera_offset || !y_scale || !out_views)))
        return MDR_EINVAL;
    if (views_per_world == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    FeederJob job;
    memset(&job, 0, sizeof(job));
    for (int k = 0; k < 3; k++) job.world_offset[k] = world_offset ? world_offset[k] : 0.f;
    job.views_per_world = views_per_world;
    job.view_rows = view_rows;
    job.camera_offset = camera_offset;
    job.y_scale = y_scale;
    job.views = out_views;
    const size_t total = (size_t)p->cfg.num_worlds * views_per_world;
    const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    physFeedViewsKernel<<<blocks, 256, 0, p->stream>>>(p->params.t, job);
    REQ_CUDA(cudaGetLastError());
    return MDR_OK;
}

int32_t mdr_phys_trace_rays_device(mdr_phys *p, int32_t rays_per_world, const float *origins, const float *dirs,
                                   const float *t_max, float *out_t, float *out_normal, int32_t *out_entity,
                                   int32_t *out_row)
{
    if (!p || rays_per_world < 0 || (rays_per_world && (!origins || !dirs || !out_t || !out_normal)))
        return MDR_EINVAL;
    if (rays_per_world == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    RayJob job;
    job.origins = origins; job.dirs = dirs; job.t_max = t_max;
    job.out_t = out_t; job.out_normal = out_normal; job.out_entity = out_entity; job.out_row = out_row;
    job.rays_per_world = rays_per_world;
    const size_t smem = sizeof(RayLeaf) * (size_t)p->params.t.stride;
    if (smem > 48 * 1024)
        REQ_CUDA(cudaFuncSetAttribute(physTraceRaysKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int threads = rays_per_world >= 128 ? 128 : (rays_per_world + 31) / 32 * 32;
    physTraceRaysKernel<<<p->cfg.num_worlds, threads < 32 ? 32 : threads, smem, p->stream>>>(p->params.t, p->params.a, job);
    REQ_CUDA(cudaGetLastError());
    return MDR_OK;
}

int32_t mdr_phys_trace_rays(mdr_phys *p, int32_t rays_per_world, const float *origins, const float *dirs,
                            const float *t_max, float *out_t, float *out_normal, int32_t *out_entity,
                            int32_t *out_row)
{
    if (!p || rays_per_world < 0 || (rays_per_world && (!origins || !dirs || !out_t || !out_normal)))
        return MDR_EINVAL;
    if (rays_per_world == 0) return MDR_OK;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const size_t n = (size_t)p->cfg.num_worlds * rays_per_world;
    // origins 3, dirs 3, t_max 1, t 1, normal 3, entity 2, row 1 words per ray
    const size_t need = n * 14 * 4;
    if (need > p->ray_buf_bytes) {
        REQ_CUDA(cudaStreamSynchronize(p->stream));
        if (p->ray_buf) REQ_CUDA(cudaFree(p->ray_buf));
        REQ_CUDA(cudaMalloc(&p->ray_buf, need));
        p->ray_buf_bytes = need;
    }
    float *d_o = (float *)p->ray_buf, *d_d = d_o + 3 * n, *d_tm = d_d + 3 * n, *d_t = d_tm + n, *d_n = d_t + n;
    int32_t *d_e = (int32_t *)(d_n + 3 * n), *d_r = d_e + 2 * n;
    REQ_CUDA(cudaMemcpyAsync(d_o, origins, 12 * n, cudaMemcpyHostToDevice, p->stream));
    REQ_CUDA(cudaMemcpyAsync(d_d, dirs, 12 * n, cudaMemcpyHostToDevice, p->stream));
    if (t_max) REQ_CUDA(cudaMemcpyAsync(d_tm, t_max, 4 * n, cudaMemcpyHostToDevice, p->stream));
    int32_t rc = mdr_phys_trace_rays_device(p, rays_per_world, d_o, d_d, t_max ? d_tm : nullptr, d_t, d_n, d_e, d_r);
    if (rc != MDR_OK) return rc;
    REQ_CUDA(cudaMemcpyAsync(out_t, d_t, 4 * n, cudaMemcpyDeviceToHost, p->stream));
    REQ_CUDA(cudaMemcpyAsync(out_normal, d_n, 12 * n, cudaMemcpyDeviceToHost, p->stream));
    if (out_entity) REQ_CUDA(cudaMemcpyAsync(out_entity, d_e, 8 * n, cudaMemcpyDeviceToHost, p->stream));
    if (out_row) REQ_CUDA(cudaMemcpyAsync(out_row, d_r, 4 * n, cudaMemcpyDeviceToHost, p->stream));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    return MDR_OK;
}

int32_t mdr_phys_probe_step(mdr_phys *p, int32_t world, int32_t max_pairs, int32_t *num_pairs,
                            int32_t *pairs, int32_t *num_contacts, mdr_contact *contacts)
{
    static_assert(sizeof(mdr_contact) == sizeof(ContactDev), "contact layout");
    if (!p || world < 0 || world >= p->cfg.num_worlds || max_pairs <= 0 || !num_pairs || !pairs ||
        !num_contacts || !contacts)
        return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    const int S = p->cfg.num_substeps, MC = p->cfg.max_contacts;
    ProbeDev &d = p->d_probe;
    if (!d.num_pairs || p->probe_pairs_cap < max_pairs) {
        d.num_pairs = p->dalloc<int32_t>(1);
        d.pairs = p->dalloc<int32_t>(2 * (size_t)max_pairs);
        d.num_contacts = p->dalloc<int32_t>(S);
        d.contacts = p->dalloc<ContactDev>((size_t)S * MC);
        p->probe_pairs_cap = max_pairs;
    }
    d.world = world; d.max_pairs = max_pairs; d.max_contacts = MC;
    REQ_CUDA(cudaMemset(d.contacts, 0, sizeof(ContactDev) * (size_t)S * MC));
    REQ_CUDA(cudaMemset(d.num_pairs, 0, 4));
    p->params.probe = d;
    int32_t rc = mdr_phys_step(p, 1);
    p->params.probe.world = -1;
    REQ_CUDA(cudaMemcpy(num_pairs, d.num_pairs, 4, cudaMemcpyDeviceToHost));
    REQ_CUDA(cudaMemcpy(pairs, d.pairs, 8 * (size_t)max_pairs, cudaMemcpyDeviceToHost));
    REQ_CUDA(cudaMemcpy(contacts, d.contacts, sizeof(ContactDev) * (size_t)S * MC, cudaMemcpyDeviceToHost));
    compactProbeContacts((ContactDev *)contacts, num_contacts, S, MC);
    return rc;
}

int32_t mdr_phys_stats(mdr_phys *p, int32_t *pairs, int32_t *contacts)
{
    if (!p) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    const size_t b = 4 * (size_t)p->cfg.num_worlds;
    if (pairs) REQ_CUDA(cudaMemcpy(pairs, p->params.world_pairs, b, cudaMemcpyDeviceToHost));
    if (contacts) REQ_CUDA(cudaMemcpy(contacts, p->params.world_contacts, b, cudaMemcpyDeviceToHost));
    return MDR_OK;
}

int32_t mdr_phys_read_bvh(mdr_phys *p, float *slot, float *leaf, uint16_t *order)
{
    if (!p) return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(p->cfg.gpu_id));
    REQ_CUDA(cudaStreamSynchronize(p->stream));
    const size_t n = p->cfg.max_bodies_per_world, st = p->params.t.stride, W = p->cfg.num_worlds;
    if (slot) REQ_CUDA(cudaMemcpy2D(slot, n * 24, p->params.t.slot_box, st * 24, n * 24, W, cudaMemcpyDeviceToHost));
    if (leaf) REQ_CUDA(cudaMemcpy2D(leaf, n * 24, p->params.t.leaf_aabb, st * 24, n * 24, W, cudaMemcpyDeviceToHost));
    if (order) REQ_CUDA(cudaMemcpy2D(order, n * 2, p->params.t.order, st * 2, n * 2, W, cudaMemcpyDeviceToHost));
    return MDR_OK;
}

int32_t mdr_phys_get_hull(mdr_phys *p, int32_t obj, uint32_t *half_edges, uint32_t *edges,
                          uint32_t *polys, float *planes, int32_t *counts)
{
    if (!p || !counts || obj < 0 || obj >= (int)p->assets.objs.size()) return MDR_EINVAL;
    const ObjDev &o = p->assets.objs[obj];
    counts[0] = o.num_hedges; counts[1] = o.num_edges; counts[2] = o.num_faces; counts[3] = o.num_verts;
    if (o.type != MDR_PRIM_HULL) return MDR_OK;
    if (half_edges) memcpy(half_edges, p->assets.hedges.data() + o.hedge_off, 16 * (size_t)o.num_hedges);
    if (edges) memcpy(edges, p->assets.edges.data() + o.edge_off, 4 * (size_t)o.num_edges);
    if (polys) memcpy(polys, p->assets.polys.data() + o.face_off, 4 * (size_t)o.num_faces);
    if (planes) memcpy(planes, p->assets.planes.data() + 4 * (size_t)o.face_off, 16 * (size_t)o.num_faces);
    return MDR_OK;
}

static int32_t handOverHull(const ObjHull &h, bool ok, float *verts, int32_t *num_verts, uint32_t *poly_words,
                            int32_t *num_poly_words, float aabb[6], char *err, int32_t err_cap)
{
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, "%s", h.error.c_str());
    if (!ok) return MDR_EINVAL;
    const int32_t nv = (int32_t)(h.verts.size() / 3), nw = (int32_t)h.words.size();
    const bool query = !verts || !poly_words;
    const bool fits = query || (*num_verts >= nv && *num_poly_words >= nw);
    *num_verts = nv;
    *num_poly_words = nw;
    if (aabb) memcpy(aabb, h.aabb, sizeof(h.aabb));
    if (query) return MDR_OK;
    if (!fits) return MDR_EOVERFLOW;
    memcpy(verts, h.verts.data(), 12 * (size_t)nv);
    memcpy(poly_words, h.words.data(), 4 * (size_t)nw);
    return MDR_OK;
}

int32_t mdr_load_hull_obj(const char *obj_text, int64_t num_bytes, float *verts, int32_t *num_verts,
                          uint32_t *poly_words, int32_t *num_poly_words, float aabb[6], char *err, int32_t err_cap)
{
    if (!obj_text || num_bytes < 0 || !num_verts || !num_poly_words) return MDR_EINVAL;
    ObjHull h;
    const bool ok = loadObjHull(obj_text, (size_t)num_bytes, h);
    return handOverHull(h, ok, verts, num_verts, poly_words, num_poly_words, aabb, err, err_cap);
}

int32_t mdr_load_hull_obj_file(const char *path, float *verts, int32_t *num_verts, uint32_t *poly_words,
                               int32_t *num_poly_words, float aabb[6], char *err, int32_t err_cap)
{
    if (!path || !num_verts || !num_poly_words) return MDR_EINVAL;
    ObjHull h;
    std::string text;
    bool ok = false;
    const char *dot = strrchr(path, '.');
    if (!dot || strcmp(dot + 1, "obj") != 0) {
        h.error = "not an .obj path (importObject reads nothing else, importer.cpp:423-434)";
    } else if (FILE *f = fopen(path, "rb")) {
        char buf[1 << 16];
        size_t n;
        while ((n = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, n);
        fclose(f);
        ok = loadObjHull(text.data(), text.size(), h);
    } else {
        h.error = "cannot open the file";
    }
    return handOverHull(h, ok, verts, num_verts, poly_words, num_poly_words, aabb, err, err_cap);
}

int32_t mdr_phys_launch_info(mdr_phys *p, int32_t info[6])
{
    if (!p || !info) return MDR_EINVAL;
    info[0] = p->blocks; info[1] = p->threads; info[2] = (int32_t)p->smem_bytes; info[3] = p->blocks_per_sm;
    info[4] = p->threads_per_world; info[5] = p->worlds_per_block;
    return MDR_OK;
}

int32_t mdr_phys_last_launches(mdr_phys *p) { return p ? p->last_call_launches : MDR_EINVAL; }

float mdr_phys_kernel_ms(mdr_phys *p, int32_t *launches)
{
    if (!p) return -1.f;
    closeTiming(p);
    int n = p->timed_launches;
    double ms = p->timed_ms;
    p->timed_launches = 0;
    p->timed_ms = 0.0;
    if (launches) *launches = n;
    return n > 0 ? (float)(ms / n) : -1.f;
}

}  // extern "C"
