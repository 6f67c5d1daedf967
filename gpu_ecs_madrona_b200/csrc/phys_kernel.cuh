// The persistent step kernel (one block steps `slots` worlds side by side, worlds strided over
// the grid) and the table of its compiled register budgets. Instantiated by phys_inst*.cu, a few
// variants per translation unit so that the units build in parallel; physics.cu picks one.
#pragma once
#include <cuda_runtime.h>

#include "phys_step.cuh"

namespace mdr {


// kHullSlots: 1 = hulls of up to 32 vertices / faces, 2 = up to 64 (phys_sat_warp.cuh); the big-hull variant carries
// the joint solve as well (one extra instantiation per register budget instead of two).
template <int kMaxThreads, int kMinBlocks, bool kJoints, int kHullSlots = 1, typename Layout = RuntimeLayout>
__global__ void __launch_bounds__(kMaxThreads, kMinBlocks)
physStepKernel(const __grid_constant__ StepParams p, int threads_per_world, int slot_bytes,
               int overflow_per_slot, int lockstep, int first_world, int end_world, int steps, uint32_t step_base,
               int block_base, int lane)
{
    __shared__ int group_level;
    __shared__ BlockSolve block_solve;
    constexpr bool kFixed = Layout::kNp > 0;
    if (kFixed) {                 // the host launches this variant only with these values
        threads_per_world = Layout::kTpw;
        lockstep = 1;
    }
    const int slots = blockDim.x / threads_per_world;
    const int slot = threadIdx.x / threads_per_world;
    BlockCtx ctx;
    ctx.tid_ = threadIdx.x - slot * threads_per_world;
    ctx.nth_ = threads_per_world;
    ctx.bar_ = lockstep ? 0 : slot + 1;
    WorldView v;
    if (kFixed) {
        carveWorld(v, mdr_smem + (size_t)slot * carveWorld(v, nullptr, Layout::kNp, Layout::kPairs, Layout::kContacts),
                   Layout::kNp, Layout::kPairs, Layout::kContacts);
    } else {
        carveWorld(v, mdr_smem + (size_t)slot * slot_bytes, p.t.stride, p.max_pairs, p.smem_contacts);
    }
    // lane, block_base = lane x grid capacity: a launch that may run BESIDE others of this executor (the world ranges
    // of mdr_phys_step_io_async rotate over a few streams) has its own overflow slots and task counter
    v.overflow = p.overflow + ((size_t)(block_base + blockIdx.x) * slots + slot) * overflow_per_slot;
    // dense solver lanes need block-wide barriers, pair slots in 10 bits and a view per slot
    BlockSolve *bs = kFixed ? &block_solve
                            : (lockstep && slots <= kDenseSlots && p.max_pairs <= 1024 ? &block_solve : nullptr);
    v.level_any = bs ? block_solve.level_any : nullptr;
    if (bs && ctx.tid_ == 0) { v.n = 0; v.active = false; bs->views[slot] = v; }
    __syncthreads();
    // Every slot of the block walks the same barriers (the task is block-uniform); a slot past the last world idles
    // through them.
    // TASKS = (step, world group), step-major, handed out by a counter instead of strided over the grid: a block that
    // drew light worlds (or shares its SM with a block that did) takes more of them, so the launch ends when the work
    // ends, not when the unluckiest block does: +5.5 % on collisions-64 (profiles/r02_summary.md section 8). The
    // first task of a block is its own index; the next one is drawn while the current one is stepped; the last block
    // out resets the counter for the next launch (launches of one executor are ordered on a stream: they share the
    // tables anyway). Which block steps a world does not change the world's result.
    // steps > 1 (mdr_phys_step_async(p, n > 1), a batch nobody can observe in between): step s of a group waits for
    // the group's step s - 1, published in group_done by the block that ran it; the task before it by one whole step
    // of all groups, so it has long finished and the wait is one load. The idle tail of a launch (blocks finishing
    // within one group time of each other) is then paid once per batch instead of once per step. Deadlock-free:
    // tasks are drawn in increasing order by resident blocks, and a task waits only for a smaller one.
    uint32_t *const work = p.work + 2 * lane;
    const int groups = (end_world - first_world + slots - 1) / slots;
    const uint32_t total = (uint32_t)groups * (uint32_t)steps;
    uint32_t task = blockIdx.x;
    while (task < total) {
        if (threadIdx.x == 0) block_solve.next_task = gridDim.x + atomicAdd(work, 1u);
        const uint32_t s = steps > 1 ? task / (uint32_t)groups : 0u;
        const int g = (int)(task - s * (uint32_t)groups);
        if (s > 0u && threadIdx.x == 0) {
            const volatile uint32_t *done = p.group_done + g;
            const uint32_t want = step_base + s;
            uint32_t spins = 0u;
            while (*done != want) {
                __nanosleep(100);
                if (++spins > (1u << 22)) { atomicOr(p.error_flags, MDR_ERR_STALL); break; }   // never hang the device
            }
            __threadfence();
        }
        const int w = first_world + g * slots + slot;
        v.active = w < end_world;
        stepWorld<kJoints, kHullSlots, Layout>(ctx, v, p, v.active ? w : first_world, &group_level, bs, slot);
        if (steps > 1) __threadfence();          // this thread's table writes, before the group is published
        __syncthreads();
        if (steps > 1 && threadIdx.x == 0) *(volatile uint32_t *)(p.group_done + g) = step_base + s + 1u;
        task = block_solve.next_task;
        __syncthreads();
    }
    if (threadIdx.x == 0 && atomicAdd(work + 1, 1u) == gridDim.x - 1) {
        atomicExch(work, 0u);
        atomicExch(work + 1, 0u);
    }
}


typedef void (*StepKernel)(const StepParams, int, int, int, int, int, int, int, uint32_t, int, int);
// {max block threads, min resident blocks}; fn_joints is the same kernel with the joint
// solve compiled in (mdr_phys_set_joints), so that batches without joints do not carry it.
struct KernelVariant {
    int threads, min_blocks;
    StepKernel fn, fn_joints, fn_big;
    StepKernel fn_fixed;                                   // plain kernel with a compile-time layout, or null
    int fixed_np, fixed_pairs, fixed_contacts, fixed_tpw;  // the shape fn_fixed was compiled for
};
#define MDR_STEP_VARIANT(T, B) { T, B, mdr::physStepKernel<T, B, false>, mdr::physStepKernel<T, B, true>, \
                                 mdr::physStepKernel<T, B, true, 2>, nullptr, 0, 0, 0, 0 }
// the same plus the collisions-64 shape as a fixed-layout instantiation
#define MDR_STEP_VARIANT_C64(T, B) { T, B, mdr::physStepKernel<T, B, false>, mdr::physStepKernel<T, B, true>, \
                                     mdr::physStepKernel<T, B, true, 2>, \
                                     mdr::physStepKernel<T, B, false, 1, mdr::Collisions64Layout>, \
                                     mdr::Collisions64Layout::kNp, mdr::Collisions64Layout::kPairs, \
                                     mdr::Collisions64Layout::kContacts, mdr::Collisions64Layout::kTpw }

}  // namespace mdr

const mdr::KernelVariant *mdrStepVariants0(int *n);
const mdr::KernelVariant *mdrStepVariants1(int *n);
const mdr::KernelVariant *mdrStepVariants2(int *n);
const mdr::KernelVariant *mdrStepVariants3(int *n);
const mdr::KernelVariant *mdrStepVariants4(int *n);
