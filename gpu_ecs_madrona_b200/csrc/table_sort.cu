// Global sort / compaction of a shared (not per-world) archetype table by a u32 key
// column: the reference's SortArchetypeNodeBase graph
// (src/mw/device/sort_archetype.cpp:960-1392: sortSetup, zeroBins, histogram, binScan,
// prepareOnesweep + onesweep per 8-bit digit, resizeTable, copyKeys, stageColumn +
// rearrangeColumn per column, rearrangeEntities) as five hand-written sm_100a kernels
// behind the mdr_sort_* C ABI.
//
// What it computes (the contract the parity tests hold it to): a STABLE sort of the rows
// by the low 8*passes bits of the key, passes = ceil(bits(num_worlds + 1) / 8) for a
// WorldID sort (sort_archetype.cpp:1290-1299) else 4; every column permuted the same way;
// for a WorldID sort the table is cut to the rows whose top digit is below 255
// (resizeTable, :1216-1221: rows of destroyed entities carry WorldID -1 and sort last);
// the entity -> row map is rewritten for the surviving rows (rearrangeEntities, :1243-1263).
//
// B200 shape: all traffic is streaming and HBM-bound (no tensor cores). One histogram
// pass builds every digit's bins at once; one single-pass kernel per digit ranks a 4096-key
// tile with warp match-any (stable), chains tiles with decoupled look-back, reorders the
// tile in shared memory and writes whole digit runs (coalesced); ONE gather kernel moves
// every column (the reference stages and copies back each column in two node pairs).
// Columns are gathered from `columns_in` to `columns_out`: callers that double-buffer
// their tables pay one read and one write per byte; a column whose in and out pointers
// are the same goes through a staging buffer like the reference does.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/madrona_b200.h"

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

constexpr int kDigits = 256;          // RADIX_BITS = 8 (sort_archetype.cpp sortConsts)
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 16;            // keys per thread in a tile
constexpr int kTile = kThreads * kItems;
constexpr int kMaxPasses = 4;
constexpr uint32_t kFlagAggregate = 1u << 30, kFlagInclusive = 2u << 30, kValueMask = (1u << 30) - 1;

// ---- 1. all digit histograms in one pass over the keys -----------------------
__global__ void __launch_bounds__(kThreads)
sortHistogramKernel(const uint32_t *__restrict__ keys, int n, int passes, uint32_t *__restrict__ bins)
{
    __shared__ uint32_t s_bins[kMaxPasses][kDigits];
    for (int p = 0; p < passes; p++) s_bins[p][threadIdx.x] = 0;
    __syncthreads();
    const int n4 = n / 4;
    const uint4 *keys4 = reinterpret_cast<const uint4 *>(keys);
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < n4; i += gridDim.x * kThreads) {
        const uint4 k = __ldg(keys4 + i);
        const uint32_t ks[4] = { k.x, k.y, k.z, k.w };
#pragma unroll
        for (int j = 0; j < 4; j++)
            for (int p = 0; p < passes; p++) atomicAdd(&s_bins[p][(ks[j] >> (8 * p)) & 255u], 1u);
    }
    if (blockIdx.x == 0 && threadIdx.x < n - 4 * n4) {
        const uint32_t k = keys[4 * n4 + threadIdx.x];
        for (int p = 0; p < passes; p++) atomicAdd(&s_bins[p][(k >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int p = 0; p < passes; p++) {
        const uint32_t c = s_bins[p][threadIdx.x];
        if (c) atomicAdd(&bins[p * kDigits + threadIdx.x], c);
    }
}

// Exclusive scan of 256 values held one per thread of a 256-thread block.
__device__ __forceinline__ uint32_t blockExclusiveScan(uint32_t x, uint32_t *s_warp /*[kWarps]*/, uint32_t *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) {
        const uint32_t s = s_warp[w];
        if (w < warp) before += s;
        all += s;
    }
    if (total) *total = all;
    __syncthreads();
    return before + incl - x;
}

// ---- 2. bins -> exclusive offsets, one block per digit position ---------------
// Also keeps the count of the last digit's value 255 out of the way: the caller reads the
// new row count as the offset of digit 255 of the last pass (resizeTable).
__global__ void __launch_bounds__(kThreads) sortScanKernel(uint32_t *bins)
{
    __shared__ uint32_t s_warp[kWarps];
    uint32_t *b = bins + blockIdx.x * kDigits;
    const uint32_t x = b[threadIdx.x];
    b[threadIdx.x] = blockExclusiveScan(x, s_warp, nullptr);
}

// ---- 3. one digit: rank, look back, reorder, scatter --------------------------
struct OnesweepSmem {
    uint32_t warp_cnt[kWarps][kDigits];
    uint32_t tile_off[kDigits];
    uint32_t gbase[kDigits];
    uint32_t keys[kTile];
    int32_t vals[kTile];
    uint32_t s_warp[kWarps];
    uint32_t tile;
};

__global__ void __launch_bounds__(kThreads, 4)
sortOnesweepKernel(const uint32_t *__restrict__ src_keys, const int32_t *__restrict__ src_vals,
                   uint32_t *__restrict__ dst_keys, int32_t *__restrict__ dst_vals,
                   const uint32_t *__restrict__ bins, uint32_t *lookback, uint32_t *tile_counter,
                   int n, int shift)
{
    extern __shared__ __align__(16) char smem_raw[];
    OnesweepSmem &s = *reinterpret_cast<OnesweepSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // Tiles are handed out in launch order, so a tile only ever waits for tiles whose blocks
    // are already running (forward progress of the look-back).
    if (tid == 0) s.tile = atomicAdd(tile_counter, 1u);
#pragma unroll
    for (int w = 0; w < kWarps; w++) s.warp_cnt[w][tid] = 0;
    __syncthreads();
    const uint32_t tile = s.tile;
    const int base = (int)tile * kTile;
    const int tile_n = min(kTile, n - base);

    // (a) stable rank inside the warp's 512 consecutive keys: match-any per round of 32.
    // All key loads are issued before the first rank (16 independent loads in flight).
    uint32_t key[kItems];
    uint16_t rank[kItems];
    const uint32_t lt = (1u << lane) - 1u;
    const int first = warp * (kItems * 32) + lane;
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int local = first + r * 32;
        key[r] = local < tile_n ? __ldg(src_keys + base + local) : 0xffffffffu;
    }
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const bool valid = first + r * 32 < tile_n;
        const uint32_t digit = (key[r] >> shift) & 255u;
        const uint32_t mask = __match_any_sync(0xffffffffu, valid ? digit : (0x100u | lane));
        uint32_t *cnt = &s.warp_cnt[warp][digit];
        const uint32_t before = valid ? *cnt : 0u;
        __syncwarp();
        const uint32_t mine = __popc(mask & lt);
        rank[r] = (uint16_t)(before + mine);
        if (valid && mine == 0) *cnt = before + __popc(mask);
        __syncwarp();
    }
    __syncthreads();

    // (b) thread d owns digit d: warp counts -> warp offsets, tile count
    uint32_t tile_count = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) {
        const uint32_t c = s.warp_cnt[w][tid];
        s.warp_cnt[w][tid] = tile_count;
        tile_count += c;
    }
    // (c) decoupled look-back over the tiles before this one, per digit
    volatile uint32_t *lb = lookback;
    uint32_t excl = 0;
    if (tile == 0) {
        lb[tid] = kFlagInclusive | tile_count;
    } else {
        lb[(size_t)tile * kDigits + tid] = kFlagAggregate | tile_count;
        for (int t = (int)tile - 1;; t--) {
            uint32_t v;
            do { v = lb[(size_t)t * kDigits + tid]; } while ((v >> 30) == 0u);
            excl += v & kValueMask;
            if ((v >> 30) == 2u) break;
        }
        lb[(size_t)tile * kDigits + tid] = kFlagInclusive | (excl + tile_count);
    }
    // (d) where each digit's run starts inside the tile and in the output
    const uint32_t off = blockExclusiveScan(tile_count, s.s_warp, nullptr);
    s.tile_off[tid] = off;
    s.gbase[tid] = __ldg(bins + tid) + excl - off;
    __syncthreads();

    // (e) reorder the tile in shared memory, then write digit runs with consecutive threads
#pragma unroll
    for (int r = 0; r < kItems; r++) {
        const int local = first + r * 32;
        if (local < tile_n) {
            const uint32_t digit = (key[r] >> shift) & 255u;
            const uint32_t pos = s.tile_off[digit] + s.warp_cnt[warp][digit] + rank[r];
            s.keys[pos] = key[r];
            s.vals[pos] = src_vals ? __ldg(src_vals + base + local) : base + local;   // pass 0: identity
        }
    }
    __syncthreads();
    for (int k = tid; k < tile_n; k += kThreads) {
        const uint32_t kk = s.keys[k];
        const uint32_t dst = s.gbase[(kk >> shift) & 255u] + (uint32_t)k;
        dst_keys[dst] = kk;
        dst_vals[dst] = s.vals[k];
    }
}

// ---- 4. every column through the permutation, one launch ----------------------
struct ColumnDesc {
    const char *src;
    char *dst;
    int32_t words;        // 4-byte words per row (column bytes are a multiple of 4)
    int32_t is_key;       // the key column is written from the sorted keys
};

// blockIdx.y = column. Thread = one 4-byte word of the output: consecutive threads write
// consecutive words (coalesced) and read runs of `words` consecutive words per source row.
// The row of a word is a division by the column's word count: a compile-time constant per
// instantiation (a variable 64-bit division cost more than the memory traffic), four
// independent words per thread and iteration in flight.
template <int kWords>
__device__ __forceinline__ void gatherColumn(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst,
                                             const int32_t *__restrict__ perm, uint32_t total, int words_rt)
{
    const uint32_t words = kWords > 0 ? (uint32_t)kWords : (uint32_t)words_rt;
    const uint32_t stride = gridDim.x * blockDim.x;
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    for (; j + 3u * stride < total; j += 4u * stride) {
        uint32_t v[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t jj = j + (uint32_t)u * stride;
            const uint32_t row = jj / words;
            v[u] = __ldg(src + (size_t)__ldg(perm + row) * words + (jj - row * words));
        }
#pragma unroll
        for (int u = 0; u < 4; u++) dst[j + (uint32_t)u * stride] = v[u];
    }
    for (; j < total; j += stride) {
        const uint32_t row = j / words;
        dst[j] = __ldg(src + (size_t)__ldg(perm + row) * words + (j - row * words));
    }
}

__global__ void __launch_bounds__(256)
tableGatherKernel(const ColumnDesc *__restrict__ cols, const int32_t *__restrict__ perm,
                  const uint32_t *__restrict__ sorted_keys, int rows)
{
    const ColumnDesc c = cols[blockIdx.y];
    const uint32_t total = (uint32_t)rows * (uint32_t)c.words;
    uint32_t *dst = reinterpret_cast<uint32_t *>(c.dst);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(c.src);
    if (c.is_key) {
        for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < total; j += gridDim.x * blockDim.x)
            dst[j] = __ldg(sorted_keys + j);
        return;
    }
    switch (c.words) {                   // block-uniform
    case 1: gatherColumn<1>(src, dst, perm, total, 1); break;
    case 2: gatherColumn<2>(src, dst, perm, total, 2); break;
    case 3: gatherColumn<3>(src, dst, perm, total, 3); break;
    case 4: gatherColumn<4>(src, dst, perm, total, 4); break;
    case 6: gatherColumn<6>(src, dst, perm, total, 6); break;
    case 7: gatherColumn<7>(src, dst, perm, total, 7); break;
    default: gatherColumn<0>(src, dst, perm, total, c.words); break;
    }
}

// ---- 5. entity -> row map of the rows that survive (rearrangeEntities) --------
__global__ void entityRemapKernel(const int32_t *__restrict__ entities /*[rows] (gen, id)*/, int rows,
                                  const uint32_t *__restrict__ cut, int32_t *__restrict__ entity_row, int32_t max_ids)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows || (cut && (uint32_t)i >= __ldg(cut))) return;
    const int32_t id = __ldg(entities + 2 * (int64_t)i + 1);
    if (id >= 0 && id < max_ids) entity_row[id] = i;
}

int passesFor(int num_worlds)
{
    if (num_worlds <= 0) return 4;                       // generic key: all 32 bits
    int bits = 0;
    for (uint32_t v = (uint32_t)num_worlds + 1u; v; v >>= 1) bits++;   // 32 - clz(num_worlds + 1)
    return (bits + 7) / 8;
}

}  // namespace

struct mdr_sort {
    int gpu_id = 0;
    int max_rows = 0, num_columns = 0, key_col = 0, passes = 0, world_sort = 0;
    std::vector<int32_t> col_bytes;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint32_t *keys_a = nullptr, *keys_b = nullptr;
    int32_t *vals_a = nullptr, *vals_b = nullptr;
    uint32_t *bins = nullptr, *lookback = nullptr, *counters = nullptr;
    ColumnDesc *d_cols = nullptr;
    std::vector<ColumnDesc> h_cols;
    char *staging = nullptr;
    size_t staging_bytes = 0;
    int sms = 0;
    float last_ms = 0.f;
    int last_launches = 0;
    uint32_t host_cut = 0;
};

extern "C" {

int32_t mdr_sort_num_passes(int32_t num_worlds) { return passesFor(num_worlds); }

int32_t mdr_sort_create(int32_t gpu_id, int32_t max_rows, int32_t num_columns, const int32_t *column_bytes,
                        int32_t key_column, int32_t num_worlds, mdr_sort **out)
{
    if (!out || max_rows <= 0 || num_columns <= 0 || num_columns > 128 || !column_bytes ||
        key_column < 0 || key_column >= num_columns || column_bytes[key_column] != 4)
        return MDR_EINVAL;
    size_t max_col = 0;
    for (int c = 0; c < num_columns; c++) {
        if (column_bytes[c] <= 0 || column_bytes[c] % 4) return MDR_EINVAL;
        if ((size_t)column_bytes[c] > max_col) max_col = column_bytes[c];
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || gpu_id < 0 || gpu_id >= ndev) return MDR_ENODEVICE;
    REQ_CUDA(cudaSetDevice(gpu_id));
    mdr_sort *s = new mdr_sort;
    s->gpu_id = gpu_id;
    s->max_rows = max_rows;
    s->num_columns = num_columns;
    s->key_col = key_column;
    s->passes = passesFor(num_worlds);
    s->world_sort = num_worlds > 0;
    s->col_bytes.assign(column_bytes, column_bytes + num_columns);
    cudaDeviceProp prop;
    REQ_CUDA(cudaGetDeviceProperties(&prop, gpu_id));
    s->sms = prop.multiProcessorCount;
    const size_t tiles = ((size_t)max_rows + kTile - 1) / kTile;
    // bins | tile counters | look-back words in one block: one memset per run
    bool ok = true;
    auto take = [&](void **ptr, size_t bytes) {
        if (ok && cudaMalloc(ptr, bytes) != cudaSuccess) { cudaGetLastError(); ok = false; }
    };
    take((void **)&s->keys_a, 4 * (size_t)max_rows);
    take((void **)&s->keys_b, 4 * (size_t)max_rows);
    take((void **)&s->vals_a, 4 * (size_t)max_rows);
    take((void **)&s->vals_b, 4 * (size_t)max_rows);
    take((void **)&s->bins, 4 * ((size_t)kMaxPasses * kDigits + kDigits + tiles * kDigits * kMaxPasses));
    take((void **)&s->d_cols, sizeof(ColumnDesc) * (size_t)num_columns);
    if (!ok) {      // the scratch of a table this size does not fit the device
        mdr_sort_destroy(s);
        return MDR_ECAPACITY;
    }
    s->counters = s->bins + kMaxPasses * kDigits;
    s->lookback = s->counters + kDigits;
    s->staging_bytes = max_col * (size_t)max_rows;
    REQ_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    REQ_CUDA(cudaEventCreate(&s->ev0));
    REQ_CUDA(cudaEventCreate(&s->ev1));
    REQ_CUDA(cudaFuncSetAttribute(sortOnesweepKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(OnesweepSmem)));
    *out = s;
    return MDR_OK;
}

void mdr_sort_destroy(mdr_sort *s)
{
    if (!s) return;
    cudaSetDevice(s->gpu_id);
    if (s->stream) { cudaStreamSynchronize(s->stream); cudaStreamDestroy(s->stream); }
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    for (void *p : { (void *)s->keys_a, (void *)s->keys_b, (void *)s->vals_a, (void *)s->vals_b, (void *)s->bins,
                     (void *)s->d_cols, (void *)s->staging })
        if (p) cudaFree(p);
    delete s;
}

int32_t mdr_sort_run(mdr_sort *s, const void *const *columns_in, void *const *columns_out, int32_t num_rows,
                     int32_t *entity_row, int32_t max_entity_ids, int32_t *out_num_rows)
{
    if (!s || !columns_in || !columns_out || num_rows < 0 || num_rows > s->max_rows) return MDR_EINVAL;
    for (int c = 0; c < s->num_columns; c++)
        if (!columns_in[c] || !columns_out[c] || ((uintptr_t)columns_in[c] & 15u) || ((uintptr_t)columns_out[c] & 3u))
            return MDR_EINVAL;
    REQ_CUDA(cudaSetDevice(s->gpu_id));
    cudaStream_t st = s->stream;
    int new_rows = num_rows;
    s->last_launches = 0;
    REQ_CUDA(cudaEventRecord(s->ev0, st));
    if (num_rows > 0) {
        const uint32_t *keys_col = (const uint32_t *)columns_in[s->key_col];
        const int tiles = (num_rows + kTile - 1) / kTile;
        REQ_CUDA(cudaMemsetAsync(s->bins, 0, 4 * ((size_t)kMaxPasses * kDigits + kDigits + (size_t)tiles * kDigits * s->passes), st));
        int hgrid = (num_rows / 4 + kThreads - 1) / kThreads;
        if (hgrid > 8 * s->sms) hgrid = 8 * s->sms;
        if (hgrid < 1) hgrid = 1;
        sortHistogramKernel<<<hgrid, kThreads, 0, st>>>(keys_col, num_rows, s->passes, s->bins);
        sortScanKernel<<<s->passes, kThreads, 0, st>>>(s->bins);
        s->last_launches += 2;
        const uint32_t *src_k = keys_col;
        const int32_t *src_v = nullptr;                  // pass 0 reads the identity permutation
        for (int p = 0; p < s->passes; p++) {
            uint32_t *dst_k = p % 2 == 0 ? s->keys_a : s->keys_b;
            int32_t *dst_v = p % 2 == 0 ? s->vals_a : s->vals_b;
            sortOnesweepKernel<<<tiles, kThreads, sizeof(OnesweepSmem), st>>>(
                src_k, src_v, dst_k, dst_v, s->bins + p * kDigits, s->lookback + (size_t)p * tiles * kDigits,
                s->counters + p, num_rows, 8 * p);
            s->last_launches++;
            src_k = dst_k;
            src_v = dst_v;
        }
        // Columns: out-of-place ones straight into columns_out in ONE launch; a column whose in and
        // out pointers are the same goes through the staging buffer (stageColumn + rearrangeColumn).
        // Rows past the cut keep their sorted (dead) content, as in the reference.
        std::vector<ColumnDesc> descs;
        std::vector<int> inplace;
        for (int c = 0; c < s->num_columns; c++) {
            const bool is_key = c == s->key_col;
            if (columns_in[c] == columns_out[c] && !is_key) { inplace.push_back(c); continue; }
            { ColumnDesc d; memset(&d, 0, sizeof(d)); d.src = (const char *)columns_in[c]; d.dst = (char *)columns_out[c]; d.words = s->col_bytes[c] / 4; d.is_key = is_key; descs.push_back(d); }
        }
        const int num_direct = (int)descs.size();
        for (const ColumnDesc &d : descs)
            if ((int64_t)d.words * num_rows >= (int64_t)1 << 32) return MDR_ECAPACITY;   // word index is 32-bit
        if (!inplace.empty() && !s->staging && cudaMalloc((void **)&s->staging, s->staging_bytes) != cudaSuccess) {
            cudaGetLastError();
            s->staging = nullptr;
            return MDR_ECAPACITY;       // no room for the staging copy of an in-place column
        }
        for (int c : inplace)
            { ColumnDesc d; memset(&d, 0, sizeof(d)); d.src = (const char *)columns_in[c]; d.dst = s->staging; d.words = s->col_bytes[c] / 4; descs.push_back(d); }
        if (descs.size() != s->h_cols.size() || memcmp(descs.data(), s->h_cols.data(), sizeof(ColumnDesc) * descs.size())) {
            s->h_cols = descs;      // same buffers as last run (the usual case): nothing to upload
            REQ_CUDA(cudaMemcpyAsync(s->d_cols, s->h_cols.data(), sizeof(ColumnDesc) * descs.size(), cudaMemcpyHostToDevice, st));
        }
        auto launch = [&](int first, int count) {
            int64_t max_words = 0;
            for (int k = first; k < first + count; k++)
                if ((int64_t)descs[k].words * num_rows > max_words) max_words = (int64_t)descs[k].words * num_rows;
            int64_t gx = (max_words + 255) / 256;
            if (gx > 16 * (int64_t)s->sms) gx = 16 * (int64_t)s->sms;
            if (gx < 1) gx = 1;
            tableGatherKernel<<<dim3((unsigned)gx, (unsigned)count), 256, 0, st>>>(s->d_cols + first, src_v, src_k, num_rows);
            s->last_launches++;
        };
        if (num_direct) launch(0, num_direct);
        for (size_t k = 0; k < inplace.size(); k++) {
            launch(num_direct + (int)k, 1);
            REQ_CUDA(cudaMemcpyAsync((void *)columns_out[inplace[k]], s->staging,
                                     (size_t)s->col_bytes[inplace[k]] * num_rows, cudaMemcpyDeviceToDevice, st));
        }
        // resizeTable: the rows before digit 255 of the last pass survive a WorldID sort
        const uint32_t *cut = s->world_sort ? s->bins + (s->passes - 1) * kDigits + 255 : nullptr;
        if (entity_row) {
            entityRemapKernel<<<(num_rows + 255) / 256, 256, 0, st>>>((const int32_t *)columns_out[0], num_rows, cut,
                                                                      entity_row, max_entity_ids);
            s->last_launches++;
        }
        if (cut) REQ_CUDA(cudaMemcpyAsync(&s->host_cut, cut, 4, cudaMemcpyDeviceToHost, st));
    }
    REQ_CUDA(cudaEventRecord(s->ev1, st));
    REQ_CUDA(cudaStreamSynchronize(st));
    REQ_CUDA(cudaEventElapsedTime(&s->last_ms, s->ev0, s->ev1));
    if (num_rows > 0 && s->world_sort) new_rows = (int)s->host_cut;
    if (out_num_rows) *out_num_rows = new_rows;
    return MDR_OK;
}

// ---- allocation-free variant (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
// forbid it): the key sort of mdr_sort_run -- histogram, scan, one Onesweep pass per digit -- with ALL its scratch
// (one key / permutation ping-pong pair, digit bins, tile counters, look-back words) in the file-scope device array
// below instead of buffers from mdr_sort_create. The array costs kStaticScratchBytes on every GPU that loads the
// library, until exit; every call shares it, so calls that overlap on the GPU race (one call at a time per device,
// or calls ordered on one stream). It keeps the last call's contents: the words that must start at zero (bins,
// counters, look-back) are cleared by every call; the ping-pong pair is fully overwritten before it is read.
constexpr int kStaticMaxRows = MDR_SORT_STATIC_MAX_ROWS;
constexpr int kStaticMaxTiles = (kStaticMaxRows + kTile - 1) / kTile;
constexpr size_t kStaticBinWords = (size_t)kMaxPasses * kDigits + kDigits + (size_t)kStaticMaxTiles * kDigits * kMaxPasses;
constexpr size_t kStaticScratchWords = 2 * (size_t)kStaticMaxRows + kStaticBinWords;
static __device__ __align__(256) uint32_t g_sort_static_scratch[kStaticScratchWords];

// In host code `g_sort_static_scratch` is a host-side stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, once per device, with that device current.
static uint32_t *staticScratchOn(int gpu_id)
{
    static uint32_t *ptrs[64] = {};
    static bool ready[64] = {};
    if (gpu_id < 0 || gpu_id >= 64) return nullptr;
    if (!ready[gpu_id]) {
        void *ptr = nullptr;
        if (cudaGetSymbolAddress(&ptr, g_sort_static_scratch) != cudaSuccess || !ptr) { cudaGetLastError(); return nullptr; }
        if (cudaFuncSetAttribute(sortOnesweepKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)sizeof(OnesweepSmem)) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        ptrs[gpu_id] = (uint32_t *)ptr;
        ready[gpu_id] = true;
    }
    return ptrs[gpu_id];
}

int32_t mdr_sort_static_max_rows(void) { return kStaticMaxRows; }
int64_t mdr_sort_static_scratch_bytes(void) { return (int64_t)(kStaticScratchWords * 4); }

int32_t mdr_sort_pairs_static(int32_t gpu_id, const uint32_t *keys_in, uint32_t *keys_out, int32_t *perm_out,
                              int32_t num_rows, int32_t num_key_bits, void *stream)
{
    if (!keys_in || !keys_out || !perm_out || num_rows < 0 || num_key_bits < 1 || num_key_bits > 32 ||
        (const void *)keys_in == (const void *)keys_out || ((uintptr_t)keys_in & 15u) || ((uintptr_t)keys_out & 3u) ||
        ((uintptr_t)perm_out & 3u))
        return MDR_EINVAL;
    if (num_rows > kStaticMaxRows) {
        fprintf(stderr, "madrona_b200: mdr_sort_pairs_static sorts at most %d rows (asked for %d): its scratch is a fixed "
                        "device array; use mdr_sort_create / mdr_sort_run for larger tables\n", kStaticMaxRows, num_rows);
        return MDR_ECAPACITY;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || gpu_id < 0 || gpu_id >= ndev) { cudaGetLastError(); return MDR_ENODEVICE; }
    if (cudaSetDevice(gpu_id) != cudaSuccess) { cudaGetLastError(); return MDR_ENODEVICE; }
    uint32_t *scratch = staticScratchOn(gpu_id);
    if (!scratch) return MDR_ESTATE;
    if (num_rows == 0) return MDR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int passes = (num_key_bits + 7) / 8;
    const int tiles = (num_rows + kTile - 1) / kTile;
    uint32_t *keys_s = scratch;
    int32_t *vals_s = (int32_t *)(scratch + kStaticMaxRows);
    uint32_t *bins = scratch + 2 * (size_t)kStaticMaxRows;
    uint32_t *counters = bins + kMaxPasses * kDigits;
    uint32_t *lookback = counters + kDigits;
    if (cudaMemsetAsync(bins, 0, 4 * ((size_t)kMaxPasses * kDigits + kDigits + (size_t)tiles * kDigits * passes), st) != cudaSuccess) {
        cudaGetLastError();
        return MDR_ESTATE;
    }
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, gpu_id);
    int hgrid = (num_rows / 4 + kThreads - 1) / kThreads;
    if (hgrid > 8 * sms) hgrid = 8 * sms;
    if (hgrid < 1) hgrid = 1;
    sortHistogramKernel<<<hgrid, kThreads, 0, st>>>(keys_in, num_rows, passes, bins);
    sortScanKernel<<<passes, kThreads, 0, st>>>(bins);
    const uint32_t *src_k = keys_in;
    const int32_t *src_v = nullptr;                      // pass 0 reads the identity permutation
    for (int p = 0; p < passes; p++) {
        const bool to_out = (passes - 1 - p) % 2 == 0;   // the last pass lands in the caller's buffers
        uint32_t *dst_k = to_out ? keys_out : keys_s;
        int32_t *dst_v = to_out ? perm_out : vals_s;
        sortOnesweepKernel<<<tiles, kThreads, sizeof(OnesweepSmem), st>>>(
            src_k, src_v, dst_k, dst_v, bins + p * kDigits, lookback + (size_t)p * tiles * kDigits, counters + p,
            num_rows, 8 * p);
        src_k = dst_k;
        src_v = dst_v;
    }
    return cudaGetLastError() == cudaSuccess ? MDR_OK : MDR_ESTATE;
}

float mdr_sort_last_ms(mdr_sort *s, int32_t *launches)
{
    if (!s) return 0.f;
    if (launches) *launches = s->last_launches;
    return s->last_ms;
}

}  // extern "C"
