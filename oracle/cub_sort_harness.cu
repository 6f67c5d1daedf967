// TEST INFRASTRUCTURE ONLY: pins the archetype sort on the third-party algorithm the reference's sort nodes
// adapt. src/mw/device/sort_archetype.cpp:960-1392 is CUB's Onesweep radix sort (agent_radix_sort_onesweep.cuh)
// cut into task-graph nodes: (key, row index) pairs sorted by the low 8 * passes bits of a u32 key. This harness
// runs cub::DeviceRadixSort::SortPairs(keys, row indices, n, begin_bit = 0, end_bit) on the same keys and dumps
// sorted keys + permutation. The CUB checkout vendored by the reference (external/cub @ 09686d2) does not compile
// with this image's CUDA 12.9 (its Thrust refuses foreign CUB trees, and util_device.cuh no longer matches the
// runtime API); the CUB shipped WITH CUDA 12.9 is the same library, later release, and is what is compiled here.
//   cub_sort_ref <in.bin> <out.bin>     in: i32 n, i32 end_bit, u32 keys[n]     out: u32 keys[n], i32 perm[n]
#include <cub/device/device_radix_sort.cuh>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 2; } } while (0)

int main(int argc, char **argv)
{
    if (argc != 3) { fprintf(stderr, "usage: %s in.bin out.bin\n", argv[0]); return 1; }
    FILE *f = fopen(argv[1], "rb");
    if (!f) return 1;
    int32_t n = 0, end_bit = 32;
    if (fread(&n, 4, 1, f) != 1 || fread(&end_bit, 4, 1, f) != 1 || n < 0) return 1;
    std::vector<uint32_t> keys((size_t)n);
    if (n && fread(keys.data(), 4, (size_t)n, f) != (size_t)n) return 1;
    fclose(f);
    std::vector<int32_t> idx((size_t)n);
    for (int32_t i = 0; i < n; i++) idx[(size_t)i] = i;
    uint32_t *dk_in = nullptr, *dk_out = nullptr;
    int32_t *dv_in = nullptr, *dv_out = nullptr;
    const size_t nb = (size_t)(n > 0 ? n : 1) * 4;
    CK(cudaMalloc(&dk_in, nb)); CK(cudaMalloc(&dk_out, nb)); CK(cudaMalloc(&dv_in, nb)); CK(cudaMalloc(&dv_out, nb));
    CK(cudaMemcpy(dk_in, keys.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dv_in, idx.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
    if (n > 0) {
        size_t tmp_bytes = 0;
        CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk_in, dk_out, dv_in, dv_out, n, 0, end_bit));
        void *tmp = nullptr;
        CK(cudaMalloc(&tmp, tmp_bytes ? tmp_bytes : 1));
        CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, dk_in, dk_out, dv_in, dv_out, n, 0, end_bit));
        CK(cudaDeviceSynchronize());
    }
    CK(cudaMemcpy(keys.data(), dk_out, (size_t)n * 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(idx.data(), dv_out, (size_t)n * 4, cudaMemcpyDeviceToHost));
    f = fopen(argv[2], "wb");
    if (!f) return 1;
    fwrite(keys.data(), 4, (size_t)n, f);
    fwrite(idx.data(), 4, (size_t)n, f);
    fclose(f);
    return 0;
}
