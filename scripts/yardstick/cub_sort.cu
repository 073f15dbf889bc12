// Library yardstick for the sort stage of the LBVH build: cub::DeviceRadixSort::SortPairs on the
// same kind of input k_sort_hist + k_sort_pass x4 see — n (30-bit Morton key, u32 id) pairs of uniformly random points.
// NOT part of the product: built into a scratch binary by scripts/gpu.sh cubsort, never linked into librtb200.so.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a scripts/yardstick/cub_sort.cu -o gpurun_out/cub_sort && gpurun_out/cub_sort [n]
#include <cub/device/device_radix_sort.cuh>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); return 1; } } while (0)

__device__ __forceinline__ uint32_t spread3(uint32_t v) {  // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__global__ void k_fill(uint32_t *keys, uint32_t *vals, uint32_t n, uint64_t seed) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (i + 1);  // splitmix64
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const uint32_t x = (uint32_t)(z & 1023u), y = (uint32_t)((z >> 10) & 1023u), w = (uint32_t)((z >> 20) & 1023u);
    keys[i] = spread3(x) | (spread3(y) << 1) | (spread3(w) << 2);
    vals[i] = i;
}

int main(int argc, char **argv) {
    const uint32_t n = argc > 1 ? (uint32_t)strtoull(argv[1], nullptr, 10) : 10000010u;
    const int reps = 20;
    uint32_t *k0, *k1, *v0, *v1, *src_k, *src_v;
    CK(cudaMalloc(&k0, 4ull * n)); CK(cudaMalloc(&k1, 4ull * n)); CK(cudaMalloc(&v0, 4ull * n)); CK(cudaMalloc(&v1, 4ull * n));
    CK(cudaMalloc(&src_k, 4ull * n)); CK(cudaMalloc(&src_v, 4ull * n));
    k_fill<<<(n + 255) / 256, 256>>>(src_k, src_v, n, 0x5EED0001ull);
    CK(cudaGetLastError());
    // a buffer larger than L2 to flush between repetitions (the build's sort input comes from a kernel that wrote it: warm-ish,
    // but 80 MB of keys + ids do not fit beside the rest of the build's traffic; both figures are printed)
    char *flush; const size_t flush_bytes = 512ull << 20;
    CK(cudaMalloc(&flush, flush_bytes));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int end_bit : {30, 32}) {
        for (int cold = 0; cold < 2; cold++) {
            cub::DoubleBuffer<uint32_t> dk(k0, k1), dv(v0, v1);
            size_t tmp_bytes = 0; void *tmp = nullptr;
            CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, dk, dv, (int)n, 0, end_bit));
            CK(cudaMalloc(&tmp, tmp_bytes));
            std::vector<float> ms;
            for (int r = 0; r < reps + 3; r++) {
                cub::DoubleBuffer<uint32_t> a(k0, k1), b(v0, v1);
                CK(cudaMemcpyAsync(k0, src_k, 4ull * n, cudaMemcpyDeviceToDevice));
                CK(cudaMemcpyAsync(v0, src_v, 4ull * n, cudaMemcpyDeviceToDevice));
                if (cold) CK(cudaMemsetAsync(flush, r, flush_bytes));
                CK(cudaEventRecord(e0));
                CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, a, b, (int)n, 0, end_bit));
                CK(cudaEventRecord(e1));
                CK(cudaEventSynchronize(e1));
                float t; CK(cudaEventElapsedTime(&t, e0, e1));
                if (r >= 3) ms.push_back(t);
                if (r == reps + 2) {  // sortedness check of the last repetition
                    std::vector<uint32_t> h(n);
                    CK(cudaMemcpy(h.data(), a.Current(), 4ull * n, cudaMemcpyDeviceToHost));
                    if (!std::is_sorted(h.begin(), h.end())) { fprintf(stderr, "not sorted\n"); return 2; }
                }
            }
            std::sort(ms.begin(), ms.end());
            printf("cub::DeviceRadixSort::SortPairs n=%u bits [0,%d) %s: median %.4f ms  min %.4f ms  (%.1f Gkeys/s, temp %.1f MB)\n", n, end_bit,
                   cold ? "L2 flushed" : "L2 warm  ", ms[ms.size() / 2], ms[0], n / (ms[ms.size() / 2] * 1e-3) / 1e9, tmp_bytes / 1e6);
            CK(cudaFree(tmp));
        }
    }
    return 0;
}
