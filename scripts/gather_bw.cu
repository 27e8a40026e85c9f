// Microbenchmark (development aid, not part of the product): what HBM delivers for the
// TRAFFIC PATTERN of the batched lookup with no decode work at all -- every warp reads 32
// compressed-row-sized pieces (144 B each, 16 B aligned) at random or sequential row
// positions of a 317 MB table and writes 32 x 1200 B of contiguous fp32 output with
// 128-bit streaming stores.  This is the practical ceiling the decode kernel is up against.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/gather_bw.bin_ scripts/gather_bw.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr uint32_t ROW_UNITS = 9;     // 144 B per compressed row
constexpr uint32_t OUT_UNITS = 75;    // 1200 B per output row

// mode 0: gather + write, mode 1: write only, mode 2: gather only (one 16 B store per row)
__global__ void traffic(const uint4* __restrict__ table, const int32_t* __restrict__ ids, uint64_t n, float4* out,
                        int mode)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const uint64_t warps = uint64_t(gridDim.x) * blockDim.x / 32;
    for (uint64_t t = warp; t < n / 32; t += warps) {
        uint32_t acc = 0;
        if (mode != 1) {
            const uint4* src = table + uint64_t(ids[t * 32 + lane]) * ROW_UNITS;
            uint4 v[ROW_UNITS];
#pragma unroll
            for (uint32_t k = 0; k < ROW_UNITS; ++k) {
                asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(v[k].x), "=r"(v[k].y), "=r"(v[k].z), "=r"(v[k].w)
                             : "l"(src + k));
            }
#pragma unroll
            for (uint32_t k = 0; k < ROW_UNITS; ++k) {
                acc ^= v[k].x ^ v[k].y ^ v[k].z ^ v[k].w;
            }
        }
        float4* dst = out + t * 32 * OUT_UNITS;
        const float4 val = make_float4(__uint_as_float(acc), 1.f, 2.f, 3.f);
        if (mode == 2) {
            __stcs(dst + lane * OUT_UNITS, val);
        } else {
            for (uint32_t u = lane; u < 32 * OUT_UNITS; u += 32) {
                __stcs(dst + u, val);
            }
        }
    }
}

int main()
{
    const uint64_t rows = 2200000, n = 1000000 / 32 * 32;
    uint4* table;
    float4* out;
    int32_t* ids;
    cudaMalloc(&table, rows * ROW_UNITS * 16);
    cudaMalloc(&out, n * OUT_UNITS * 16);
    cudaMalloc(&ids, n * 4);
    cudaMemset(table, 1, rows * ROW_UNITS * 16);
    std::vector<int32_t> h(n);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int pattern = 0; pattern < 2; ++pattern) {
        uint64_t s = 88172645463325252ull;
        for (uint64_t i = 0; i < n; ++i) {
            s ^= s << 13;
            s ^= s >> 7;
            s ^= s << 17;
            h[i] = pattern == 0 ? int32_t(s % rows) : int32_t(i);
        }
        cudaMemcpy(ids, h.data(), n * 4, cudaMemcpyHostToDevice);
        for (int mode = 0; mode < 3; ++mode) {
            for (int warpsPerSm : {16, 32, 64}) {
                const int block = 256, grid = 148 * warpsPerSm * 32 / block;
                for (int i = 0; i < 3; ++i) {
                    traffic<<<grid, block>>>(table, ids, n, out, mode);
                }
                const int reps = 20;
                cudaEventRecord(e0);
                for (int i = 0; i < reps; ++i) {
                    traffic<<<grid, block>>>(table, ids, n, out, mode);
                }
                cudaEventRecord(e1);
                cudaEventSynchronize(e1);
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= reps;
                const double rd = mode == 1 ? 0 : double(n) * (ROW_UNITS * 16 + 4);
                const double wr = mode == 2 ? double(n) * 16 : double(n) * OUT_UNITS * 16;
                printf("%s ids, %s, %2d warps/SM: %.4f ms  %.0f GB/s (read %.0f MB, write %.0f MB)  %.3g rows/s\n",
                       pattern == 0 ? "random    " : "sequential",
                       mode == 0 ? "gather+write" : mode == 1 ? "write only  " : "gather only ", warpsPerSm, ms,
                       (rd + wr) / ms / 1e6, rd / 1e6, wr / 1e6, n / (ms / 1e3));
            }
        }
    }
    if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("CUDA error\n");
        return 1;
    }
    return 0;
}
