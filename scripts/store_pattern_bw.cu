// Microbenchmark (development aid): HBM write rate of the decode kernel's STORE PATTERN in
// isolation.  A warp owns a tile of 32 consecutive 1200 B rows.  It emits the tile chunk
// by chunk: in one store instruction `seg` lanes cover seg*16 contiguous, line-aligned
// bytes of a row (rows lag by their start offset modulo 128 B, like the kernel's flush) and
// 32/seg rows are covered at once; between two chunks the warp idles for `delay` cycles
// (the decode of the next chunk).  Compared with the same bytes written tile-contiguously.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/store_pattern_bw.bin_ scripts/store_pattern_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#ifndef ROW_UNITS_DEF
#define ROW_UNITS_DEF 75
#endif
constexpr int ROW_UNITS = ROW_UNITS_DEF;  // 75: 1200 B

// POLICY 0: st.global.cs (streaming), 1: plain st.global (write-back), 2: st.global.cg, 3: st.global.wt
template <int POLICY>
__device__ __forceinline__ void store16(float4* p, float4 v)
{
    if (POLICY == 0) {
        __stcs(p, v);
    } else if (POLICY == 1) {
        *p = v;
    } else if (POLICY == 2) {
        __stcg(p, v);
    } else {
        __stwt(p, v);
    }
}

template <int POLICY>
__global__ void pattern(float4* out, uint64_t rows, int seg, int delay)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const uint64_t warps = uint64_t(gridDim.x) * blockDim.x / 32;
    const int fr = lane / seg, fq = lane % seg, rowsPer = 32 / seg;
    const float4 v = make_float4(1.f, 2.f, 3.f, float(lane));
    for (uint64_t t = warp; t < rows / 32; t += warps) {
        float4* tile = out + t * 32 * ROW_UNITS;
        const int chunks = (ROW_UNITS + 7 + seg - 1) / seg;  // + lag drain
        for (int c = 0; c < chunks; ++c) {
            for (int it = 0; it < 32 / rowsPer; ++it) {
                const int r = it * rowsPer + fr;
                float4* row = tile + r * ROW_UNITS;
                const int a = int((reinterpret_cast<uintptr_t>(row) >> 4) & 7);
                const int j = c * seg - a + fq;
                if (j >= 0 && j < ROW_UNITS) {
                    store16<POLICY>(row + j, v);
                }
            }
            if (delay) {
                const long long t0 = clock64();
                while (clock64() - t0 < delay) {
                }
            }
        }
    }
}

__global__ void contiguous(float4* out, uint64_t rows)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const uint64_t warps = uint64_t(gridDim.x) * blockDim.x / 32;
    const float4 v = make_float4(1.f, 2.f, 3.f, float(lane));
    for (uint64_t t = warp; t < rows / 32; t += warps) {
        float4* tile = out + t * 32 * ROW_UNITS;
        for (uint32_t u = lane; u < 32 * ROW_UNITS; u += 32) {
            __stcs(tile + u, v);
        }
    }
}

template <class F>
float timeIt(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) launch();
    cudaEventRecord(e0);
    for (int i = 0; i < 20; ++i) launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / 20;
}

int main()
{
    const uint64_t rows = 1000000 / 32 * 32;
    float4* out;
    cudaMalloc(&out, rows * ROW_UNITS * 16);
    const double gb = double(rows) * ROW_UNITS * 16 / 1e9;
    for (int warps : {10, 20}) {
        const int block = 320, grid = 148 * warps / 10;
        float ms = timeIt([&] { contiguous<<<grid, block>>>(out, rows); });
        printf("contiguous tiles,            %2d warps/SM: %.4f ms  %.0f GB/s\n", warps, ms, gb / ms * 1e3);
        for (int seg : {8, 16, 32}) {
            for (int delay : {0, 1000, 3000}) {
                ms = timeIt([&] { pattern<0><<<grid, block>>>(out, rows, seg, delay); });
                printf("seg %2d lanes (%3d B runs), delay %4d, %2d warps/SM: %.4f ms  %.0f GB/s\n", seg, seg * 16, delay,
                       warps, ms, gb / ms * 1e3);
                if (seg == 8 && delay == 1000) {
                    const float m1 = timeIt([&] { pattern<1><<<grid, block>>>(out, rows, seg, delay); });
                    const float m2 = timeIt([&] { pattern<2><<<grid, block>>>(out, rows, seg, delay); });
                    const float m3 = timeIt([&] { pattern<3><<<grid, block>>>(out, rows, seg, delay); });
                    printf("    same with st.global %.0f GB/s, st.global.cg %.0f GB/s, st.global.wt %.0f GB/s\n",
                           gb / m1 * 1e3, gb / m2 * 1e3, gb / m3 * 1e3);
                }
            }
        }
    }
    const cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        printf("CUDA error: %s\n", cudaGetErrorString(err));
        return 1;
    }
    return 0;
}
