// Microbenchmark (development aid, not part of the product): latency of a dependent
// LDS chain on one SM as a function of resident warps and of concurrent global stores.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/lds_latency scripts/lds_latency.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void chain(uint32_t* out, long long* cycles, int iters, int storeEvery, float4* sink, int dynPad)
{
    extern __shared__ uint32_t lut[];  // 512 entries x 32 replicas
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t i = threadIdx.x; i < 512 * 32; i += blockDim.x) {
        lut[i] = (i / 32) * 2654435761u;
    }
    __syncthreads();
    uint32_t hi = threadIdx.x * 747796405u + blockIdx.x, lo = hi * 31u;
    uint32_t acc = 0;
    float4* dst = sink + (size_t)(blockIdx.x * blockDim.x + threadIdx.x);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        const uint32_t e = lut[((hi >> 23) << 5) | lane];
        hi = __funnelshift_l(lo, hi, e);
        lo = lo << (e & 31) | (e >> 7);
        acc += e;
        if (storeEvery && (i & (storeEvery - 1)) == 0) {
            const float4 v = make_float4(__uint_as_float(i), 0.f, 0.f, 0.f);
            if (dynPad == 1) {
                __stcs(dst, v);
            } else if (dynPad == 2) {
                *dst = v;
            } else if (dynPad == 3) {
                *reinterpret_cast<float*>(dst) = v.x;
            } else if (dynPad == 4) {
                __stwt(dst, v);
            } else if (dynPad == 5) {
                asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
            }
            dst += gridDim.x * blockDim.x;
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc + hi;
    if (threadIdx.x == 0) {
        cycles[blockIdx.x] = t1 - t0;
    }
}

int main()
{
    uint32_t* out;
    long long* cyc;
    float4* sink;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 148 * 8);
    const size_t sinkBytes = size_t(4) << 30;
    cudaMalloc(&sink, sinkBytes);
    const int iters = 4096;
    cudaFuncSetAttribute(chain, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int mode : {0, 1, 2, 3, 4, 5}) {
        const int storeEvery = mode ? 4 : 0;
        for (int warps : {1, 8, 24}) {
            for (int smemKB : {200}) {
                chain<<<148, warps * 32, smemKB * 1024>>>(out, cyc, iters, storeEvery, sink, mode);
                cudaDeviceSynchronize();
                long long h[148];
                cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
                double avg = 0;
                for (int i = 0; i < 148; ++i) avg += double(h[i]);
                avg /= 148;
                printf("mode %d storeEvery %d warps %2d smem %3d KB: %.1f cycles/iter (warp 0), %s\n", mode, storeEvery, warps, smemKB,
                       avg / iters, cudaGetErrorString(cudaGetLastError()));
            }
        }
    }
    return 0;
}
