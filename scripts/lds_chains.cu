// Microbenchmark (development aid): throughput of dependent LOP3 -> LDS -> SHF chains on one
// SM as a function of resident warps and of independent chains per lane (the decode kernel's
// per-symbol chain; is it latency or the shared-memory pipe that bounds it?).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/lds_chains.bin_ scripts/lds_chains.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int CHAINS>
__global__ void chains(uint32_t* out, long long* cycles, int iters)
{
    extern __shared__ uint32_t lut[];  // 512 entries x 32 bank replicas
    const uint32_t lane = threadIdx.x & 31;
    for (uint32_t i = threadIdx.x; i < 512 * 32; i += blockDim.x) {
        lut[i] = ((i / 32) * 2654435761u) | 3u;
    }
    __syncthreads();
    uint32_t lo[CHAINS], hi[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
        hi[c] = (threadIdx.x + 977 * c) * 747796405u + blockIdx.x;
        lo[c] = hi[c] * 31u;
    }
    const uint32_t base = static_cast<uint32_t>(__cvta_generic_to_shared(lut)) | (lane * 4);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
            uint32_t e;
            const uint32_t a = (lo[c] & (511u << 7)) | base;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(a));
            lo[c] = __funnelshift_r(lo[c], hi[c], e);
            hi[c] = __funnelshift_r(hi[c], lo[c] ^ e, e);
        }
    }
    const long long t1 = clock64();
    uint32_t acc = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) acc += lo[c] + hi[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int CHAINS>
void run(uint32_t* out, long long* cyc)
{
    const int iters = 4096;
    cudaFuncSetAttribute(chains<CHAINS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int warps : {1, 4, 10, 16, 32}) {
        chains<CHAINS><<<148, warps * 32, 64 * 1024>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0;
        for (int i = 0; i < 148; ++i) avg += double(h[i]);
        avg /= 148;
        printf("chains/lane %d, warps/SM %2d: %.1f cycles per round, %.3f LDS per cycle and SM  %s\n", CHAINS, warps,
               avg / iters, double(CHAINS) * warps * iters / avg, cudaGetErrorString(cudaGetLastError()));
    }
}

int main()
{
    uint32_t* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 148 * 8);
    run<1>(out, cyc);
    run<2>(out, cyc);
    run<4>(out, cyc);
    return 0;
}
