// Microbenchmark (development aid): shared-memory load throughput per SM for 32-, 64- and
// 128-bit conflict-free loads (lane l reads its own bank group), instructions and bytes per cycle.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/lds_width.bin_ scripts/lds_width.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

template <int WORDS>
__global__ void loads(uint32_t* out, long long* cycles, int iters)
{
    extern __shared__ uint32_t buf[];
    for (uint32_t i = threadIdx.x; i < 16384; i += blockDim.x) buf[i] = i * 2654435761u;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t base = static_cast<uint32_t>(__cvta_generic_to_shared(buf)) + lane * WORDS * 4;
    uint32_t acc = 0, off = (threadIdx.x / 32) * 128 * WORDS;
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const uint32_t a = base + ((off + u * 128 * WORDS) & (16384 * 4 - 128 * WORDS));
            if (WORDS == 1) {
                uint32_t x;
                asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x) : "r"(a));
                acc ^= x;
            } else if (WORDS == 2) {
                uint32_t x, y;
                asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(x), "=r"(y) : "r"(a));
                acc ^= x ^ y;
            } else {
                uint32_t x, y, z, w;
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(a));
                acc ^= x ^ y ^ z ^ w;
            }
        }
        off += 8 * 128 * WORDS;
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int WORDS>
void run(uint32_t* out, long long* cyc)
{
    const int iters = 2048;
    cudaFuncSetAttribute(loads<WORDS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int warps : {4, 10, 16, 32}) {
        loads<WORDS><<<148, warps * 32, 64 * 1024>>>(out, cyc, iters);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0;
        for (int i = 0; i < 148; ++i) avg += double(h[i]);
        avg /= 148;
        const double instr = double(warps) * iters * 8;
        printf("LDS.%-3d warps/SM %2d: %.3f instr/cycle/SM, %.1f B/cycle/SM  %s\n", WORDS * 32, warps, instr / avg,
               instr * 32 * WORDS * 4 / avg, cudaGetErrorString(cudaGetLastError()));
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
