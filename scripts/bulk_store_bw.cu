// Microbenchmark (development aid): can the flush of the decode kernel go through the TMA
// engine?  Every lane owns one output row of its warp's 32-row tile, fills a chunk of it in
// its own shared-memory staging row with 128-bit STS, and hands the chunk to
// cp.async.bulk.global.shared::cta (one bulk copy per lane and chunk).  Compared with plain
// coalesced st.global.cs.v4 of the same bytes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/bulk_store_bw.bin_ scripts/bulk_store_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr uint32_t ROW_BYTES = 1200;

__device__ __forceinline__ uint32_t smemAddr(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// chunkBytes: 128 or 256 (last chunk of a row is shorter).  Two staging buffers per warp.
template <uint32_t CHUNK>
__global__ void bulkWriter(unsigned char* out, uint64_t rows)
{
    extern __shared__ __align__(128) unsigned char smem[];
    constexpr uint32_t STRIDE = CHUNK + 16;  // 128-bit stores of 8 lanes cover all 32 banks
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x / 32, warpsPerCta = blockDim.x / 32;
    const uint32_t stage0 = smemAddr(smem) + warp * 2 * 32 * STRIDE;
    const uint64_t tiles = rows / 32;
    uint32_t parity = 0;
    for (uint64_t t = uint64_t(blockIdx.x) * warpsPerCta + warp; t < tiles; t += uint64_t(gridDim.x) * warpsPerCta) {
        unsigned char* rowPtr = out + (t * 32 + lane) * ROW_BYTES;
        for (uint32_t c0 = 0; c0 < ROW_BYTES; c0 += CHUNK) {
            const uint32_t bytes = min(CHUNK, ROW_BYTES - c0);
            const uint32_t buf = stage0 + parity * 32 * STRIDE + lane * STRIDE;
            // the buffer we are about to fill was handed to the TMA two chunks ago
            asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            for (uint32_t b = 0; b < bytes; b += 16) {
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(buf + b), "f"(1.f), "f"(2.f),
                             "f"(float(c0)), "f"(float(lane)));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(rowPtr + c0), "r"(buf),
                         "r"(bytes)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            parity ^= 1;
        }
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

__global__ void stgWriter(float4* out, uint64_t rows)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const uint64_t warps = uint64_t(gridDim.x) * blockDim.x / 32;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    for (uint64_t t = warp; t < rows / 32; t += warps) {
        float4* dst = out + t * 32 * (ROW_BYTES / 16);
        for (uint32_t u = lane; u < 32 * ROW_BYTES / 16; u += 32) {
            __stcs(dst + u, v);
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
    unsigned char* out;
    cudaMalloc(&out, rows * ROW_BYTES);
    const double gb = double(rows) * ROW_BYTES / 1e9;
    for (int warps : {4, 8, 10, 16}) {
        const int block = warps * 32;
        {
            const size_t smem = size_t(warps) * 2 * 32 * (128 + 16);
            cudaFuncSetAttribute(bulkWriter<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            const float ms = timeIt([&] { bulkWriter<128><<<148, block, smem>>>(out, rows); });
            printf("bulk 128 B/lane/chunk, %2d warps/SM: %.4f ms  %.0f GB/s  %.3g rows/s\n", warps, ms, gb / ms * 1e3,
                   rows / (ms / 1e3));
        }
        {
            const size_t smem = size_t(warps) * 2 * 32 * (256 + 16);
            cudaFuncSetAttribute(bulkWriter<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
            const float ms = timeIt([&] { bulkWriter<256><<<148, block, smem>>>(out, rows); });
            printf("bulk 256 B/lane/chunk, %2d warps/SM: %.4f ms  %.0f GB/s  %.3g rows/s\n", warps, ms, gb / ms * 1e3,
                   rows / (ms / 1e3));
        }
        {
            const float ms = timeIt([&] { stgWriter<<<148, block>>>(reinterpret_cast<float4*>(out), rows); });
            printf("st.global.cs.v4,       %2d warps/SM: %.4f ms  %.0f GB/s  %.3g rows/s\n", warps, ms, gb / ms * 1e3,
                   rows / (ms / 1e3));
        }
    }
    const cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        printf("CUDA error: %s\n", cudaGetErrorString(err));
        return 1;
    }
    return 0;
}
