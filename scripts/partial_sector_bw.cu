// Microbenchmark (development aid): what a half-written 32 B sector per row costs.  Rows of
// `units` x 16 B are written at a stride of 2400 B (the concatenate layout of two dim-300
// readers: each launch writes 1200 B = 37.5 sectors of every 2400 B union row).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/partial_sector_bw.bin_ scripts/partial_sector_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__global__ void rowsAtStride(float4* out, uint64_t rows, int units, int strideUnits, int firstUnit)
{
    const uint32_t lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const uint64_t warps = uint64_t(gridDim.x) * blockDim.x / 32;
    const float4 v = make_float4(1.f, 2.f, 3.f, float(lane));
    for (uint64_t t = warp; t < rows / 32; t += warps) {
        for (uint32_t u = lane; u < 32u * units; u += 32) {
            const uint32_t r = u / units, c = u % units;
            __stcs(out + (t * 32 + r) * strideUnits + firstUnit + c, v);
        }
    }
}

int main()
{
    const uint64_t rows = 1000000 / 32 * 32;
    float4* out;
    cudaMalloc(&out, rows * 150 * 16);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    struct Case { int units, stride, first; const char* what; };
    const Case cases[] = {
        {75, 75, 0, "1200 B rows, contiguous (plain decode)"},
        {75, 150, 0, "1200 B at stride 2400, offset 0    (concatenate, reader 0: last sector half written)"},
        {75, 150, 75, "1200 B at stride 2400, offset 1200 (concatenate, reader 1: first sector half written)"},
        {76, 150, 0, "1216 B at stride 2400, offset 0    (whole sectors)"},
        {74, 150, 76, "1184 B at stride 2400, offset 1216 (whole sectors)"},
        {72, 150, 0, "1152 B at stride 2400, offset 0    (whole 128 B lines)"},
    };
    for (const Case& c : cases) {
        for (int i = 0; i < 3; ++i) rowsAtStride<<<148 * 2, 320>>>(out, rows, c.units, c.stride, c.first);
        cudaEventRecord(e0);
        for (int i = 0; i < 20; ++i) rowsAtStride<<<148 * 2, 320>>>(out, rows, c.units, c.stride, c.first);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        ms /= 20;
        printf("%-95s %.4f ms  %.0f GB/s\n", c.what, ms, double(rows) * c.units * 16 / ms / 1e6);
    }
    return cudaDeviceSynchronize() != cudaSuccess;
}
