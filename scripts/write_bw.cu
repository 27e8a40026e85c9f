// Microbenchmark (development aid): pure-write HBM bandwidth on one GPU, for the
// store patterns the decode kernel could use.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/write_bw.bin_ scripts/write_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

// mode 0: fully coalesced float4 streaming stores (512 B per warp instruction)
// mode 1: same, default cache policy
// mode 2: decode-kernel pattern: a warp owns 32 rows of `dim` floats; one store
//         instruction writes 128 B of each of 4 rows; 32-column chunks left to right
// mode 3: like 2 but a warp writes one full row (dim floats) at a time, 512 B per instruction
__global__ void writer(float4* out, size_t rows, uint32_t dim, int mode)
{
    const uint32_t lane = threadIdx.x & 31;
    const size_t warp = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) / 32;
    const size_t warps = size_t(gridDim.x) * blockDim.x / 32;
    const float4 v = make_float4(1.f, 2.f, 3.f, 4.f);
    if (mode <= 1) {
        const size_t total = rows * dim / 4;
        for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
            if (mode == 0) __stcs(out + i, v); else out[i] = v;
        }
    } else if (mode == 2 || mode >= 4) {
        // seg lanes cover one row segment (seg*16 B); 32/seg rows per store instruction
        const uint32_t seg = mode == 2 ? 8 : mode == 4 ? 16 : 32;
        const uint32_t fr = lane / seg, fq = lane % seg, rowsPer = 32 / seg;
        for (size_t t = warp; t < rows / 32; t += warps) {
            for (uint32_t c0 = 0; c0 < dim; c0 += seg * 4) {
                if (c0 + fq * 4 < dim) {
                    for (uint32_t it = 0; it < 32 / rowsPer; ++it) {
                        const size_t r = t * 32 + it * rowsPer + fr;
                        __stcs(out + (r * dim + c0) / 4 + fq, v);
                    }
                }
            }
        }
    } else {
        for (size_t r = warp; r < rows; r += warps) {
            for (uint32_t c = lane; c < dim / 4; c += 32) {
                __stcs(out + r * (dim / 4) + c, v);
            }
        }
    }
}

__global__ void copier(const float4* in, float4* out, size_t total)
{
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
        __stcs(out + i, __ldcs(in + i));
    }
}

int main()
{
  for (uint32_t dim : {320u}) {
    printf("---- dim %u\n", dim);
    const size_t rows = 2200000 / 32 * 32;
    const size_t bytes = rows * dim * 4;
    float4 *out, *in;
    cudaMalloc(&out, bytes);
    cudaMalloc(&in, bytes);
    cudaMemset(in, 1, bytes);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    auto timeit = [&](const char* name, auto fn, double gb) {
        for (int i = 0; i < 3; ++i) fn();
        cudaEventRecord(a);
        for (int i = 0; i < 10; ++i) fn();
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        ms /= 10;
        printf("%-44s %.3f ms  %.1f GB/s  (%s)\n", name, ms, gb / (ms / 1e3), cudaGetErrorString(cudaGetLastError()));
    };
    const double gb = bytes / 1e9;
    timeit("cudaMemsetAsync", [&] { cudaMemsetAsync(out, 0, bytes); }, gb);
    for (int blocks : {148 * 2, 148 * 6}) {
        for (int mode = 0; mode < 6; ++mode) {
            char name[64];
            snprintf(name, sizeof(name), "writer mode %d grid %d x 256", mode, blocks);
            timeit(name, [&] { writer<<<blocks, 256>>>(out, rows, dim, mode); }, gb);
        }
    }
    timeit("copy read+write (bytes counted twice)", [&] { copier<<<148 * 16, 256>>>(in, out, bytes / 16); }, 2 * gb);
    cudaFree(out);
    cudaFree(in);
  }
    return 0;
}
