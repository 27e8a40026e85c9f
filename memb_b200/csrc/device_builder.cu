// GPU converter: the bulk passes of the trained compressor on the device.
//
//   src/trained_compression.cpp:40-101   predict every value, Huffman-encode every row byte-aligned
//   src/kmeans.cpp:65-79                 predict = number of split points below the value
//   src/huffman_encoder.cpp:43-117       symbol frequencies -> code lengths -> canonical codes
//   src/bit_stream.h:18-34               MSB-first packing, zero padded to a byte
//
// What stays on the host, and why: the k-means FIT runs on the first 10 000 rows only
// (src/trained_compression.cpp:21,45-52) and its running-mean update is order dependent in fp32
// (src/kmeans.cpp:86-102) -- a parallel reduction would change centroids in the last bits and with
// them the file; the Huffman construction works on <= 255 counts.  Everything that touches all
// n x dim values runs here, and the result is BYTE-IDENTICAL to the host encoder's
// (tests/test_gpu_parity.py::test_gpu_converter_matches_cpu_builder): predict is the same
// lower_bound over the same fp32 split points, counts are integers, and the first-occurrence order
// of the symbols (which fixes the tie-breaks of the Huffman construction) is an exact minimum.
#include "builder.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <stdexcept>
#include <string>

namespace memb_b200 {
namespace {

void cudaOrThrow(cudaError_t err, const char* what)
{
    if (err != cudaSuccess) {
        throw std::runtime_error(std::string("GPU converter: ") + what + ": " + cudaGetErrorString(err));
    }
}
#define ENC_CHECK(expr) cudaOrThrow((expr), #expr)

constexpr int ENC_THREADS = 256;

// Pass 1.  symbol = number of split points strictly below the value (std::lower_bound with `<`,
// so a NaN maps to 0 exactly like on the host); per-symbol count and first position.
__global__ void encode_predict_kernel(const float* __restrict__ values, uint64_t total, const float* __restrict__ splits,
                                      uint32_t numSplits, uint8_t* __restrict__ symbols,
                                      unsigned long long* __restrict__ counts, unsigned long long* __restrict__ first)
{
    __shared__ float sSplits[256];
    __shared__ unsigned int sCounts[256];
    __shared__ unsigned long long sFirst[256];
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
        sSplits[i] = i < numSplits ? splits[i] : 0.f;
        sCounts[i] = 0;
        sFirst[i] = ~0ull;
    }
    __syncthreads();
    // a block walks one contiguous span, so that its per-symbol counter cannot overflow 32 bits
    const uint64_t perBlock = (total + gridDim.x - 1) / gridDim.x;
    const uint64_t begin = blockIdx.x * perBlock;
    const uint64_t end = min(total, begin + perBlock);
    for (uint64_t i = begin + threadIdx.x; i < end; i += blockDim.x) {
        const float v = values[i];
        uint32_t lo = 0, len = numSplits;
        while (len > 0) {
            const uint32_t half = len >> 1;
            if (sSplits[lo + half] < v) {
                lo += half + 1;
                len -= half + 1;
            } else {
                len = half;
            }
        }
        symbols[i] = static_cast<uint8_t>(lo);
        atomicAdd(&sCounts[lo], 1u);
        atomicMin(&sFirst[lo], static_cast<unsigned long long>(i));
    }
    __syncthreads();
    for (uint32_t s = threadIdx.x; s < 256; s += blockDim.x) {
        if (sCounts[s]) {
            atomicAdd(&counts[s], static_cast<unsigned long long>(sCounts[s]));
            atomicMin(&first[s], sFirst[s]);
        }
    }
}

struct DeviceCode {
    uint16_t codeOf[256];
    uint8_t lengthOf[256];
};

// Pass 2.  Bytes of every row: ceil(sum of its code lengths / 8).  One warp per row.
__global__ void encode_row_bytes_kernel(const uint8_t* __restrict__ symbols, uint64_t rows, uint32_t dim,
                                        const DeviceCode* __restrict__ code, uint32_t* __restrict__ rowBytes)
{
    __shared__ uint8_t sLen[256];
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
        sLen[i] = code->lengthOf[i];
    }
    __syncthreads();
    const uint64_t row = (blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x) / 32;
    const uint32_t lane = threadIdx.x & 31;
    if (row >= rows) {
        return;
    }
    const uint8_t* src = symbols + row * dim;
    uint32_t bits = 0;
    for (uint32_t i = lane; i < dim; i += 32) {
        bits += sLen[src[i]];
    }
    for (int off = 16; off; off >>= 1) {
        bits += __shfl_xor_sync(0xffffffffu, bits, off);
    }
    if (lane == 0) {
        rowBytes[row] = (bits + 7) / 8;
    }
}

// Pass 3.  MSB-first concatenation of the row's codes, zero padded to a byte, at rowStart[row].
// One thread per row with a 64-bit accumulator, exactly the host loop.
__global__ void encode_pack_kernel(const uint8_t* __restrict__ symbols, uint64_t rows, uint32_t dim,
                                   const DeviceCode* __restrict__ code, const uint64_t* __restrict__ rowStart,
                                   uint8_t* __restrict__ packed)
{
    __shared__ uint16_t sCode[256];
    __shared__ uint8_t sLen[256];
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
        sCode[i] = code->codeOf[i];
        sLen[i] = code->lengthOf[i];
    }
    __syncthreads();
    const uint64_t row = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (row >= rows) {
        return;
    }
    const uint8_t* src = symbols + row * dim;
    uint8_t* dst = packed + rowStart[row];
    uint64_t acc = 0;
    uint32_t have = 0;
    for (uint32_t i = 0; i < dim; ++i) {
        const uint32_t s = src[i];
        acc = (acc << sLen[s]) | sCode[s];
        have += sLen[s];
        while (have >= 8) {
            *dst++ = static_cast<uint8_t>(acc >> (have - 8));
            have -= 8;
        }
    }
    if (have) {
        *dst = static_cast<uint8_t>(acc << (8 - have));
    }
}

template <class T>
struct DeviceBuffer {
    T* ptr = nullptr;
    explicit DeviceBuffer(size_t count) { ENC_CHECK(cudaMalloc(reinterpret_cast<void**>(&ptr), std::max<size_t>(count, 1) * sizeof(T))); }
    ~DeviceBuffer() { cudaFree(ptr); }
    DeviceBuffer(const DeviceBuffer&) = delete;
    DeviceBuffer& operator=(const DeviceBuffer&) = delete;
};

}  // namespace

int deviceEncoderCount()
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return count;
}

TrainedEncoding encodeTrainedOnDevice(int device, const float* values, size_t n, size_t dim, const KMeans1D& km)
{
    if (device < 0 || device >= deviceEncoderCount()) {
        throw std::runtime_error("GPU converter: no CUDA device " + std::to_string(device));
    }
    int previous = 0;
    ENC_CHECK(cudaGetDevice(&previous));
    ENC_CHECK(cudaSetDevice(device));
    struct Restore {
        int device;
        ~Restore() { cudaSetDevice(device); }
    } restore{previous};

    const uint64_t total = static_cast<uint64_t>(n) * dim;
    if (total == 0) {
        throw std::runtime_error("Cannot save an empty trained index");
    }
    if (km.splits.size() > 255) {
        throw std::runtime_error("GPU converter: more than 256 clusters");
    }
    int sms = 148;
    ENC_CHECK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));

    // ---- pass 1: symbols, counts, first occurrences ----
    DeviceBuffer<float> dValues(total);
    ENC_CHECK(cudaMemcpy(dValues.ptr, values, total * sizeof(float), cudaMemcpyHostToDevice));
    DeviceBuffer<float> dSplits(256);
    if (!km.splits.empty()) {
        ENC_CHECK(cudaMemcpy(dSplits.ptr, km.splits.data(), km.splits.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    DeviceBuffer<uint8_t> dSymbols(total);
    DeviceBuffer<unsigned long long> dCounts(256), dFirst(256);
    ENC_CHECK(cudaMemset(dCounts.ptr, 0, 256 * sizeof(unsigned long long)));
    ENC_CHECK(cudaMemset(dFirst.ptr, 0xff, 256 * sizeof(unsigned long long)));
    // spans of < 2^32 elements per block (32-bit shared counters), at least 8 blocks per SM
    const unsigned blocks = static_cast<unsigned>(std::max<uint64_t>(uint64_t(sms) * 8, (total >> 31) + 1));
    encode_predict_kernel<<<blocks, ENC_THREADS>>>(dValues.ptr, total, dSplits.ptr, static_cast<uint32_t>(km.splits.size()),
                                                   dSymbols.ptr, dCounts.ptr, dFirst.ptr);
    ENC_CHECK(cudaGetLastError());
    unsigned long long counts[256], first[256];
    ENC_CHECK(cudaMemcpy(counts, dCounts.ptr, sizeof(counts), cudaMemcpyDeviceToHost));
    ENC_CHECK(cudaMemcpy(first, dFirst.ptr, sizeof(first), cudaMemcpyDeviceToHost));

    // ---- Huffman code on the host (<= 255 symbols), in first-occurrence order like the host encoder ----
    std::vector<std::pair<uint64_t, uint8_t>> seenOrder;
    for (int s = 0; s < 256; ++s) {
        if (counts[s]) {
            seenOrder.push_back({first[s], static_cast<uint8_t>(s)});
        }
    }
    std::sort(seenOrder.begin(), seenOrder.end());
    std::vector<std::pair<uint8_t, size_t>> firstSeen;
    for (const auto& p : seenOrder) {
        firstSeen.push_back({p.second, static_cast<size_t>(counts[p.second])});
    }
    TrainedEncoding enc;
    enc.code = buildHuffmanCode(firstSeen);
    DeviceCode hostCode;
    for (int s = 0; s < 256; ++s) {
        hostCode.codeOf[s] = enc.code.codeOf[s];
        hostCode.lengthOf[s] = enc.code.lengthOf[s];
    }
    DeviceBuffer<DeviceCode> dCode(1);
    ENC_CHECK(cudaMemcpy(dCode.ptr, &hostCode, sizeof(DeviceCode), cudaMemcpyHostToDevice));

    // ---- pass 2: row sizes; offsets by a host scan (n entries) ----
    DeviceBuffer<uint32_t> dRowBytes(n);
    encode_row_bytes_kernel<<<static_cast<unsigned>((uint64_t(n) * 32 + ENC_THREADS - 1) / ENC_THREADS), ENC_THREADS>>>(
        dSymbols.ptr, n, static_cast<uint32_t>(dim), dCode.ptr, dRowBytes.ptr);
    ENC_CHECK(cudaGetLastError());
    std::vector<uint32_t> rowBytes(n);
    ENC_CHECK(cudaMemcpy(rowBytes.data(), dRowBytes.ptr, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    enc.rowStart.assign(n + 1, 0);
    for (size_t r = 0; r < n; ++r) {
        enc.rowStart[r + 1] = enc.rowStart[r] + rowBytes[r];
    }
    if (enc.rowStart[n] > 0xffffffffULL) {
        throw std::runtime_error("packed_values exceeds the 32-bit offsets of the format");
    }

    // ---- pass 3: pack ----
    DeviceBuffer<uint64_t> dRowStart(n + 1);
    ENC_CHECK(cudaMemcpy(dRowStart.ptr, enc.rowStart.data(), (n + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice));
    DeviceBuffer<uint8_t> dPacked(enc.rowStart[n]);
    encode_pack_kernel<<<static_cast<unsigned>((n + ENC_THREADS - 1) / ENC_THREADS), ENC_THREADS>>>(
        dSymbols.ptr, n, static_cast<uint32_t>(dim), dCode.ptr, dRowStart.ptr, dPacked.ptr);
    ENC_CHECK(cudaGetLastError());
    enc.packed.resize(enc.rowStart[n]);
    if (!enc.packed.empty()) {
        ENC_CHECK(cudaMemcpy(enc.packed.data(), dPacked.ptr, enc.packed.size(), cudaMemcpyDeviceToHost));
    }
    ENC_CHECK(cudaDeviceSynchronize());
    return enc;
}

}  // namespace memb_b200
