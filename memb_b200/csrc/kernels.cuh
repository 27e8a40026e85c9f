// sm_100a kernels of the batched embedding lookup.  Compiled only with
//   nvcc -gencode arch=compute_100a,code=sm_100a
// (no other architecture, no dispatch).  See DESIGN.md for the HBM layout and
// the roofline of every kernel.
//
//   L0 measure_rows / repack_rows   load time: exact row lengths, 16 B-aligned re-pack
//   L1 hash_build                   load time: vocabulary -> open-addressing hash
//   K0 resolve_words                words -> row ids         (src/trained_compression.cpp:113-127)
//   K1 trained_decode               ids -> fp32 rows         (src/trained_compression.cpp:129-136,
//                                                             src/huffman_table_decoder.h:102-118,
//                                                             src/bit_stream_reader.h:16-31)
//   K2 uniform_decode               ids -> fp32 rows         (src/uniform_compression.cpp:58-73)
//   K3 full_gather                  ids -> fp32 rows         (src/full_compression.cpp:41-42)
//   K4 = K1/K2/K3 with a fused ReadersUnion epilogue         (python/memb/readers_union.py:16-18,30-32)
#pragma once

#include <cstdint>
#include <cuda_runtime.h>

namespace memb_b200 {
namespace kernels {

constexpr uint64_t HASH_EMPTY = ~0ull;
constexpr int WARP = 32;

// Epilogue of a decode: how a decoded value meets the output buffer.
//   WRITE       out = v                         (Reader, concatenate)
//   ACCUMULATE  out = out + v                   (average, readers 1..R-2)
//   FINALIZE    out = (out + v) / R             (average, last reader)
// np.mean(axis=0) on fp32 is a sequential fp32 add over the readers followed by
// one fp32 divide; __fadd_rn/__fdiv_rn keep nvcc from contracting.
enum Epilogue : int { EPI_WRITE = 0, EPI_ACCUMULATE = 1, EPI_FINALIZE = 2 };

struct OutputSpec {
    float* out;          // first output row
    uint64_t rowStride;  // floats between consecutive output rows
    uint32_t colOffset;  // first column written by this reader (concatenate)
    float divisor;       // R for FINALIZE
};

// ------------------------------------------------------------------ hash ----
__device__ __forceinline__ uint64_t mix64(uint64_t h)
{
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdULL;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ULL;
    h ^= h >> 33;
    return h;
}

// FNV-1a over the bytes up to (not including) the first NUL or `limit`.
__device__ __forceinline__ uint64_t hashWord(const char* s, uint32_t limit, uint32_t& len)
{
    uint64_t h = 0xcbf29ce484222325ULL;
    uint32_t i = 0;
    for (; i < limit; ++i) {
        const unsigned char c = static_cast<unsigned char>(s[i]);
        if (c == 0) {
            break;
        }
        h = (h ^ c) * 0x100000001b3ULL;
    }
    len = i;
    return mix64(h);
}

// One thread per vocabulary row: insert (tag, row) with linear probing.
__global__ void hash_build_kernel(const char* __restrict__ words, const uint32_t* __restrict__ wordOffsets,
                                  uint32_t rows, unsigned long long* __restrict__ slots, uint32_t slotMask)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) {
        return;
    }
    uint32_t len;
    const uint64_t h = hashWord(words + wordOffsets[r], 0xffffffffu, len);
    const unsigned long long entry = (h & 0xffffffff00000000ULL) | r;
    uint32_t idx = static_cast<uint32_t>(h) & slotMask;
    while (atomicCAS(&slots[idx], HASH_EMPTY, entry) != HASH_EMPTY) {
        idx = (idx + 1) & slotMask;
    }
}

// K0.  One thread per query.  Equality is an exact byte compare against the
// stored word (never fingerprint-only), so the result is the index the
// reference's lower_bound+strcmp would accept, or -1.
__global__ void resolve_words_kernel(const char* __restrict__ queryBytes,
                                     const uint64_t* __restrict__ queryOffsets, uint64_t n,
                                     const char* __restrict__ words,
                                     const uint32_t* __restrict__ wordOffsets,
                                     const unsigned long long* __restrict__ slots, uint32_t slotMask,
                                     int32_t* __restrict__ ids)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n) {
        return;
    }
    const char* q = queryBytes + queryOffsets[i];
    const uint64_t span = queryOffsets[i + 1] - queryOffsets[i];
    uint32_t len;
    const uint64_t h = hashWord(q, span > 0xffffffffULL ? 0xffffffffu : static_cast<uint32_t>(span), len);
    const uint32_t tag = static_cast<uint32_t>(h >> 32);
    uint32_t idx = static_cast<uint32_t>(h) & slotMask;
    int32_t found = -1;
    while (true) {
        const unsigned long long slot = slots[idx];
        if (slot == HASH_EMPTY) {
            break;
        }
        if (static_cast<uint32_t>(slot >> 32) == tag) {
            const uint32_t row = static_cast<uint32_t>(slot);
            const char* w = words + wordOffsets[row];
            uint32_t k = 0;
            while (k < len && w[k] == q[k]) {
                ++k;
            }
            if (k == len && w[len] == 0) {
                found = static_cast<int32_t>(row);
                break;
            }
        }
        idx = (idx + 1) & slotMask;
    }
    ids[i] = found;
}

// ----------------------------------------------------- trained: load time ----
// Device description of the prefix code, canonical form:
// for len = 1..16: firstCode[len] (left-aligned to 16 bits), count[len], and the
// index of the first key of that length.  Used by the slow path and by the
// load-time row measuring.
struct CanonicalCode {
    uint16_t firstCode[17];   // canonical code of the first key of this length (right-aligned)
    uint16_t count[17];       // number of keys with this length
    uint16_t firstIndex[17];  // index into keys[] of the first key of this length
    uint8_t keys[256];
    uint32_t minLen, maxLen;
};

// Bit-serial canonical decode of one symbol from a left-aligned 32-bit window
// that holds at least 16 valid bits.  Returns key index, sets len.
__device__ __forceinline__ uint32_t canonicalDecode(const CanonicalCode& c, uint32_t window, uint32_t& len)
{
    for (uint32_t l = c.minLen; l <= c.maxLen; ++l) {
        const uint32_t v = l ? (window >> (32 - l)) : 0;
        const uint32_t rel = v - c.firstCode[l];
        if (rel < c.count[l]) {
            len = l;
            return c.firstIndex[l] + rel;
        }
    }
    len = c.maxLen;  // unreachable for a complete code
    return 0;
}

// Big-endian 32-bit window starting at byte `pos` of the file's packed_values,
// bytes at or beyond `size` read as zero (src/bit_stream_reader.h:20-26).
__device__ __forceinline__ uint32_t fileWindow(const uint8_t* __restrict__ data, uint64_t size, uint64_t pos)
{
    uint32_t w = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        w = (w << 8) | (pos + k < size ? data[pos + k] : 0u);
    }
    return w;
}

// L0a.  One thread per row: walk the row's `dim` codes in the file's own layout
// and record how many bytes the row really occupies.  The format does not store
// row lengths; measuring them by decoding is exact for any file the
// reference can read (shared, overlapping or truncated rows included).
__global__ void measure_rows_kernel(const uint8_t* __restrict__ packed, uint64_t packedSize,
                                    const uint32_t* __restrict__ valueOffsets, uint32_t rows, uint32_t dim,
                                    const CanonicalCode* __restrict__ codePtr, uint32_t* __restrict__ rowBytes)
{
    __shared__ CanonicalCode code;
    for (uint32_t i = threadIdx.x; i < sizeof(CanonicalCode) / 4; i += blockDim.x) {
        reinterpret_cast<uint32_t*>(&code)[i] = reinterpret_cast<const uint32_t*>(codePtr)[i];
    }
    __syncthreads();
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) {
        return;
    }
    const uint64_t base = valueOffsets[r];
    uint64_t bit = 0;
    for (uint32_t i = 0; i < dim; ++i) {
        const uint64_t byte = base + (bit >> 3);
        const uint32_t sh = static_cast<uint32_t>(bit & 7);
        uint32_t w = fileWindow(packed, packedSize, byte) << sh;
        if (sh) {
            w |= (byte + 4 < packedSize ? packed[byte + 4] : 0u) >> (8 - sh);
        }
        uint32_t len;
        canonicalDecode(code, w, len);
        bit += len;
    }
    rowBytes[r] = static_cast<uint32_t>((bit + 7) >> 3);
}

// L0b.  Copy every row to its 16 B-aligned home in row-id order.  Words are
// stored byte-swapped so that a plain 32-bit shared-memory load yields the
// MSB-first bit order of src/bit_stream.h:18-34 without a PRMT in the hot loop.
// One warp per row; tail bytes (and bytes past the end of the file) are zero.
__global__ void repack_rows_kernel(const uint8_t* __restrict__ packed, uint64_t packedSize,
                                   const uint32_t* __restrict__ valueOffsets,
                                   const uint32_t* __restrict__ rowBytes,
                                   const uint32_t* __restrict__ rowOff16, uint32_t rows,
                                   uint32_t* __restrict__ blob)
{
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) / WARP;
    const uint32_t lane = threadIdx.x % WARP;
    if (warp >= rows) {
        return;
    }
    const uint64_t src = valueOffsets[warp];
    const uint32_t bytes = rowBytes[warp];
    const uint32_t words = (rowOff16[warp + 1] - rowOff16[warp]) * 4;
    uint32_t* dst = blob + static_cast<uint64_t>(rowOff16[warp]) * 4;
    for (uint32_t w = lane; w < words; w += WARP) {
        uint32_t v = 0;
#pragma unroll
        for (uint32_t k = 0; k < 4; ++k) {
            const uint32_t b = w * 4 + k;
            const uint32_t byte = (b < bytes && src + b < packedSize) ? packed[src + b] : 0u;
            v = (v << 8) | byte;
        }
        dst[w] = v;
    }
}

__global__ void pad_rows16_kernel(const uint32_t* __restrict__ rowBytes, uint32_t rows,
                                  uint32_t* __restrict__ rowUnits)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < rows) {
        rowUnits[r] = (rowBytes[r] + 15) / 16;
    }
}

// ------------------------------------------------------------ K1 trained ----
// HBM layout (built once at load):
//   blob      rows in row-id order, each padded to 16 B, 32-bit words byte-swapped
//   rowOff16  rows+1 offsets into blob in 16 B units
//   lut       2^lutBits entries, entry = (sym*4) << 16 | escape << 15 | len
//             optionally replicated 32x in shared memory so lane l only ever
//             touches bank l
//   codebook  centroids, fp32
//
// Work decomposition: one warp owns a tile of 32 consecutive OUTPUT rows; lane l
// decodes row l of the tile serially (rows are independent byte-aligned streams,
// SURVEY D1), all lanes in lock step on the symbol index.  Compressed rows are
// staged into shared memory with one cp.async.bulk per row completing on a
// per-warp mbarrier; decoded values go through a swizzled 32x32 fp32 transpose
// tile so that every global store instruction writes 4 x 128 contiguous bytes.
struct TrainedParams {
    const uint4* blob;
    const uint32_t* rowOff16;
    const uint32_t* lut;
    const float* codebook;
    const CanonicalCode* code;
    uint32_t rows;
    uint32_t dim;
    uint32_t lutBits;       // L
    uint32_t lutReplicas;   // 1 or 32
    uint32_t lutShift;      // window -> byte offset:  (hi >> lutShift) & lutMask
    uint32_t lutMask;
    uint32_t codebookSize;
    uint32_t slotBytes;     // shared-memory bytes reserved per staged row (multiple of 16)
};

__device__ __forceinline__ uint32_t smemAddr(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MEMB_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MEMB_DONE;\n"
        "bra MEMB_WAIT;\n"
        "MEMB_DONE:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP); src, dst and
// bytes are multiples of 16.
__device__ __forceinline__ void bulkCopyG2S(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

struct BitWindow {
    uint32_t hi, lo;    // 64 valid bits, MSB first
    uint32_t consumed;  // low 16 bits: bits consumed in this row so far (junk above)
};

// Slow path for a window whose code is longer than the LUT width.
__device__ __noinline__ uint32_t escapeDecode(const CanonicalCode* __restrict__ code, uint32_t window)
{
    for (uint32_t l = code->minLen; l <= code->maxLen; ++l) {
        const uint32_t rel = (l ? (window >> (32 - l)) : 0u) - code->firstCode[l];
        if (rel < code->count[l]) {
            return (static_cast<uint32_t>(code->keys[code->firstIndex[l] + rel]) << 18) | l;
        }
    }
    return 0;
}

template <bool ESCAPE>
__device__ __forceinline__ float decodeSymbol(BitWindow& w, const unsigned char* lutLane, uint32_t lutShift,
                                              uint32_t lutMask, const unsigned char* codebook,
                                              const CanonicalCode* __restrict__ code)
{
    uint32_t e = *reinterpret_cast<const uint32_t*>(lutLane + ((w.hi >> lutShift) & lutMask));
    if (ESCAPE) {
        if (e & 0x8000u) {
            e = escapeDecode(code, w.hi);
        }
    }
    const float v = *reinterpret_cast<const float*>(codebook + (e >> 16));
    w.hi = __funnelshift_l(w.lo, w.hi, e);  // shift = e & 31 = code length
    w.lo = __funnelshift_l(0u, w.lo, e);
    w.consumed += e;
    return v;
}

// Re-centre the 64-bit window on the row's current bit position: three aligned
// 32-bit loads and two funnel shifts, no data-dependent branch.
__device__ __forceinline__ void reloadWindow(BitWindow& w, const unsigned char* rowBase)
{
    const uint32_t* p = reinterpret_cast<const uint32_t*>(rowBase + ((w.consumed >> 3) & 0x1ffcu));
    const uint32_t w0 = p[0], w1 = p[1], w2 = p[2];
    w.hi = __funnelshift_l(w1, w0, w.consumed);
    w.lo = __funnelshift_l(w2, w1, w.consumed);
}

// Float4 slot of (row, quad) in the swizzled 32x32 fp32 tile.
__device__ __forceinline__ float4* tileQuad(unsigned char* tile, uint32_t row, uint32_t quad)
{
    return reinterpret_cast<float4*>(tile + row * 128 + ((quad ^ (row & 7)) << 4));
}

__device__ __forceinline__ float applyEpilogue(float v, float old, int epi, float divisor)
{
    if (epi == EPI_WRITE) {
        return v;
    }
    const float s = __fadd_rn(old, v);
    return epi == EPI_ACCUMULATE ? s : __fdiv_rn(s, divisor);
}

constexpr uint32_t TRAINED_TILE_BYTES = 32 * 32 * 4;
constexpr uint32_t TRAINED_MAX_DIM = 4095;  // consumed bits live in 16 bits: dim * 16 < 65536
constexpr uint32_t TRAINED_MAX_WARPS = 24;

__host__ __device__ inline uint32_t trainedLutBytes(uint32_t lutBits, uint32_t replicas)
{
    return (4u << lutBits) * replicas;
}
__host__ __device__ inline uint32_t trainedCodebookBytes(uint32_t codebookSize)
{
    return (codebookSize * 4 + 127) & ~127u;
}
__host__ __device__ inline uint32_t trainedWarpBytes(uint32_t slotBytes)
{
    return WARP * slotBytes + TRAINED_TILE_BYTES + 128;
}

// G = symbols decoded per window reload (4 for codes up to 16 bits, 8 for codes
// up to 8 bits).  VEC4: dim, stride and column offset are multiples of 4 and the
// output is 16 B aligned -> 128-bit shared loads and streaming 128-bit stores.
template <int G, bool VEC4, bool ESCAPE, int EPI>
__global__ void __launch_bounds__(TRAINED_MAX_WARPS * WARP, 1)
trained_decode_kernel(const TrainedParams p, const int32_t* __restrict__ ids, uint32_t rowBegin, uint64_t n,
                      const OutputSpec o)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t lane = threadIdx.x % WARP;
    const uint32_t warp = threadIdx.x / WARP;
    const uint32_t warpsPerCta = blockDim.x / WARP;

    // shared layout: [lut][codebook][per warp: slots | tile | mbarrier]
    const uint32_t lutBytes = trainedLutBytes(p.lutBits, p.lutReplicas);
    unsigned char* const codebook = smem + lutBytes;
    unsigned char* const slots = codebook + trainedCodebookBytes(p.codebookSize) + warp * trainedWarpBytes(p.slotBytes);
    unsigned char* const tile = slots + WARP * p.slotBytes;
    const uint32_t bar = smemAddr(tile + TRAINED_TILE_BYTES);

    {
        uint32_t* lutS = reinterpret_cast<uint32_t*>(smem);
        const uint32_t entries = 1u << p.lutBits;
        if (p.lutReplicas == 1) {
            for (uint32_t i = threadIdx.x; i < entries; i += blockDim.x) {
                lutS[i] = p.lut[i];
            }
        } else {
            for (uint32_t i = threadIdx.x; i < entries * WARP; i += blockDim.x) {
                lutS[i] = p.lut[i / WARP];
            }
        }
        float* cbS = reinterpret_cast<float*>(codebook);
        for (uint32_t i = threadIdx.x; i < p.codebookSize; i += blockDim.x) {
            cbS[i] = p.codebook[i];
        }
        if (lane == 0) {
            mbarInit(bar, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const unsigned char* const lutLane = smem + (p.lutReplicas == 1 ? 0u : lane * 4u);
    const uint32_t lutShift = p.lutShift, lutMask = p.lutMask;
    const uint64_t numTiles = (n + WARP - 1) / WARP;
    const uint64_t tileStride = static_cast<uint64_t>(gridDim.x) * warpsPerCta;
    unsigned char* const mySlot = slots + lane * p.slotBytes;
    const uint32_t dim = p.dim;
    uint32_t parity = 0;

    // Stage the 32 rows of tile t; returns the lane's "row present" flag.
    auto stage = [&](uint64_t t) -> bool {
        const uint64_t i = t * WARP + lane;
        int64_t row = -1;
        if (i < n) {
            row = ids ? static_cast<int64_t>(ids[i]) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
        }
        const bool present = row >= 0 && row < static_cast<int64_t>(p.rows);
        uint32_t o16 = 0, bytes = 0;
        if (present) {
            o16 = p.rowOff16[row];
            bytes = (p.rowOff16[row + 1] - o16) * 16;
        }
        uint32_t total = bytes;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            total += __shfl_xor_sync(0xffffffffu, total, s);
        }
        if (lane == 0) {
            mbarExpectTx(bar, total);
        }
        __syncwarp();
        if (bytes) {
            bulkCopyG2S(smemAddr(mySlot), p.blob + o16, bytes, bar);
        }
        return present;
    };

    uint64_t t = static_cast<uint64_t>(blockIdx.x) * warpsPerCta + warp;
    bool present = false;
    if (t < numTiles) {
        present = stage(t);
    }
    for (; t < numTiles; t += tileStride) {
        const uint32_t presentMask = __ballot_sync(0xffffffffu, present);
        // lanes without a row decode a copy of some present row so that their
        // bit position stays inside a staged slot; their output is forced to 0.
        const uint32_t donor = presentMask ? (__ffs(presentMask) - 1) : lane;
        const unsigned char* const rowBase = present ? mySlot : slots + donor * p.slotBytes;
        mbarWait(bar, parity);
        parity ^= 1;

        BitWindow w;
        w.hi = w.lo = 0;
        w.consumed = 0;
        const uint64_t row0 = t * WARP;
        const uint32_t rowsHere = static_cast<uint32_t>(min(static_cast<uint64_t>(WARP), n - row0));
        bool nextPresent = false;

        for (uint32_t c0 = 0; c0 < dim; c0 += 32) {
            const uint32_t nsym = min(32u, dim - c0);
            if (presentMask) {
                if (nsym == 32) {
#pragma unroll
                    for (int g = 0; g < 32 / G; ++g) {
                        reloadWindow(w, rowBase);
#pragma unroll
                        for (int h = 0; h < G / 4; ++h) {
                            float4 v;
                            v.x = decodeSymbol<ESCAPE>(w, lutLane, lutShift, lutMask, codebook, p.code);
                            v.y = decodeSymbol<ESCAPE>(w, lutLane, lutShift, lutMask, codebook, p.code);
                            v.z = decodeSymbol<ESCAPE>(w, lutLane, lutShift, lutMask, codebook, p.code);
                            v.w = decodeSymbol<ESCAPE>(w, lutLane, lutShift, lutMask, codebook, p.code);
                            *tileQuad(tile, lane, g * (G / 4) + h) = v;
                        }
                    }
                } else {
                    for (uint32_t s = 0; s < nsym; s += 4) {
                        reloadWindow(w, rowBase);
                        float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                        for (uint32_t k = 0; k < 4; ++k) {
                            if (s + k < nsym) {
                                v[k] = decodeSymbol<ESCAPE>(w, lutLane, lutShift, lutMask, codebook, p.code);
                            }
                        }
                        *tileQuad(tile, lane, s / 4) = make_float4(v[0], v[1], v[2], v[3]);
                    }
                }
            }
            __syncwarp();
            // all compressed bytes of this tile are consumed once the last chunk
            // is decoded: start staging the next tile under the final flush
            if (c0 + 32 >= dim && t + tileStride < numTiles) {
                nextPresent = stage(t + tileStride);
            }
            if (VEC4) {
                const uint32_t quads = nsym / 4;
                const uint32_t q = lane & 7;
#pragma unroll
                for (uint32_t it = 0; it < 8; ++it) {
                    const uint32_t r = it * 4 + (lane >> 3);
                    if (q < quads && r < rowsHere) {
                        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                        if ((presentMask >> r) & 1u) {
                            v = *tileQuad(tile, r, q);
                        }
                        float4* dst = reinterpret_cast<float4*>(o.out + (row0 + r) * o.rowStride + o.colOffset + c0) + q;
                        if (EPI != EPI_WRITE) {
                            const float4 old = *dst;
                            v.x = applyEpilogue(v.x, old.x, EPI, o.divisor);
                            v.y = applyEpilogue(v.y, old.y, EPI, o.divisor);
                            v.z = applyEpilogue(v.z, old.z, EPI, o.divisor);
                            v.w = applyEpilogue(v.w, old.w, EPI, o.divisor);
                        }
                        __stcs(dst, v);
                    }
                }
            } else {
                for (uint32_t r = 0; r < rowsHere; ++r) {
                    if (lane < nsym) {
                        float v = 0.f;
                        if ((presentMask >> r) & 1u) {
                            v = reinterpret_cast<const float*>(tileQuad(tile, r, lane >> 2))[lane & 3];
                        }
                        float* dst = o.out + (row0 + r) * o.rowStride + o.colOffset + c0 + lane;
                        if (EPI != EPI_WRITE) {
                            v = applyEpilogue(v, *dst, EPI, o.divisor);
                        }
                        __stcs(dst, v);
                    }
                }
            }
            __syncwarp();
        }
        present = nextPresent;
    }
}

// Generic trained decode for shapes the tiled kernel does not take (dim > 4095
// or a row too long to stage 32 of them in shared memory): one thread per row,
// straight from HBM, scalar stores.  Correct for everything, fast for nothing.
template <int EPI>
__global__ void trained_decode_generic_kernel(const TrainedParams p, const int32_t* __restrict__ ids,
                                              uint32_t rowBegin, uint64_t n, const OutputSpec o)
{
    const uint64_t i = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    if (i >= n) {
        return;
    }
    const int64_t row = ids ? static_cast<int64_t>(ids[i]) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
    const bool present = row >= 0 && row < static_cast<int64_t>(p.rows);
    float* dst = o.out + i * o.rowStride + o.colOffset;
    const uint32_t* words = present ? reinterpret_cast<const uint32_t*>(p.blob + p.rowOff16[row]) : nullptr;
    const uint64_t limitWords =
        present ? static_cast<uint64_t>(p.rowOff16[row + 1] - p.rowOff16[row]) * 4 : 0;
    uint64_t bit = 0;
    for (uint32_t c = 0; c < p.dim; ++c) {
        float v = 0.f;
        if (present) {
            const uint64_t wi = bit >> 5;
            const uint32_t sh = static_cast<uint32_t>(bit & 31);
            const uint32_t w0 = wi < limitWords ? words[wi] : 0u;
            const uint32_t w1 = wi + 1 < limitWords ? words[wi + 1] : 0u;
            const uint32_t window = __funnelshift_l(w1, w0, sh);
            uint32_t len;
            const uint32_t idx = canonicalDecode(*p.code, window, len);
            bit += len;
            v = p.codebook[p.code->keys[idx]];
        }
        if (EPI != EPI_WRITE) {
            v = applyEpilogue(v, dst[c], EPI, o.divisor);
        }
        dst[c] = v;
    }
}

// ------------------------------------------------------------ K2 uniform ----
// values: rows x dim bytes (dense, flattened at load from the per-word tables),
// minmax: rows x (min, max).  out[i] = min + (max - min) * float(v) / levels,
// evaluated sub -> mul -> div -> add in binary32 round-to-nearest exactly like
// src/uniform_compression.cpp:68-72 (no FMA contraction, no fast division).
struct UniformParams {
    const uint8_t* values;
    const float2* minmax;
    uint32_t rows;
    uint32_t dim;
    float levels;
};

__device__ __forceinline__ float uniformDequant(float mn, float range, uint32_t v, float levels)
{
    return __fadd_rn(mn, __fdiv_rn(__fmul_rn(range, static_cast<float>(v)), levels));
}

// One thread per output float4 (VEC4) or per output float; output elements are
// numbered row-major so every warp store is 512 contiguous bytes.
template <bool VEC4, int EPI>
__global__ void uniform_decode_kernel(const UniformParams p, const int32_t* __restrict__ ids, uint32_t rowBegin,
                                      uint64_t n, const OutputSpec o)
{
    const uint32_t perRow = VEC4 ? p.dim / 4 : p.dim;
    const uint64_t total = n * perRow;
    for (uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; e < total;
         e += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint64_t i = e / perRow;
        const uint32_t c = static_cast<uint32_t>(e - i * perRow);
        const int64_t row = ids ? static_cast<int64_t>(ids[i]) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
        const bool present = row >= 0 && row < static_cast<int64_t>(p.rows);
        if (VEC4) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (present) {
                const uint32_t packed = reinterpret_cast<const uint32_t*>(p.values + static_cast<uint64_t>(row) * p.dim)[c];
                const float2 mm = p.minmax[row];
                const float range = __fsub_rn(mm.y, mm.x);
                v.x = uniformDequant(mm.x, range, packed & 0xff, p.levels);
                v.y = uniformDequant(mm.x, range, (packed >> 8) & 0xff, p.levels);
                v.z = uniformDequant(mm.x, range, (packed >> 16) & 0xff, p.levels);
                v.w = uniformDequant(mm.x, range, packed >> 24, p.levels);
            }
            float4* dst = reinterpret_cast<float4*>(o.out + i * o.rowStride + o.colOffset) + c;
            if (EPI != EPI_WRITE) {
                const float4 old = *dst;
                v.x = applyEpilogue(v.x, old.x, EPI, o.divisor);
                v.y = applyEpilogue(v.y, old.y, EPI, o.divisor);
                v.z = applyEpilogue(v.z, old.z, EPI, o.divisor);
                v.w = applyEpilogue(v.w, old.w, EPI, o.divisor);
            }
            __stcs(dst, v);
        } else {
            float v = 0.f;
            if (present) {
                const float2 mm = p.minmax[row];
                v = uniformDequant(mm.x, __fsub_rn(mm.y, mm.x), p.values[static_cast<uint64_t>(row) * p.dim + c],
                                   p.levels);
            }
            float* dst = o.out + i * o.rowStride + o.colOffset + c;
            if (EPI != EPI_WRITE) {
                v = applyEpilogue(v, *dst, EPI, o.divisor);
            }
            __stcs(dst, v);
        }
    }
}

// --------------------------------------------------------------- K3 full ----
struct FullParams {
    const float* values;  // rows x dim
    uint32_t rows;
    uint32_t dim;
};

template <bool VEC4, int EPI>
__global__ void full_gather_kernel(const FullParams p, const int32_t* __restrict__ ids, uint32_t rowBegin,
                                   uint64_t n, const OutputSpec o)
{
    const uint32_t perRow = VEC4 ? p.dim / 4 : p.dim;
    const uint64_t total = n * perRow;
    for (uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; e < total;
         e += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint64_t i = e / perRow;
        const uint32_t c = static_cast<uint32_t>(e - i * perRow);
        const int64_t row = ids ? static_cast<int64_t>(ids[i]) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
        const bool present = row >= 0 && row < static_cast<int64_t>(p.rows);
        if (VEC4) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (present) {
                v = reinterpret_cast<const float4*>(p.values + static_cast<uint64_t>(row) * p.dim)[c];
            }
            float4* dst = reinterpret_cast<float4*>(o.out + i * o.rowStride + o.colOffset) + c;
            if (EPI != EPI_WRITE) {
                const float4 old = *dst;
                v.x = applyEpilogue(v.x, old.x, EPI, o.divisor);
                v.y = applyEpilogue(v.y, old.y, EPI, o.divisor);
                v.z = applyEpilogue(v.z, old.z, EPI, o.divisor);
                v.w = applyEpilogue(v.w, old.w, EPI, o.divisor);
            }
            __stcs(dst, v);
        } else {
            float v = present ? p.values[static_cast<uint64_t>(row) * p.dim + c] : 0.f;
            float* dst = o.out + i * o.rowStride + o.colOffset + c;
            if (EPI != EPI_WRITE) {
                v = applyEpilogue(v, *dst, EPI, o.divisor);
            }
            __stcs(dst, v);
        }
    }
}

}  // namespace kernels
}  // namespace memb_b200
