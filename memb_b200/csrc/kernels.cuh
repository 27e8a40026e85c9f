// sm_100a kernels of the batched embedding lookup.  Compiled only with
//   nvcc -gencode arch=compute_100a,code=sm_100a
// (no other architecture, no dispatch).  See DESIGN.md for the HBM layout and
// the roofline of every kernel.
//
//   L0 measure_rows / repack_rows   load time: exact row lengths, re-pack into fixed-stride records (or 16 B-aligned rows)
//      interleave_records           load time: the records once more, interleaved per 32 rows (one bulk copy per tile)
//   L1 hash_build                   load time: vocabulary -> open-addressing hash
//   K0 resolve_words                words -> row ids         (src/trained_compression.cpp:113-127)
//   K1 trained_decode               ids -> fp32 rows         (src/trained_compression.cpp:129-136,
//                                                             src/huffman_table_decoder.h:102-118,
//                                                             src/bit_stream_reader.h:16-31)
//   K2 uniform_decode               ids -> fp32 rows         (src/uniform_compression.cpp:58-73)
//   K3 full_gather                  ids -> fp32 rows         (src/full_compression.cpp:41-42)
//   K4 = K1/K2/K3 with a fused ReadersUnion epilogue         (python/memb/readers_union.py:16-18,30-32)
// trained_rows.cuh: trained_rows_kernel, whole-row tiles + multi-symbol LUT, a union of two trained
// readers in one launch (measured slower, selectable); device_builder.cu: the GPU converter's kernels.
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
#ifndef MEMB_UNIT_BYTES
#define MEMB_UNIT_BYTES 16
#endif
// A re-packed row is a whole number of units; one unit = one global load of a lane.
constexpr uint32_t TRAINED_UNIT_BYTES = MEMB_UNIT_BYTES;  // 16: 128-bit loads, 32: 256-bit loads (one sector)
constexpr uint32_t TRAINED_UNIT_WORDS = TRAINED_UNIT_BYTES / 4;
static_assert(TRAINED_UNIT_BYTES == 16 || TRAINED_UNIT_BYTES == 32, "row unit is one 128- or 256-bit load");
struct alignas(TRAINED_UNIT_BYTES) RowUnit {
    uint32_t w[TRAINED_UNIT_WORDS];
};

enum Epilogue : int { EPI_WRITE = 0, EPI_ACCUMULATE = 1, EPI_FINALIZE = 2 };

struct OutputSpec {
    float* out;          // first output row
    uint64_t rowStride;  // floats between consecutive output rows
    uint32_t colOffset;  // first column written by this reader (concatenate)
    float divisor;       // R for FINALIZE
    // concatenate: a column block that begins or ends in the middle of a 32 B sector shares that
    // sector with its neighbour, which another launch writes.  A half-written sector costs a
    // read-modify-write in DRAM (scripts/partial_sector_bw.cu: 3.2 TB/s against 5.1 TB/s with
    // whole sectors).  When set, the first / last 16 B unit of row i goes to sideHead[i] /
    // sideTail[i] instead and stitch_sectors_kernel writes the shared sector whole.
    float4* sideHead = nullptr;
    float4* sideTail = nullptr;
    // average of trained readers: reader 0 does not write fp32 rows at all -- its launch leaves
    // the SYMBOL words of its rows (dim bytes per row instead of 4 * dim) in `sym`, and the
    // launch of reader 1 gathers reader 0's centroids from them while it flushes its own
    // (SYM instantiations of the trained kernel): the output is written once, 1800 instead of
    // 3600 bytes move per row.  Row i of the launch lives at sym + i * symWords.
    uint32_t* sym = nullptr;
    uint32_t symWords = 0;
    // `out` is an NVLS MULTICAST address (one store lands in the same place of every GPU's buffer:
    // the fused decode -> all-gather of a sharded full-vocabulary decode): stores are multimem.st
    uint32_t multicast = 0;
    const float* symCodebook = nullptr;  // reader 0's centroids (SYM == 2)
    uint32_t symCodebookSize = 0;
    uint32_t symFieldShift = 0;          // reader 0's symbol field -> byte offset into its codebook
    // SYM == 2 with small codebooks: the merged value depends on the symbol PAIR only, so the launch
    // builds T[s0][s1] = epilogue(c0[s0], c1[s1]) once per CTA ((size0 + 1) x pairStride entries, the
    // +0.0 slots of missing words included) and the flush gathers ONE value per element instead of two
    // centroids + add + scale.  Same device arithmetic as the two-gather flush: bit-identical.
    // pairReplicas == 32: one copy of the table per shared-memory bank (entry i of lane l at word i * 32 + l), like
    // the LUT -- the gathers of a warp never conflict (a single copy spans every bank: ncu r02 counted 43 M
    // bank-conflict wavefronts of 75 M in the average's second launch).  Needs entries * 128 B <= 64 KB.
    uint32_t pairEntries = 0;            // (size0 + 1) * pairStride logical entries; 0: two-gather flush
    uint32_t pairStride = 0;             // entries per s0 row (>= size1 + 1)
    uint32_t pairReplicas = 1;           // 1 or 32
    uint32_t straightPairs = 0;          // middle chunks take the predicate-free pair flush (option union_pair_table = 2, default)
};

// Where the 16 B unit `unit` of output row i goes (see OutputSpec::sideHead).
__device__ __forceinline__ float4* redirectUnit(const OutputSpec& o, float4* dst, uint64_t i, uint32_t unit,
                                                uint32_t lastUnit)
{
    if (o.sideHead && unit == 0) {
        return o.sideHead + i;
    }
    if (o.sideTail && unit == lastUnit) {
        return o.sideTail + i;
    }
    return dst;
}

// One shared sector per row and block boundary: lower half = last unit of the block on the
// left, upper half = first unit of the block on the right; two lanes write one whole sector.
__global__ void stitch_sectors_kernel(const float4* __restrict__ tailOfLeft, const float4* __restrict__ headOfRight,
                                      uint64_t n, float* out, uint64_t rowStride, uint32_t boundaryColumn)
{
    const uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x;
    const uint64_t i = e >> 1;
    if (i >= n) {
        return;
    }
    const bool upper = e & 1;
    const float4 v = upper ? headOfRight[i] : tailOfLeft[i];
    float4* dst = reinterpret_cast<float4*>(out + i * rowStride + boundaryColumn) + (upper ? 0 : -1);
    __stcs(dst, v);
}

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

// L0b.  Copy every row to its 16 B-aligned home in row-id order, re-ordering the
// bits for the decode kernel: the file packs codes MSB-first into bytes
// (src/bit_stream.h:18-34); the blob holds the same bit sequence LSB-first in
// 32-bit words and shifted up by padBits, i.e. stream bit k is bit (k + padBits) % 32
// of word (k + padBits) / 32.  The decoder then indexes its LUT with `window & mask`
// (no shift) and advances with one funnel shift.  One warp per row; tail bits (and
// bytes past the end of the file, src/bit_stream_reader.h:22-25) are zero.
__global__ void repack_rows_kernel(const uint8_t* __restrict__ packed, uint64_t packedSize,
                                   const uint32_t* __restrict__ valueOffsets,
                                   const uint32_t* __restrict__ rowBytes,
                                   const uint32_t* __restrict__ rowOff16, uint32_t recUnits, uint32_t rows,
                                   uint32_t padBits, uint32_t* __restrict__ blob)
{
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) / WARP;
    const uint32_t lane = threadIdx.x % WARP;
    if (warp >= rows) {
        return;
    }
    const uint64_t src = valueOffsets[warp];
    const uint32_t bytes = rowBytes[warp];
    // recUnits != 0: fixed-stride records (TrainedParams::recUnits), else the offset table
    const uint32_t words = (recUnits ? recUnits : rowOff16[warp + 1] - rowOff16[warp]) * TRAINED_UNIT_WORDS;
    uint32_t* dst = blob + (recUnits ? static_cast<uint64_t>(warp) * recUnits : static_cast<uint64_t>(rowOff16[warp])) * TRAINED_UNIT_WORDS;
    // big-endian word `wi` of the row as stored in the file (zero outside the row)
    auto fileWord = [&](int64_t wi) -> uint32_t {
        uint32_t v = 0;
        if (wi >= 0) {
#pragma unroll
            for (uint32_t k = 0; k < 4; ++k) {
                const uint64_t b = static_cast<uint64_t>(wi) * 4 + k;
                const uint32_t byte = (b < bytes && src + b < packedSize) ? packed[src + b] : 0u;
                v = (v << 8) | byte;
            }
        }
        return v;
    };
    for (uint32_t w = lane; w < words; w += WARP) {
        // 32 stream bits starting at stream bit 32*w - padBits, MSB-first, then bit-reversed
        const uint32_t cur = fileWord(w);
        const uint32_t msbFirst = padBits ? __funnelshift_l(cur, fileWord(static_cast<int64_t>(w) - 1), 32 - padBits) : cur;
        dst[w] = __brev(msbFirst);
    }
}

// L0c.  Fixed-stride records once more, interleaved per group of 32 rows (TrainedParams::blobTiled):
// output word o = (g * recWords + w) * 32 + l holds word w of row 32 g + l (zero past the last row).
__global__ void interleave_records_kernel(const uint32_t* __restrict__ blob, uint32_t rows, uint32_t recWords,
                                          uint32_t* __restrict__ tiled, uint64_t totalWords)
{
    for (uint64_t o = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; o < totalWords;
         o += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint64_t group = o / (recWords * 32u);
        const uint32_t rem = static_cast<uint32_t>(o - group * (recWords * 32u));
        const uint64_t row = group * 32u + (rem & 31u);
        tiled[o] = row < rows ? blob[row * recWords + (rem >> 5)] : 0u;
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
//   blob      rows in row-id order, each padded to 16 B.  Bits are stored LSB-first:
//             stream bit k of a row (k-th bit in the MSB-first order of
//             src/bit_stream.h:18-34) is bit (k + padBits) % 32 of word (k + padBits) / 32
//   rowOff16  rows+1 offsets into blob in 16 B units
//   lut       2^lutBits entries indexed by the next lutBits stream bits, first bit in
//             bit 0 (i.e. the bit-reversed code prefix);
//             entry = (sym << symShift) << 16 | escape << 15 | len
//   codebook  centroids, fp32
//
// Work decomposition: one warp owns a tile of 32 consecutive OUTPUT rows; lane l
// decodes row l of the tile serially (rows are independent byte-aligned streams,
// SURVEY D1), all lanes in lock step on the symbol index.
//
// The per-symbol dependency chain is what bounds this kernel (ncu r01: 35 % of
// the stall samples are the LDS of the LUT lookup, 22 % fixed-latency ALU), so
// the layout is chosen to make that chain LOP3 -> LDS -> SHF:
//   * LSB-first bits: the LUT index is `window & mask`, no shift; the window
//     advances with one funnel shift right;
//   * padBits (7 for the bank-replicated LUT, 2 for a single copy) parks the
//     window so that `window & mask` already is the byte offset of the entry;
//   * the LUT sits at a shared-memory address aligned to its own size, so the
//     base and the lane's replica are OR-ed in by the same LOP3.
// Shared memory is the scarce resource, so every structure is laid out so that
// lane l only ever touches bank l:
//   codebook  1 KiB, 1 KiB-aligned   fp32 centroids (address = base | sym * 4)
//   lut       aligned to its size    entry i of replica l at i*128 + l*4 (32 replicas),
//                                    or one plain copy when the code is too long
//   per warp  rows[w][l]             word w of lane l's compressed row at (w*32+l)*4
//             tile[r]                ring of the last 128 decoded SYMBOLS of row r, 132 B stride
// Warp buffers fill the space below and above the LUT.  Compressed rows reach
// shared memory through registers: each lane pulls its own row with 128-bit
// loads (a tile ahead) and scatters the words down its bank.  The decode loop
// keeps a 64-bit window in two registers, re-centred from shared memory every G
// symbols (G*maxLen <= 64 - padBits), looks a symbol up with one conflict-free
// LDS and packs four symbols into one word of the tile.  The fp32 gather from the
// codebook happens on the way OUT of the transpose tile, so the serial
// per-symbol chain carries no float and every global store is a 128-bit
// streaming store covering 128 contiguous bytes per 8 lanes.
struct TrainedParams {
    const RowUnit* blob;
    const RowUnit* blobTiled;   // fixed-stride records once more, interleaved per group of 32 rows: word w of row
                                // 32 g + l at word (g * recUnits * 4 + w) * 32 + l -- the staged layout itself, so a
                                // range decode stages a whole tile with ONE bulk copy (nullptr: not built)
    const uint32_t* rowOff16;   // (name from the 16 B layout) rows+1 offsets into blob in UNITS; unused when recUnits != 0
    uint32_t recUnits;          // != 0: fixed-stride records, row r at blob + r * recUnits (no offset table)
    const uint32_t* lut;
    const uint32_t* lut2;       // second level: entries of codes longer than the LUT is wide (lut2Entries of them)
    uint32_t lut2Entries;
    const float* codebook;
    const CanonicalCode* code;
    uint32_t rows;
    uint32_t dim;
    uint32_t lutBits;       // L
    uint32_t lutReplicas;   // 1 or 32
    uint32_t padBits;       // 7 (32 replicas) or 2 (single copy)
    uint32_t lutMask;       // ((1 << L) - 1) << padBits
    uint32_t symShift;      // 2 when the codebook has <= 64 entries (symbol stored as byte offset), else 0
    uint32_t codebookSize;
    uint32_t rowWords;      // words reserved per staged row (max row units * 4 + 4)
    uint32_t hasZeroSlot;   // the shared codebook has a slot past the last centroid (it holds +0.0) ...
    uint32_t zeroWord;      // ... and this is its symbol field, four times: what a lane WITHOUT a row writes to the tile
};

constexpr uint32_t TRAINED_CODEBOOK_BYTES = 2048;  // 256 centroids, twice (the second copy: OutputSpec::symCodebook)
constexpr uint32_t TRAINED_CHUNK = 32;             // symbols decoded between two flushes
constexpr uint32_t TRAINED_CHUNK_UNITS = TRAINED_CHUNK / 4;       // 16 B output units (= tile words) per chunk and row
constexpr uint32_t TRAINED_RING_UNITS = 2 * TRAINED_CHUNK_UNITS;  // the tile keeps the last two chunks of every row
constexpr uint32_t TRAINED_FLUSH_ROWS = 32 / TRAINED_CHUNK_UNITS;  // rows covered by one store instruction
// Symbol tile of a warp: 32 rows x ring of 16 words.  Word w of row r lives at word
//   (r >> 2) * 65 + (w >> 3) * 32 + (r & 3) * 8 + (w & 7)
// so that (a) the decode, where lane l writes the same word w of row l, hits bank
// (l >> 2) + 8 * (l & 3) + (w & 7): 32 distinct banks; and (b) the flush, where 8 lanes read
// 8 consecutive ring words of each of the rows 4*it .. 4*it+3 starting anywhere, hits
// banks it + 8 * (r & 3) + {0..7}: 32 distinct banks whatever the rows' start offsets
// (the skewed-stride layout of the first versions lost 13 % of its shared-memory
// wavefronts to conflicts there, ncu r01).
constexpr uint32_t TRAINED_TILE_GROUP_BYTES = 65 * 4;  // four rows (+ one word of skew)
static_assert(TRAINED_CHUNK_UNITS == 8 && TRAINED_FLUSH_ROWS == 4, "tile layout is written for 32-symbol chunks");
constexpr uint32_t TRAINED_TILE_BYTES = 17 * 128;  // 8 groups of 260 B, rounded up to 128 B
constexpr uint32_t TRAINED_MAX_DIM = 4095;  // consumed bits live in 16 bits: dim * 16 < 65536
constexpr uint32_t TRAINED_MAX_WARPS = 10;  // measured optimum (sweep r01: 8 -> 2.98, 10 -> 3.22, 12 -> 3.10 G rows/s) ...
constexpr uint32_t TRAINED_MAX_REGS = 168;  // ... up to three warps on a scheduler, which owns 16384 registers: 3 x 32 x 168 = 16128
#ifndef MEMB_PREFETCH_BYTES
#define MEMB_PREFETCH_BYTES 192
#endif
constexpr uint32_t TRAINED_PREFETCH_UNITS = MEMB_PREFETCH_BYTES / MEMB_UNIT_BYTES;  // bytes of the next row per lane held in registers

__host__ __device__ inline uint32_t trainedLutBytes(uint32_t lutBits, uint32_t replicas)
{
    return (4u << lutBits) * replicas;
}
// per warp: staged rows (bank-per-lane), symbol tile, and -- bulk-async staging (TMA) -- the
// landing buffer the warp's 32 records are copied into by cp.async.bulk
// (sized like the staged rows themselves, window slack included: in the tiled mode it IS a
// second staged-rows buffer)
__host__ __device__ inline uint32_t trainedLandingBytes(uint32_t recUnits)
{
    return (recUnits * TRAINED_UNIT_WORDS + 4u) * 128u;
}
__host__ __device__ inline uint32_t trainedWarpBytes(uint32_t rowWords, uint32_t landingBytes = 0)
{
    return rowWords * 128 + TRAINED_TILE_BYTES + landingBytes;
}

// Where things live in the CTA's shared window, as ABSOLUTE shared addresses
// (`base` = address of the first dynamic shared byte, 1 KiB on sm_100).
struct TrainedLayout {
    uint32_t codebook;   // 1 KiB aligned
    uint32_t lut2;       // second-level entries, right after the codebooks
    uint32_t mbar;       // one mbarrier per warp (bulk-async staging)
    uint32_t lut;        // aligned to lutBytes
    uint32_t lowStart;   // warps [0, lowWarps) at lowStart + w * perWarp
    uint32_t lowWarps;
    uint32_t highStart;  // warps [lowWarps, ...) at highStart + (w - lowWarps) * perWarp
    uint32_t end;        // first byte not used
};
__host__ __device__ inline TrainedLayout trainedLayout(uint32_t base, uint32_t lutBytes, uint32_t perWarp,
                                                       uint32_t warps, uint32_t lut2Entries)
{
    TrainedLayout l;
    l.codebook = (base + 1023u) & ~1023u;
    l.lut2 = l.codebook + TRAINED_CODEBOOK_BYTES;
    l.mbar = (l.lut2 + lut2Entries * 4u + 7u) & ~7u;
    l.lowStart = (l.mbar + 8u * TRAINED_MAX_WARPS + 127u) & ~127u;
    l.lut = (l.lowStart + lutBytes - 1) / lutBytes * lutBytes;
    const uint32_t fit = (l.lut - l.lowStart) / perWarp;
    l.lowWarps = fit < warps ? fit : warps;
    l.highStart = l.lut + lutBytes;
    l.end = l.highStart + (warps - l.lowWarps) * perWarp;
    return l;
}

__device__ __forceinline__ uint32_t tileRowBase(uint32_t tile, uint32_t r)
{
    return tile + (r >> 2) * TRAINED_TILE_GROUP_BYTES + (r & 3u) * 32u;
}
// byte offset of ring word w (any integer, taken modulo the ring) within its row
__device__ __forceinline__ uint32_t tileWordOffset(int w)
{
    return ((static_cast<uint32_t>(w) & 8u) << 4) + ((static_cast<uint32_t>(w) & 7u) << 2);
}

__device__ __forceinline__ uint32_t smemAddr(const void* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// First dynamic shared byte of a kernel without static shared memory (the host
// sizes the launch with it; the decode kernel re-derives and checks it).
__global__ void smem_base_probe_kernel(uint32_t* out)
{
    extern __shared__ __align__(128) unsigned char probeSmem[];
    *out = smemAddr(probeSmem);
}

// ---- shared-memory accessors on 32-bit shared-window addresses ----
// All hot-loop traffic goes through these instead of generic pointers: with
// generic pointers nvcc re-derives the shared window base (S2UR SR_CgaCtaId /
// ULEA) and adds it to every gather address; with explicit shared addresses the
// loads compile to LDS [R + imm].  `volatile` keeps them ordered with respect to
// each other and to __syncwarp; ptxas still schedules them for latency.
__device__ __forceinline__ uint32_t ldsU32(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
template <int OFFSET>
__device__ __forceinline__ uint32_t ldsU32At(uint32_t addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1+%2];" : "=r"(v) : "r"(addr), "n"(OFFSET));
    return v;
}
__device__ __forceinline__ float ldsF32(uint32_t addr)
{
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void stsU32(uint32_t addr, uint32_t v)
{
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v));
}
// 128-bit streaming store, executed only when `pred` (no branch).
__device__ __forceinline__ void stcsF4If(bool pred, void* ptr, float4 v)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %5, 0;\n"
        "@p st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};\n"
        "}\n" ::"l"(ptr),
        "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(static_cast<uint32_t>(pred))
        : "memory");
}
// 128-bit load of the value an average epilogue adds to, executed only when `pred`.
__device__ __forceinline__ float4 ldF4If(bool pred, const void* ptr)
{
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %5, 0;\n"
        "@p ld.global.v4.f32 {%0, %1, %2, %3}, [%4];\n"
        "}\n"
        : "+f"(v.x), "+f"(v.y), "+f"(v.z), "+f"(v.w)
        : "l"(ptr), "r"(static_cast<uint32_t>(pred))
        : "memory");
    return v;
}
__device__ __forceinline__ void prefetchL2If(bool pred, const void* ptr)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %1, 0;\n"
        "@p prefetch.global.L2 [%0];\n"
        "}\n" ::"l"(ptr),
        "r"(static_cast<uint32_t>(pred)));
}
__device__ __forceinline__ void stcsF4(void* ptr, float4 v)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(ptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// ---- mbarrier + 1-D bulk copy (cp.async.bulk, the non-tensor TMA path) ----
__device__ __forceinline__ void mbarInit(uint32_t mbar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint32_t mbar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t mbar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MEMB_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MEMB_MBAR_DONE;\n"
        "bra MEMB_MBAR_WAIT;\n"
        "MEMB_MBAR_DONE:\n"
        "}\n" ::"r"(mbar),
        "r"(parity)
        : "memory");
}
// 1-D bulk copy global -> shared (bytes a multiple of 16, both addresses 16 B aligned),
// completion counted on the mbarrier.  SASS: UBLKCP.S.G
__device__ __forceinline__ void bulkLoad(uint32_t dst, const void* src, uint32_t bytes, uint32_t mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(mbar)
                 : "memory");
}

// One elected lane of the (converged) warp arms the mbarrier with `bytes` and, when `copy`, issues ONE
// bulk copy -- as predicated instructions, not under a divergent branch.
__device__ __forceinline__ void electExpectAndCopy(uint32_t mbar, uint32_t bytes, bool copy, uint32_t dst, const void* src)
{
    asm volatile(
        "{\n"
        ".reg .pred leader, go;\n"
        "elect.sync _|leader, 0xffffffff;\n"
        "setp.ne.and.u32 go, %4, 0, leader;\n"
        "@leader mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
        "@go cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%2], [%3], %1, [%0];\n"
        "}\n" ::"r"(mbar),
        "r"(bytes), "r"(dst), "l"(src), "r"(static_cast<uint32_t>(copy))
        : "memory");
}

// all earlier generic-proxy accesses of this thread to shared memory are ordered before its later
// async-proxy operations (a bulk copy that overwrites a buffer the thread has just read)
__device__ __forceinline__ void fenceProxyAsync()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ RowUnit ldsRowUnit(uint32_t addr)
{
    RowUnit v;
    static_assert(TRAINED_UNIT_WORDS == 4, "bulk-async staging is written for 16 B units");
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]) : "r"(addr));
    return v;
}

// The flush's store.  An NVLS multicast address (OutputSpec::multicast: the NVSwitch replicates the
// store into every GPU of the group) takes the same streaming store.  NOT multimem.st: ptxas 12.9
// lowers a predicated `@p multimem.st.weak.global.v4.f32` to an UNPREDICATED STG.E.128 -- in the
// first version every flush store ran twice and the predicated-off ones of the first / last chunk
// clobbered the neighbouring rows (found r02 by a whole-matrix comparison); and a second, separately
// predicated plain store beside the streaming one cost 2.5 % of the full-vocabulary decode.
__device__ __forceinline__ void flushStoreF4If(uint32_t multicast, bool pred, void* ptr, float4 v)
{
    (void)multicast;
    stcsF4If(pred, ptr, v);
}
// The stores of a row's FIRST and LAST chunk: they hold the row's two partial 128 B lines, and for every
// other row (1200 B rows) a 32 B sector shared with the neighbouring row, whose half arrives a tile time
// (~14 us, ~55 MB of other stores) later.  Build switch MEMB_BOUNDARY_STORE: 0 = the streaming store like
// everything else, 1 = plain st.global (evict-normal beside the evict-first body), 2 = L2::evict_last
// cache hint on both, 3 = evict_last on the first chunk only (within a tile the head of row r is written
// at the START of the tile and the tail of row r - 1, which completes the line, at its END), 4 = on the
// last chunk only (control) -- does the half-written sector survive in L2 until its other half arrives?
// Measured r02, 1 M random ids, interleaved repeats of 200 steps (each figure repeats to 0.1 % on one box;
// boxes differ by more than the effect): box A 0.2959 (0) / 0.3044 (1) / 0.2864 (2); boxes B, C 0.2981-0.2988
// (0) / 0.2974-0.2980 (2) / 0.2963-0.2968 (3) / 0.2999 (4).  Sequential ids: no difference.  Default 3.
#ifndef MEMB_BOUNDARY_STORE
#define MEMB_BOUNDARY_STORE 3
#endif
__device__ __forceinline__ uint64_t boundaryStorePolicy()
{
    uint64_t policy = 0;
#if MEMB_BOUNDARY_STORE >= 2
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
#endif
    return policy;
}
// FIRST: the store belongs to the row's first chunk (else to its last)
template <bool FIRST>
__device__ __forceinline__ void boundaryStoreF4If(uint64_t policy, bool pred, void* ptr, float4 v)
{
    constexpr bool hinted = MEMB_BOUNDARY_STORE == 2 || (MEMB_BOUNDARY_STORE == 3 && FIRST) || (MEMB_BOUNDARY_STORE == 4 && !FIRST);
    if (MEMB_BOUNDARY_STORE == 1) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "setp.ne.u32 p, %5, 0;\n"
            "@p st.global.v4.f32 [%0], {%1, %2, %3, %4};\n"
            "}\n" ::"l"(ptr),
            "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(static_cast<uint32_t>(pred))
            : "memory");
    } else if (hinted) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "setp.ne.u32 p, %5, 0;\n"
            "@p st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %6;\n"
            "}\n" ::"l"(ptr),
            "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(static_cast<uint32_t>(pred)), "l"(policy)
            : "memory");
    } else {
        stcsF4If(pred, ptr, v);
    }
}

// Four centroids for the four symbol fields packed in s4.  cb = shared address of
// the codebook (1 KiB aligned).  SMALLCB: a field already is the byte offset
// (sym * 4 <= 252), so one PRMT splices it under the upper bytes of cb.
template <bool SMALLCB>
__device__ __forceinline__ float4 gatherCentroids(uint32_t cb, uint32_t s4)
{
    float4 v;
    if (SMALLCB) {
        v.x = ldsF32(__byte_perm(s4, cb, 0x7650));
        v.y = ldsF32(__byte_perm(s4, cb, 0x7651));
        v.z = ldsF32(__byte_perm(s4, cb, 0x7652));
        v.w = ldsF32(__byte_perm(s4, cb, 0x7653));
    } else {
        v.x = ldsF32(cb | ((s4 << 2) & 0x3fcu));
        v.y = ldsF32(cb | ((s4 >> 6) & 0x3fcu));
        v.z = ldsF32(cb | ((s4 >> 14) & 0x3fcu));
        v.w = ldsF32(cb | ((s4 >> 22) & 0x3fcu));
    }
    return v;
}

// The same for symbol words of ANOTHER reader (field << shift = byte offset into its codebook).
__device__ __forceinline__ float4 gatherCentroidsShift(uint32_t cb, uint32_t s4, uint32_t shift)
{
    float4 v;
    v.x = ldsF32(cb + ((s4 & 0xffu) << shift));
    v.y = ldsF32(cb + (((s4 >> 8) & 0xffu) << shift));
    v.z = ldsF32(cb + (((s4 >> 16) & 0xffu) << shift));
    v.w = ldsF32(cb + ((s4 >> 24) << shift));
    return v;
}
// Four merged values for four symbol pairs: f0 / f1 = packed symbol fields of reader 0 / reader 1,
// m0 / m1 turn a field into its byte offset contribution (OutputSpec::pairEntries).
__device__ __forceinline__ float4 gatherPairs(uint32_t pairBase, uint32_t f0, uint32_t f1, uint32_t m0, uint32_t m1)
{
    float4 v;
    v.x = ldsF32(pairBase + (f0 & 0xffu) * m0 + (f1 & 0xffu) * m1);
    v.y = ldsF32(pairBase + ((f0 >> 8) & 0xffu) * m0 + ((f1 >> 8) & 0xffu) * m1);
    v.z = ldsF32(pairBase + ((f0 >> 16) & 0xffu) * m0 + ((f1 >> 16) & 0xffu) * m1);
    v.w = ldsF32(pairBase + (f0 >> 24) * m0 + (f1 >> 24) * m1);
    return v;
}
// The same through 16-bit lanes: two elements per ALU instruction.  e = field bytes spread to 16-bit
// halves (PRMT), offset pair = e0 * m0 + e1 * m1 (no carry between the halves: the table is < 64 KB).
__device__ __forceinline__ float4 gatherPairsWide(uint32_t pairBase, uint32_t f0, uint32_t f1, uint32_t m0, uint32_t m1)
{
    const uint32_t lo = __byte_perm(f0, 0u, 0x4140) * m0 + __byte_perm(f1, 0u, 0x4140) * m1;  // elements 0, 1
    const uint32_t hi = __byte_perm(f0, 0u, 0x4342) * m0 + __byte_perm(f1, 0u, 0x4342) * m1;  // elements 2, 3
    float4 v;
    v.x = ldsF32(pairBase + (lo & 0xffffu));
    v.y = ldsF32(pairBase + (lo >> 16));
    v.z = ldsF32(pairBase + (hi & 0xffffu));
    v.w = ldsF32(pairBase + (hi >> 16));
    return v;
}
__device__ __forceinline__ uint32_t ldU32If(bool pred, const uint32_t* ptr)
{
    uint32_t v = 0;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %2, 0;\n"
        "@p ld.global.u32 %0, [%1];\n"
        "}\n"
        : "+r"(v)
        : "l"(ptr), "r"(static_cast<uint32_t>(pred))
        : "memory");
    return v;
}
__device__ __forceinline__ void stU32If(bool pred, uint32_t* ptr, uint32_t v)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %2, 0;\n"
        "@p st.global.u32 [%0], %1;\n"
        "}\n" ::"l"(ptr),
        "r"(v), "r"(static_cast<uint32_t>(pred))
        : "memory");
}

// Decode state of one lane.  lutLane / rowLane are shared-window addresses.
struct LaneDecoder {
    uint32_t lo, hi;     // window: bit 0 of lo = bit `consumed` of the re-packed row
    uint32_t consumed;   // low 16 bits: stream bits consumed in this row so far (junk above)
    uint32_t lutLane;    // LUT base | lane * 4 (bank-replicated LUT) or LUT base (single copy)
    uint32_t lutMask;
    uint32_t rowLane;    // word w of this lane's compressed row lives at rowLane + w * 128
    uint32_t lut2;       // second-level table (ESCAPE)
    uint32_t padL;       // padBits + lutBits: where the bits after the LUT window start
};

// Re-centre the 64-bit window on the row's current bit position: three loads
// down the lane's own bank and two funnel shifts, no data-dependent branch.
__device__ __forceinline__ void reloadWindow(LaneDecoder& d)
{
    const uint32_t a = d.rowLane + ((d.consumed << 2) & 0x3ff80u);
    const uint32_t w0 = ldsU32At<0>(a), w1 = ldsU32At<128>(a), w2 = ldsU32At<256>(a);
    d.lo = __funnelshift_r(w0, w1, d.consumed);
    d.hi = __funnelshift_r(w1, w2, d.consumed);
}

// A code longer than the LUT is wide: second-level table indexed by the bits that follow the
// window (src/huffman_table_decoder.h:58-93,108-116).  Deliberately NOT inlined: inlined, ptxas
// predicates these nine instructions into every symbol of the chain (r02: 0.52 ms against
// 0.29 ms per 1 M rows); as a call it is a branch that a warp takes only when a lane has a long code.
__device__ __noinline__ uint32_t secondLevelEntry(uint32_t e, uint32_t lo, uint32_t padL, uint32_t lut2)
{
    const uint32_t idx = (e >> 16) + ((lo >> padL) & ((1u << (e & 15u)) - 1u));
    return ldsU32(lut2 + idx * 4u);
}

// One symbol: LUT entry for the next bits of the window, window advanced by the
// code length.  Returns the entry (symbol field in bits 16..23).
template <bool ESCAPE>
__device__ __forceinline__ uint32_t decodeSymbol(LaneDecoder& d, const TrainedParams& p)
{
    uint32_t e = ldsU32((d.lo & d.lutMask) | d.lutLane);
    if (ESCAPE) {
        if (e & 0x8000u) {
            e = secondLevelEntry(e, d.lo, d.padL, d.lut2);
        }
    }
    d.lo = __funnelshift_r(d.lo, d.hi, e);  // shift = e & 31 = code length
    d.hi = __funnelshift_r(d.hi, 0u, e);
    d.consumed += e;
    return e;
}

// NW words of four symbols each, fully unrolled, window re-centred every G
// symbols; word q goes to tileWord + 4*q.
template <int NW, int G, bool ESCAPE>
__device__ __forceinline__ void decodeWords(LaneDecoder& d, const TrainedParams& p, uint32_t tileWord, uint32_t symAnd,
                                            uint32_t symOr)
{
#pragma unroll
    for (int q = 0; q < NW; ++q) {
        uint32_t acc = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if ((q * 4 + k) % G == 0) {
                reloadWindow(d);
            }
            const uint32_t e = decodeSymbol<ESCAPE>(d, p);
            // byte 2 of e (the symbol field) -> byte k of acc
            acc = __byte_perm(acc, e, k == 0 ? 0x3216 : k == 1 ? 0x3260 : k == 2 ? 0x3610 : 0x6210);
        }
        stsU32(tileWord + 4 * q, (acc & symAnd) | symOr);
    }
}

// sum / R of np.mean.  R a power of two (two readers!): sum * 2^-k is the same correctly
// rounded real number as sum / 2^k, in one instruction instead of the IEEE divide sequence
// (which tripled the instruction count of the FINALIZE launch: 0.74 -> ms per 1 M rows).
// R = 3, 5, 6, 7: the constant-divisor sequence of divideByLevels (q0 = s * RN(1/R), one FMA
// residual, one FMA correction), verified EXHAUSTIVELY on the host against s / R for every
// float with 1e-20 < |s| < 1e20 and for +0 (scripts/divconst_exhaustive.c: 0 mismatches of
// 2.23e9 per divisor; -0 is NOT in the set: the sequence returns +0 for it).  Sums outside
// that set, and every other R, take the IEEE divide.  RN(1/R) is formed once per thread
// (it depends on the launch parameter only).
__device__ __forceinline__ float divideByReaders(float s, float divisor)
{
    const uint32_t bits = __float_as_uint(divisor);
    if ((bits & 0x007fffffu) == 0u) {
        return __fmul_rn(s, __uint_as_float(0x7f000000u - bits));  // 1 / 2^k
    }
    if (divisor <= 7.0f) {  // 3, 5, 6, 7 (R is an integer >= 2 and not a power of two here)
        const float inv = __frcp_rn(divisor);
        const float q0 = __fmul_rn(s, inv);
        const float e = __fmaf_rn(-q0, divisor, s);
        const float q1 = __fmaf_rn(e, inv, q0);
        const float as = fabsf(s);
        if ((as > 1e-20f && as < 1e20f) || __float_as_uint(s) == 0u) {
            return q1;
        }
    }
    return __fdiv_rn(s, divisor);
}

__device__ __forceinline__ float applyEpilogue(float v, float old, int epi, float divisor)
{
    if (epi == EPI_WRITE) {
        return v;
    }
    const float s = __fadd_rn(old, v);
    return epi == EPI_ACCUMULATE ? s : divideByReaders(s, divisor);
}

// np.mean's add chain on four lanes of a row: old + v, and for the last reader / R.
__device__ __forceinline__ float4 applyEpilogue4(float4 v, float4 old, int epi, float divisor)
{
    v.x = __fadd_rn(old.x, v.x);
    v.y = __fadd_rn(old.y, v.y);
    v.z = __fadd_rn(old.z, v.z);
    v.w = __fadd_rn(old.w, v.w);
    if (epi == EPI_FINALIZE) {
        v.x = divideByReaders(v.x, divisor);
        v.y = divideByReaders(v.y, divisor);
        v.z = divideByReaders(v.z, divisor);
        v.w = divideByReaders(v.w, divisor);
    }
    return v;
}

__device__ __forceinline__ RowUnit loadRowUnit(const RowUnit* p)
{
    RowUnit v;
    if (TRAINED_UNIT_WORDS == 8) {
        asm volatile("ld.global.nc.v8.u32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3]), "=r"(v.w[4 % TRAINED_UNIT_WORDS]),
                       "=r"(v.w[5 % TRAINED_UNIT_WORDS]), "=r"(v.w[6 % TRAINED_UNIT_WORDS]), "=r"(v.w[7 % TRAINED_UNIT_WORDS])
                     : "l"(p));
    } else {
        asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];"
                     : "=r"(v.w[0]), "=r"(v.w[1]), "=r"(v.w[2]), "=r"(v.w[3])
                     : "l"(p));
    }
    return v;
}
// scatter a unit down the lane's own bank: word w of the unit to addr + w * 128
__device__ __forceinline__ void stageRowUnit(uint32_t addr, const RowUnit& v)
{
#pragma unroll
    for (uint32_t w = 0; w < TRAINED_UNIT_WORDS; ++w) {
        stsU32(addr + w * 128, v.w[w]);
    }
}

__device__ __forceinline__ uint32_t shflFrom(uint32_t v, uint32_t srcLane)
{
    uint32_t r;
    asm volatile("shfl.sync.idx.b32 %0, %1, %2, 0x1f, 0xffffffff;" : "=r"(r) : "r"(v), "r"(srcLane));
    return r;
}

// G = symbols decoded per window reload (G * maxLen <= 64 - padBits).  VEC4: dim,
// stride and column offset are multiples of 4 and the output is 16 B aligned ->
// 128-bit streaming stores.  ESCAPE: some code is longer than the LUT is wide.
// SMALLCB: symShift == 2 (see gatherCentroids).
// RMW: the launch carries an average epilogue (ACCUMULATE / FINALIZE): the fast flush
// reads the running sum, adds, divides for the last reader, and stores.
// SIDE: the launch redirects first / last units to side buffers (OutputSpec::sideHead).
// SYM: 3 = 2 with the merged values taken from the pair table (OutputSpec::pairEntries != 0);
// SYM: 1 = the launch stores symbol words instead of fp32 rows, 2 = the running sum it adds to
// is gathered from another reader's symbol words (OutputSpec::sym).
// TMA: the warp's 32 records reach shared memory through cp.async.bulk (one 1-D bulk copy per
// lane into a landing buffer, completion on the warp's mbarrier), a tile ahead, instead of
// through 128-bit loads into registers: no gathered LDG in the LSU pipe, no prefetch registers.
// Needs fixed-stride records (TrainedParams::recUnits).  Measured r02 (1 M rows, 2.2 M x 300 4-bit):
// 32 bulk copies of 144 B per tile are SLOWER than the register path (0.305 vs 0.296 ms random,
// 0.304 vs 0.277 ms sequential): the copy engine's per-operation cost, not bytes, is the limit.
// TMA == 2 (range decodes whose first row is a multiple of 32): the tile's 32 rows are one
// contiguous block of TrainedParams::blobTiled, already in the staged layout -- ONE bulk copy per
// tile straight into a second staged-rows buffer, no copy-out, no row ids (4 bit: 0.576 vs 0.582 ms
// for the 2.2 M-row table; 8 bit, 272 B records: 0.706 vs 0.775 ms).
// Option trained_tma: 0 off, 1 (default) tiled range decodes + one copy per row for records longer
// than the prefetch registers, 2 one copy per row everywhere.
template <int G, bool VEC4, bool ESCAPE, bool SMALLCB, bool RMW, bool SIDE, int SYM, int TMA = 0>
__global__ void __maxnreg__(TRAINED_MAX_REGS)
trained_decode_kernel(const TrainedParams p, const int32_t* __restrict__ ids, uint32_t rowBegin, uint64_t n,
                      const OutputSpec o, const int epi, uint32_t* __restrict__ tileCounter)
{
    // SYM == 3 is SYM == 2 with the pair table as a compile-time fact: the two-gather flush, its adds and the divide
    // sequences are not in the code at all (ncu r02 of the 82 KB SYM == 2 kernel, both flavours compiled in: stall
    // no_instruction 1.16 per issue, 18 % of its samples, in the rarely executed first / last chunk flush)
    constexpr int SYMK = SYM == 3 ? 2 : SYM;
    constexpr bool PAIRS = SYM == 3;
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t lane = threadIdx.x % WARP;
    const uint32_t warp = threadIdx.x / WARP;
    const uint32_t warpsPerCta = blockDim.x / WARP;

    const uint32_t lutBytes = trainedLutBytes(p.lutBits, p.lutReplicas);
    const uint32_t base = smemAddr(smem);
    const uint32_t perWarp = trainedWarpBytes(p.rowWords, TMA ? trainedLandingBytes(p.recUnits) : 0u);
    // (the region behind the codebooks holds the second-level table, or -- SYMK == 2, never with long codes -- the pair table)
    const TrainedLayout lay = trainedLayout(base, lutBytes, perWarp, warpsPerCta, SYMK == 2 ? o.pairEntries * o.pairReplicas : p.lut2Entries);
    {
        uint32_t dynBytes;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dynBytes));
        if (lay.end - base > dynBytes) {
            __trap();  // the host sized the launch for another shared-window base
        }
        float* cbS = reinterpret_cast<float*>(smem + (lay.codebook - base));
        uint32_t* lutS = reinterpret_cast<uint32_t*>(smem + (lay.lut - base));
        const uint32_t entries = 1u << p.lutBits;
        if (p.lutReplicas == 1) {
            for (uint32_t i = threadIdx.x; i < entries; i += blockDim.x) {
                lutS[i] = p.lut[i];
            }
        } else {
            // every thread fetches its entries once (independent loads, one round trip for
            // all of them) and writes the 32 bank replicas of each with eight 128-bit stores;
            // one global load per replica word had the prologue wait on ~50 dependent L2
            // round trips per thread (ncu r01: 2 % of the kernel's stall samples)
            for (uint32_t e0 = threadIdx.x; e0 < entries; e0 += 4 * blockDim.x) {
                uint32_t v[4];
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    const uint32_t e = e0 + k * blockDim.x;
                    v[k] = e < entries ? __ldg(p.lut + e) : 0u;
                }
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    const uint32_t e = e0 + k * blockDim.x;
                    if (e < entries) {
                        uint4* dst = reinterpret_cast<uint4*>(lutS + e * WARP);
#pragma unroll
                        for (uint32_t q = 0; q < WARP / 4; ++q) {
                            dst[q] = make_uint4(v[k], v[k], v[k], v[k]);
                        }
                    }
                }
            }
        }
        if (PAIRS) {
            float* pairS = reinterpret_cast<float*>(smem + (lay.lut2 - base));
            for (uint32_t i = threadIdx.x; i < o.pairEntries; i += blockDim.x) {
                const uint32_t s0 = i / o.pairStride, s1 = i % o.pairStride;
                const float c0 = s0 < o.symCodebookSize ? o.symCodebook[s0] : 0.f;
                const float c1 = s1 < p.codebookSize ? p.codebook[s1] : 0.f;
                const float t = applyEpilogue(c1, c0, epi, o.divisor);  // c0 + c1 in reader order, / R for the last reader
                if (o.pairReplicas == 1) {
                    pairS[i] = t;
                } else {
                    float4* dst = reinterpret_cast<float4*>(pairS + i * WARP);  // (lay.lut2 is 1 KiB aligned: replica l sits in bank l)
#pragma unroll
                    for (uint32_t q = 0; q < WARP / 4; ++q) {
                        dst[q] = make_float4(t, t, t, t);
                    }
                }
            }
        }
        if (ESCAPE) {
            uint32_t* lut2S = reinterpret_cast<uint32_t*>(smem + (lay.lut2 - base));
            for (uint32_t i = threadIdx.x; i < p.lut2Entries; i += blockDim.x) {
                lut2S[i] = p.lut2[i];
            }
        }
        for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
            cbS[i] = i < p.codebookSize ? p.codebook[i] : 0.f;
            if (SYMK == 2) {
                cbS[256 + i] = i < o.symCodebookSize ? o.symCodebook[i] : 0.f;
            }
        }
    }
    __syncthreads();

    // shared-window addresses
    const uint32_t cb = lay.codebook;
    const uint32_t warpBase = warp < lay.lowWarps ? lay.lowStart + warp * perWarp
                                                  : lay.highStart + (warp - lay.lowWarps) * perWarp;
    const uint32_t tile = warpBase + p.rowWords * 128;  // 32 rows x 132 B of symbols (ring of two chunks)
    const uint32_t landing = tile + TRAINED_TILE_BYTES;  // TMA: record of lane l at + l * recBytes
    const uint32_t recBytes = p.recUnits * TRAINED_UNIT_BYTES;
    const uint32_t mbar = lay.mbar + warp * 8u;
    LaneDecoder d;
    d.lutLane = lay.lut | (p.lutReplicas == 1 ? 0u : lane * 4u);
    d.lutMask = p.lutMask;
    d.rowLane = warpBase + lane * 4;
    d.lut2 = lay.lut2;
    d.padL = p.padBits + p.lutBits;
    const uint32_t tileLane = tileRowBase(tile, lane);
    const uint64_t boundaryPolicy = boundaryStorePolicy();
    const uint32_t pairBase = lay.lut2 + (SYMK == 2 && o.pairReplicas != 1 ? lane * 4u : 0u);  // lane l reads bank l
    // The flush reads ring slots it then does not store (units before the row in the
    // first chunk, units past it in the last).  With SMALLCB a symbol byte IS a shared
    // address offset, so those slots must never hold anything but multiples of 4.
    for (uint32_t i = lane; i < TRAINED_TILE_BYTES / 4; i += WARP) {
        stsU32(tile + i * 4, 0u);
    }
    __syncwarp();

    // tile numbers fit 32 bits: the host refuses launches with tiles + 4 * warps >= 2^32
    const uint32_t numTiles = static_cast<uint32_t>((n + WARP - 1) / WARP);
    const uint32_t totalWarps = gridDim.x * warpsPerCta;
    const uint32_t dim = p.dim;
    // flush role: CHUNK_UNITS lanes per output row, FLUSH_ROWS rows per store instruction
    const uint32_t fr = lane / TRAINED_CHUNK_UNITS, fq = lane % TRAINED_CHUNK_UNITS;
    const uint64_t rowBytes = o.rowStride * 4;

    // Three-stage software pipeline over the warp's tiles (every stage is one HBM
    // round trip, none of them is waited for in the tile that issues it):
    //   tile t+3  row id             (ids[i], or rowBegin + i for a range decode)
    //   tile t+2  row offsets        (rowOff16[row], rowOff16[row + 1])
    //   tile t+1  compressed row     (up to TRAINED_PREFETCH_UNITS x 16 B into registers)
    //   tile t    decode from shared memory
    // Tiles are handed out dynamically (one atomicAdd per tile, also a stage ahead)
    // after a static first round, so no SM idles while another still has a queue.
    // A lane without a row (id < 0, id out of range, past the end of the batch)
    // decodes row 0 instead, which keeps it in lock step on valid data; its output
    // is forced to zero at the flush.
    // (the loaded id is only touched one tile later: nothing here consumes it)
    auto fetchId = [&](uint32_t tt, bool& has) -> uint32_t {
        const uint64_t i = static_cast<uint64_t>(tt) * WARP + lane;
        has = tt < numTiles && i < n;
        if (!has) {
            return 0;
        }
        return ids ? static_cast<uint32_t>(__ldg(ids + i)) : rowBegin + static_cast<uint32_t>(i);
    };
    auto fetchOffsets = [&](uint32_t row, bool has, uint32_t& lo, uint32_t& hi) -> bool {
        const bool ok = has && row < p.rows;  // a negative id is a huge unsigned one
        lo = hi = 0;
        if (p.rows) {
            const uint32_t r = ok ? row : 0u;
            if (p.recUnits) {
                lo = r * p.recUnits;
                hi = lo + p.recUnits;
            } else {
                lo = __ldg(p.rowOff16 + r);
                hi = __ldg(p.rowOff16 + r + 1);
            }
        }
        return ok;
    };
    RowUnit pre[TRAINED_PREFETCH_UNITS];
    auto prefetchRows = [&](uint32_t rowO16, uint32_t rowUnits) {
        const RowUnit* src = p.blob + rowO16;
#pragma unroll
        for (uint32_t k = 0; k < TRAINED_PREFETCH_UNITS; ++k) {
            if (k < rowUnits) {
                pre[k] = loadRowUnit(src + k);
            }
        }
    };

    // tile queue of this warp: t (decoding), t1 (rows in flight), t2 (offsets in
    // flight), t3 (id in flight); `ticket` is lane 0's pending atomicAdd result
    uint32_t t = blockIdx.x * warpsPerCta + warp;
    uint32_t t1 = t + totalWarps, t2 = t1 + totalWarps, t3 = t2 + totalWarps;
    uint32_t ticket = 0, ticketLane = 0;
    auto drawTicket = [&]() {
        // elect.sync + a predicated PTX atomic: atomicAdd() under `lane == 0` compiles to
        // a warp-aggregated atomic whose broadcast shuffle waits for the L2 round trip on
        // the spot (ncu r01: 10 % of all stall samples sat on that shuffle).  This way
        // nothing touches `ticket` until it is redeemed a tile later.
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "elect.sync %1|p, 0xffffffff;\n"
            "@p atom.global.add.u32 %0, [%2], 1;\n"
            "}\n"
            : "+r"(ticket), "=r"(ticketLane)
            : "l"(tileCounter)
            : "memory");
    };
    drawTicket();
    uint32_t offLo = 0, offHi = 0, offLoNext = 0, offHiNext = 0;
    bool hasRow;
    uint32_t rowAfter = fetchId(t, hasRow);
    bool present, presentNext;
    uint32_t o16 = 0, units = 0, mbarPhase = 0;
    // TMA: every lane copies its own record (record 0 for a lane without a row) into the landing
    // buffer; lane 0 arms the warp's mbarrier with the 32 records' bytes first
    auto issueRecords = [&](uint32_t row, bool ok) {
        electExpectAndCopy(mbar, 32u * recBytes, false, 0u, nullptr);
        __syncwarp();
        bulkLoad(landing + lane * recBytes, p.blob + static_cast<uint64_t>(ok ? row : 0u) * p.recUnits, recBytes, mbar);
    };
    const uint32_t rowsBuffer = d.rowLane;  // TMA == 2: the decode alternates between this and the landing buffer
    uint32_t tiledBuffer = 0;
    auto issueTile = [&](uint32_t tt, uint32_t buffer) {
        if (tt < numTiles) {  // (warp-uniform)
            const uint32_t bytes = p.recUnits * (TRAINED_UNIT_BYTES * 32u);
            electExpectAndCopy(mbar, bytes, true, buffer ? landing : warpBase,
                               p.blobTiled + (static_cast<uint64_t>(rowBegin >> 5) + tt) * (p.recUnits * 32u));
        }
    };
    if (TMA == 2) {
        if (lane == 0) {
            mbarInit(mbar, 1);
        }
        __syncwarp();
        issueTile(t, 0);
        present = hasRow && rowAfter < p.rows;
        presentNext = false;
    } else if (TMA) {
        // pipeline: tile t+2 row id, tile t+1 records in flight (bulk copies), tile t decode
        if (lane == 0) {
            mbarInit(mbar, 1);
        }
        __syncwarp();
        present = hasRow && rowAfter < p.rows;
        issueRecords(rowAfter, present);
        rowAfter = fetchId(t1, hasRow);
        presentNext = false;
    } else {
        present = fetchOffsets(rowAfter, hasRow, offLo, offHi);
        rowAfter = fetchId(t1, hasRow);
        presentNext = fetchOffsets(rowAfter, hasRow, offLoNext, offHiNext);
        rowAfter = fetchId(t2, hasRow);
        o16 = offLo;
        units = offHi - offLo;
        prefetchRows(o16, units);
    }

    while (t < numTiles) {
        // The ticket drawn a tile ago is redeemed BEFORE this tile's loads are
        // issued: ncu r01 showed the shuffle waiting on the scoreboard it shared
        // with the loads issued just ahead of it (10 % of all stall samples).
        const uint32_t t4 = (TMA ? 3u : 4u) * totalWarps + shflFrom(ticket, ticketLane);
        const uint32_t presentMask = __ballot_sync(0xffffffffu, present);
        if (TMA == 2) {
            // ---- the tile has landed in the staged layout; the next one goes to the other buffer ----
            mbarWait(mbar, mbarPhase);
            __syncwarp();
            mbarPhase ^= 1u;
            d.rowLane = tiledBuffer ? landing + lane * 4u : rowsBuffer;
            tiledBuffer ^= 1u;
            fenceProxyAsync();  // (the other buffer was read by the previous tile's decode, which ended in a __syncwarp)
            issueTile(t1, tiledBuffer);
            drawTicket();
            rowAfter = fetchId(t1, hasRow);
            presentNext = hasRow && rowAfter < p.rows;
        } else if (TMA) {
            // ---- the records have landed: own record -> own bank; then the next tile's copies ----
            mbarWait(mbar, mbarPhase);
            mbarPhase ^= 1u;
            for (uint32_t k0 = 0; k0 < p.recUnits; k0 += 4) {
                RowUnit c[4];
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    if (k0 + k < p.recUnits) {
                        c[k] = ldsRowUnit(landing + lane * recBytes + (k0 + k) * TRAINED_UNIT_BYTES);
                    }
                }
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    if (k0 + k < p.recUnits) {
                        stageRowUnit(d.rowLane + (k0 + k) * (TRAINED_UNIT_WORDS * 128), c[k]);
                    }
                }
            }
            fenceProxyAsync();
            __syncwarp();
            presentNext = hasRow && rowAfter < p.rows;
            issueRecords(rowAfter, presentNext);
            drawTicket();
            rowAfter = fetchId(t2, hasRow);
        }
        // ---- stage: own row -> own bank ----
        if (!TMA) {
#pragma unroll
        for (uint32_t k = 0; k < TRAINED_PREFETCH_UNITS; ++k) {
            if (k < units) {
                stageRowUnit(d.rowLane + k * (TRAINED_UNIT_WORDS * 128), pre[k]);
            }
        }
        }
        if (!TMA) {
            // rows longer than the prefetch window: the rest comes straight from HBM
            const RowUnit* src = p.blob + o16;
            const uint32_t unitsMax = __reduce_max_sync(0xffffffffu, presentMask ? units : 0u);
            for (uint32_t j = TRAINED_PREFETCH_UNITS; j < unitsMax; j += 4) {
                RowUnit c[4];
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    if (j + k < units) {
                        c[k] = loadRowUnit(src + j + k);
                    }
                }
#pragma unroll
                for (uint32_t k = 0; k < 4; ++k) {
                    if (j + k < units) {
                        stageRowUnit(d.rowLane + (j + k) * (TRAINED_UNIT_WORDS * 128), c[k]);
                    }
                }
            }
        }
        // advance the pipeline: ticket of t+4, rows of t+1, offsets of t+2, id of t+3
        uint32_t o16Next = 0, unitsNext = 0, offLoAfter = 0, offHiAfter = 0;
        bool presentAfter = false;
        if (!TMA) {
            drawTicket();
            o16Next = offLoNext;
            unitsNext = offHiNext - offLoNext;
            prefetchRows(o16Next, unitsNext);
            presentAfter = fetchOffsets(rowAfter, hasRow, offLoAfter, offHiAfter);
            rowAfter = fetchId(t3, hasRow);
        }

        d.hi = d.lo = 0;
        d.consumed = 0;
        const uint64_t row0 = static_cast<uint64_t>(t) * WARP;
        const uint32_t rowsHere = static_cast<uint32_t>(min(static_cast<uint64_t>(WARP), n - row0));
        unsigned char* const outTile = reinterpret_cast<unsigned char*>(o.out + row0 * o.rowStride + o.colOffset);
        // flush geometry of this tile: lane (fr, fq) serves rows fr, fr+FLUSH_ROWS, ... ; row
        // FLUSH_ROWS*it+fr starts aBase + it*aStep 16 B units into its 128 B line (mod 8)
        unsigned char* const rowPtr0 = outTile + fr * rowBytes;
        const uint32_t aBase = static_cast<uint32_t>(reinterpret_cast<uintptr_t>(rowPtr0) >> 4);
        const uint32_t aStep = static_cast<uint32_t>(rowBytes >> 4) * TRAINED_FLUSH_ROWS;  // FLUSH_ROWS rows, in 16 B units
        const uint32_t tileRows = tileRowBase(tile, fr);  // rows 4*it + fr: + it * TRAINED_TILE_GROUP_BYTES
        // SYM: symbol words of row 4*it + fr of this tile start at symRow0 + it * symStep
        uint32_t* const symRow0 = SYMK ? o.sym + (row0 + fr) * o.symWords : nullptr;
        const uint64_t symStep = static_cast<uint64_t>(TRAINED_FLUSH_ROWS) * o.symWords;
        const uint32_t cb0 = cb + 1024;  // the other reader's codebook (SYMK == 2)
        if (SYMK == 2 && rowsHere == WARP) {
            // The tile's symbol words (32 rows x symWords, one contiguous run) are pulled into L2 now: the register
            // loads a chunk ahead of their use were still in flight when the flush wanted them (ncu r02: long
            // scoreboard on the flush's first PRMT, 10 % of the launch's samples) -- from L2 they are not.
            const char* const symTile = reinterpret_cast<const char*>(o.sym + row0 * o.symWords);
            const uint32_t symBytes = WARP * o.symWords * 4u;
            for (uint32_t b = lane * 128u; b < symBytes; b += WARP * 128u) {
                prefetchL2If(true, symTile + b);
            }
        }
        // A lane without a row (missing word) decodes row 0 like the others but writes the
        // symbol of the codebook's +0.0 slot to the tile, so the straight-line flush needs no
        // per-row predicate: only a ragged last tile (or a codebook with no free slot) takes
        // the general path.
        const bool zeroFill = p.hasZeroSlot && !present;
        const uint32_t symAnd = zeroFill ? 0u : 0xffffffffu;
        const uint32_t symOr = zeroFill ? p.zeroWord : 0u;
        const bool fastFlush = VEC4 && rowsHere == WARP && presentMask != 0u &&
                               (presentMask == 0xffffffffu || p.hasZeroSlot) &&
                               (RMW ? epi != EPI_WRITE : epi == EPI_WRITE);

        for (uint32_t c0 = 0; c0 < dim; c0 += TRAINED_CHUNK) {
            const uint32_t nsym = min(TRAINED_CHUNK, dim - c0);
            const uint32_t word0 = (c0 / 4) & (TRAINED_RING_UNITS - 1);  // ring position of this chunk's first word
            // SYMK == 2: the symbol words this chunk's flush adds to are LOADED here, before the decode, and
            // consumed after it (eight registers per lane; fp32 running sums would be 32: L2 prefetch only).
            // Measured r02 on the two-reader average, 1 M Zipf words (stable once the stream-ordered pool kept
            // its memory): 0.777 ms two-gather flush -> 0.752 with the pair table -> 0.723 with the
            // predicate-free middle-chunk flush below.  The launch's cost is its flush's INSTRUCTIONS (ncu:
            // 226 M warp instructions against 119 M for a plain decode, ~55 per flushed unit, half of them
            // pair-table address math; 43 M bank conflicts on the table, which spans every bank).
            uint32_t symAhead[TRAINED_CHUNK_UNITS];
            if (RMW && fastFlush) {
                // the flush of this chunk will read these 16 B units of the running sum
                const int ub = int(c0 / 4);
                const int ua = int((c0 + nsym) / 4);
                const bool lastChunk = c0 + TRAINED_CHUNK >= dim;
#pragma unroll
                for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                    const int a = int((aBase + it * aStep) & 7u);
                    const int j = ub - a + int(fq);
                    if (SYMK == 2) {
                        // (exactly the predicate of the flush's first pass)
                        symAhead[it] = ldU32If(j >= 0 && j < (lastChunk ? ua : ua - a), symRow0 + it * symStep + j);
                    } else {
                        prefetchL2If(j >= 0 && j < ua, reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j);
                    }
                }
            }
            if (presentMask) {
                uint32_t tw = tileLane + (word0 >> 3) * 128;  // this chunk's half of the ring
                if (nsym == TRAINED_CHUNK) {
                    decodeWords<TRAINED_CHUNK / 4, G, ESCAPE>(d, p, tw, symAnd, symOr);
                } else {
                    uint32_t nw = nsym / 4;
                    if (nw & 8) {
                        decodeWords<8, G, ESCAPE>(d, p, tw, symAnd, symOr);
                        tw += 32;
                    }
                    if (nw & 4) {
                        decodeWords<4, G, ESCAPE>(d, p, tw, symAnd, symOr);
                        tw += 16;
                    }
                    if (nw & 2) {
                        decodeWords<2, G, ESCAPE>(d, p, tw, symAnd, symOr);
                        tw += 8;
                    }
                    if (nw & 1) {
                        decodeWords<1, G, ESCAPE>(d, p, tw, symAnd, symOr);
                        tw += 4;
                    }
                    const uint32_t rest = nsym & 3;
                    if (rest) {
                        reloadWindow(d);
                        uint32_t acc = 0;
                        for (uint32_t k = 0; k < rest; ++k) {
                            const uint32_t e = decodeSymbol<ESCAPE>(d, p);
                            acc |= ((e >> 16) & 0xffu) << (8 * k);
                        }
                        stsU32(tw, (acc & symAnd) | symOr);
                    }
                }
            }
            __syncwarp();
            if (VEC4) {
                // Flush in whole 128 B lines of the OUTPUT: row r lags the decode by a_r
                // 16 B units, where a_r is its start address modulo a line, so every
                // store instruction writes two aligned 256 B runs (microbench r01:
                // line-aligned runs of >= 256 B reach the plain-copy write rate, the
                // unaligned 128 B pieces of the first version capped at 2.7 TB/s).
                const bool last = c0 + TRAINED_CHUNK >= dim;
                const int unitsBefore = int(c0 / 4);
                const int unitsAfter = int((c0 + nsym) / 4);
                if (PAIRS && fastFlush && RMW && !last && c0 != 0 && o.straightPairs) {
                    // pair-table average, middle chunk: every unit of the lagged window exists -- no predicates,
                    // no pass loop, address math on two elements per instruction (gatherPairsWide)
                    const uint32_t m0 = (o.pairStride * o.pairReplicas) << o.symFieldShift;
                    const uint32_t m1 = o.pairReplicas * (SMALLCB ? 1u : 4u);
#pragma unroll
                    for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                        const int a = int((aBase + it * aStep) & 7u);
                        const int j = unitsBefore - a + int(fq);
                        const uint32_t s4 = ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j));
                        stcsF4(reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j,
                               gatherPairsWide(pairBase, symAhead[it], s4, m0, m1));
                    }
                } else if (fastFlush && RMW) {
                    // average epilogue, straight-line: the running sums of the chunk were
                    // pulled into L2 while the chunk was being decoded; load them all, then
                    // gather, add (divide for the last reader) and store.  A second pass
                    // drains the lagging units after the last chunk.
                    const int passes = last ? 2 : 1;
                    for (int pass = 0; pass < passes; ++pass) {
                        const int shift = pass * int(TRAINED_CHUNK_UNITS);
                        const int end0 = last ? unitsAfter : 0x7fffffff;
                        float4 old[TRAINED_CHUNK_UNITS];
#pragma unroll
                        for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                            const int a = int((aBase + it * aStep) & 7u);
                            const int j = unitsBefore - a + int(fq) + shift;
                            const int end = last ? end0 : unitsAfter - a;
                            if (SYMK == 2) {
                                old[it].x = __uint_as_float(pass == 0 ? symAhead[it]
                                                                      : ldU32If(j >= 0 && j < end, symRow0 + it * symStep + j));
                            } else {
                                old[it] = ldF4If(j >= 0 && j < end, reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j);
                            }
                        }
#pragma unroll
                        for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                            const int a = int((aBase + it * aStep) & 7u);
                            const int j = unitsBefore - a + int(fq) + shift;
                            const int end = last ? end0 : unitsAfter - a;
                            const uint32_t s4 = ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j));
                            float4 v;
                            if (PAIRS) {
                                v = gatherPairs(pairBase, __float_as_uint(old[it].x), s4,
                                                (o.pairStride * o.pairReplicas) << o.symFieldShift,
                                                o.pairReplicas * (SMALLCB ? 1u : 4u));
                            } else {
                                const float4 sum = SYMK == 2 ? gatherCentroidsShift(cb0, __float_as_uint(old[it].x), o.symFieldShift)
                                                            : old[it];
                                v = applyEpilogue4(gatherCentroids<SMALLCB>(cb, s4), sum, epi, o.divisor);
                            }
                            float4* const dst = reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j;
                            if (c0 == 0) {  // (warp-uniform)
                                boundaryStoreF4If<true>(boundaryPolicy, j >= 0 && j < end, dst, v);
                            } else if (last) {
                                boundaryStoreF4If<false>(boundaryPolicy, j >= 0 && j < end, dst, v);
                            } else {
                                stcsF4If(j >= 0 && j < end, dst, v);
                            }
                        }
                    }
                } else if (fastFlush) {
                    // common case, straight-line: 16 x (1 word of symbols -> 4 centroids -> 128-bit
                    // store).  A middle chunk needs no predicate at all (the 16 units a lane
                    // group covers are exactly the chunk, shifted by the row's lag); the
                    // first chunk skips units before the row, the last one drains the lag.
                    if (!last && c0 != 0) {
#pragma unroll
                        for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                            const int a = int((aBase + it * aStep) & 7u);
                            const int j = unitsBefore - a + int(fq);
                            const uint32_t s4 = ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j));
                            if (SYMK == 1) {
                                symRow0[it * symStep + j] = s4;
                                continue;
                            }
                            const float4 v = gatherCentroids<SMALLCB>(cb, s4);
                            flushStoreF4If(o.multicast, true, reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j, v);
                        }
                    } else {
#pragma unroll
                        for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                            const int a = int((aBase + it * aStep) & 7u);
                            const int j = unitsBefore - a + int(fq);
                            const int end = last ? unitsAfter : unitsAfter - a;
                            const uint32_t s4 = ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j));
                            if (SYMK == 1) {
                                stU32If(j >= 0 && j < end, symRow0 + it * symStep + j, s4);
                                continue;
                            }
                            const float4 v = gatherCentroids<SMALLCB>(cb, s4);
                            float4* dst = reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j;
                            if (SIDE) {
                                dst = redirectUnit(o, dst, row0 + it * TRAINED_FLUSH_ROWS + fr, static_cast<uint32_t>(j), dim / 4 - 1);
                            }
                            if (c0 == 0) {  // (warp-uniform)
                                boundaryStoreF4If<true>(boundaryPolicy, j >= 0 && j < end, dst, v);
                            } else {
                                boundaryStoreF4If<false>(boundaryPolicy, j >= 0 && j < end, dst, v);
                            }
                        }
                        if (last) {
#pragma unroll
                            for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                                const int a = int((aBase + it * aStep) & 7u);
                                const int j2 = unitsBefore - a + int(fq) + int(TRAINED_CHUNK_UNITS);
                                const uint32_t s4 = ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j2));
                                if (SYMK == 1) {
                                    stU32If(j2 < unitsAfter, symRow0 + it * symStep + j2, s4);
                                    continue;
                                }
                                const float4 u = gatherCentroids<SMALLCB>(cb, s4);
                                float4* dst = reinterpret_cast<float4*>(rowPtr0 + it * TRAINED_FLUSH_ROWS * rowBytes) + j2;
                                if (SIDE) {
                                    dst = redirectUnit(o, dst, row0 + it * TRAINED_FLUSH_ROWS + fr, static_cast<uint32_t>(j2), dim / 4 - 1);
                                }
                                if (c0 == 0) {
                                    boundaryStoreF4If<true>(boundaryPolicy, j2 < unitsAfter, dst, u);
                                } else {
                                    boundaryStoreF4If<false>(boundaryPolicy, j2 < unitsAfter, dst, u);
                                }
                            }
                        }
                    }
                } else {
                    unsigned char* rowPtr = rowPtr0;
                    for (uint32_t it = 0; it < TRAINED_CHUNK_UNITS; ++it) {
                        const uint32_t r = it * TRAINED_FLUSH_ROWS + fr;
                        if (r < rowsHere) {
                            const int a = int((aBase + it * aStep) & 7u);
                            const int begin = max(unitsBefore - a, 0);
                            const int end = last ? unitsAfter : unitsAfter - a;
                            const bool rowPresent = (presentMask >> r) & 1u;
                            for (int j = begin + int(fq); j < end; j += int(TRAINED_CHUNK_UNITS)) {
                                if (SYMK == 1) {
                                    o.sym[(row0 + r) * o.symWords + j] =
                                        rowPresent ? ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j)) : p.zeroWord;
                                    continue;
                                }
                                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                                if (rowPresent) {
                                    v = gatherCentroids<SMALLCB>(
                                        cb, ldsU32(tileRows + it * TRAINED_TILE_GROUP_BYTES + tileWordOffset(j)));
                                }
                                float4* dst = reinterpret_cast<float4*>(rowPtr) + j;
                                if (SIDE) {
                                    dst = redirectUnit(o, dst, row0 + r, static_cast<uint32_t>(j), dim / 4 - 1);
                                }
                                if (epi != EPI_WRITE) {
                                    const float4 old = SYMK == 2 ? gatherCentroidsShift(cb0, o.sym[(row0 + r) * o.symWords + j],
                                                                                       o.symFieldShift)
                                                                : *dst;
                                    v.x = applyEpilogue(v.x, old.x, epi, o.divisor);
                                    v.y = applyEpilogue(v.y, old.y, epi, o.divisor);
                                    v.z = applyEpilogue(v.z, old.z, epi, o.divisor);
                                    v.w = applyEpilogue(v.w, old.w, epi, o.divisor);
                                }
                                flushStoreF4If(o.multicast, true, dst, v);
                            }
                        }
                        rowPtr += TRAINED_FLUSH_ROWS * rowBytes;
                    }
                }
            } else {
                const uint32_t fieldShift = 2u - p.symShift;  // symbol field -> byte offset into the codebook
                for (uint32_t r = 0; r < rowsHere; ++r) {
                    const bool rowPresent = (presentMask >> r) & 1u;
                    const uint32_t tileRow = tileRowBase(tile, r) + (word0 >> 3) * 128;
                    float* const dstRow = reinterpret_cast<float*>(outTile + r * rowBytes) + c0;
                    for (uint32_t c = lane; c < nsym; c += WARP) {
                        float v = 0.f;
                        if (rowPresent) {
                            const uint32_t s4 = ldsU32(tileRow + (c & ~3u));
                            v = ldsF32(cb + (((s4 >> (8 * (c & 3))) & 0xffu) << fieldShift));
                        }
                        if (epi != EPI_WRITE) {
                            v = applyEpilogue(v, dstRow[c], epi, o.divisor);
                        }
                        __stcs(dstRow + c, v);
                    }
                }
            }
            __syncwarp();
        }
        present = presentNext;
        if (TMA) {
            t = t1;
            t1 = t2;
            t2 = t4;
        } else {
            o16 = o16Next;
            units = unitsNext;
            presentNext = presentAfter;
            offLoNext = offLoAfter;
            offHiNext = offHiAfter;
            t = t1;
            t1 = t2;
            t2 = t3;
            t3 = t4;
        }
    }
    if (TMA == 1) {
        mbarWait(mbar, mbarPhase);  // the copies issued for the tile after the last: never leave with them in flight
    }
    // The tile counter is handed back CLEAN: tileCounter[1] counts the warps that are through (every warp of the
    // grid gets here, all of its ticket draws behind it); atomicInc wraps it to 0 for the last one, which resets the
    // ticket word.  Saves the cudaMemsetAsync that used to precede every launch (2-3 us of stream time: 3 % of a
    // 94 us strong-scaling launch at 8 GPUs).
    __syncwarp();
    if (lane == 0) {
        __threadfence();
        if (atomicInc(tileCounter + 1, totalWarps - 1) == totalWarps - 1) {
            atomicExch(tileCounter, 0u);
        }
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
    const uint64_t firstUnit = !present ? 0 : p.recUnits ? static_cast<uint64_t>(row) * p.recUnits : p.rowOff16[row];
    const uint32_t* words = present ? reinterpret_cast<const uint32_t*>(p.blob + firstUnit) : nullptr;
    const uint64_t limitWords =
        !present ? 0 : static_cast<uint64_t>(p.recUnits ? p.recUnits : p.rowOff16[row + 1] - p.rowOff16[row]) * TRAINED_UNIT_WORDS;
    uint64_t bit = 0;
    for (uint32_t c = 0; c < p.dim; ++c) {
        float v = 0.f;
        if (present) {
            const uint64_t pos = bit + p.padBits;  // stream bit k lives at bit k + padBits, LSB-first
            const uint64_t wi = pos >> 5;
            const uint32_t sh = static_cast<uint32_t>(pos & 31);
            const uint32_t w0 = wi < limitWords ? words[wi] : 0u;
            const uint32_t w1 = wi + 1 < limitWords ? words[wi + 1] : 0u;
            const uint32_t window = __brev(__funnelshift_r(w0, w1, sh));
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

// Row-major walk over the n x perRow output elements (16 B units or floats) of K2/K3.
// A block takes 32 consecutive output rows at a time and spreads their elements over its
// threads, so every warp store covers 512 contiguous bytes and (row, column) come from a
// multiply-high instead of a 64-bit division per element.  Elements are handled in batches
// of K per thread in two phases -- `load` issues the (dependent: id -> row -> payload)
// global loads of all K elements, `finish` computes and stores -- because with one element
// in flight per thread the kernels were bound by the two HBM round trips per element
// (r01: 3.0 TB/s of stores at 2048 threads/SM), not by bandwidth.
template <int K, class State, class Load, class Finish>
__device__ __forceinline__ void forEachElement(uint64_t n, uint32_t perRow, Load load, Finish finish)
{
    if (perRow > 1 && perRow < 2048) {
        // floor(t / perRow) == umulhi(t, ceil(2^32 / perRow)) for t < 2^21 (t < 32 * 2048 here)
        const uint32_t inv = 0xffffffffu / perRow + 1;
        for (uint64_t rowBase = static_cast<uint64_t>(blockIdx.x) * 32; rowBase < n;
             rowBase += static_cast<uint64_t>(gridDim.x) * 32) {
            const uint32_t rowsHere = static_cast<uint32_t>(min(static_cast<uint64_t>(32), n - rowBase));
            const uint32_t elems = rowsHere * perRow;
            for (uint32_t t0 = threadIdx.x; t0 < elems; t0 += K * blockDim.x) {
                State st[K];
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const uint32_t t = t0 + k * blockDim.x;
                    if (t < elems) {
                        const uint32_t r = __umulhi(t, inv);
                        st[k] = load(rowBase + r, t - r * perRow);
                    }
                }
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    if (t0 + k * blockDim.x < elems) {
                        finish(st[k]);
                    }
                }
            }
        }
    } else {
        const uint64_t total = n * perRow;
        for (uint64_t e = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; e < total;
             e += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
            const uint64_t i = e / perRow;
            finish(load(i, static_cast<uint32_t>(e - i * perRow)));
        }
    }
}

constexpr int FLAT_BATCH = 4;  // elements in flight per thread in K2/K3

// Load time: (values rows x dim, minmax rows) -> records (UniformParams).  One thread per byte.
__global__ void interleave_uniform_kernel(const uint8_t* __restrict__ values, const float2* __restrict__ minmax,
                                          uint32_t rows, uint32_t dim, uint32_t stride, uint8_t* __restrict__ records)
{
    const uint64_t total = static_cast<uint64_t>(rows) * stride;
    for (uint64_t b = blockIdx.x * static_cast<uint64_t>(blockDim.x) + threadIdx.x; b < total;
         b += static_cast<uint64_t>(gridDim.x) * blockDim.x) {
        const uint64_t row = b / stride;
        const uint32_t off = static_cast<uint32_t>(b - row * stride);
        uint8_t v = 0;
        if (off < 8) {
            v = reinterpret_cast<const uint8_t*>(minmax + row)[off];
        } else if (off - 8 < dim) {
            v = values[row * dim + (off - 8)];
        }
        records[b] = v;
    }
}

// ------------------------------------------------------------ K2 uniform ----
// values: rows x dim bytes (dense, flattened at load from the per-word tables),
// minmax: rows x (min, max).  out[i] = min + (max - min) * float(v) / levels,
// evaluated sub -> mul -> div -> add in binary32 round-to-nearest exactly like
// src/uniform_compression.cpp:68-72 (no FMA contraction, no fast division).
struct UniformParams {
    // one RECORD per row: {float min, float max, dim quantised bytes, zero padding}, `stride` bytes
    // apart (a multiple of 64 for records of 256 B and more, of 16 below).  Interleaved at load:
    // with (min, max) in an array of its own a random row cost a DRAM access of its own, and a
    // 300 B payload row at an odd offset touched 11 sectors (ncu r01, 1 M random ids of the SoA
    // layout: 474 B read per row for 312 B of use).  dim 300: 320 B records, 5 whole 64 B pairs.
    const uint8_t* records;
    uint32_t stride;
    uint32_t rows;
    uint32_t dim;
    float levels;
    float invLevels;  // mode 1: 1 / levels exactly (levels a power of two); mode 2: RN(1 / levels)
    int divMode;      // 0: IEEE divide, 1: exact multiply, 2: constant-divisor FMA sequence (levels == 255)
};

// scaled / levels, correctly rounded (binary32, round to nearest even) in every mode:
//   1  levels = 2^k: x * 2^-k is the same correctly rounded real number as x / 2^k
//   2  levels = 255 (what the format's writer stores for 8 bits, min(1 << bits, 255),
//      src/uniform_compression.cpp:5-8): q0 = x * RN(1/255), one FMA residual, one FMA
//      correction.  Verified EXHAUSTIVELY on the host against x / 255.0f for every float
//      with 1e-20 < |x| < 1e20 and for +0 (scripts/div255_exhaustive.c: 0 mismatches of
//      2.25e9); anything outside that range takes the IEEE divide.
//   0  any other divisor: IEEE divide.
// The IEEE divide sequence made the 8-bit sweep ALU-bound (r01: 57 % of HBM against 80 %
// for the power-of-two widths).
__device__ __forceinline__ float divideByLevels(float x, const float levels, const float invLevels, const int divMode)
{
    if (divMode == 1) {
        return __fmul_rn(x, invLevels);
    }
    if (divMode == 2) {
        const float q0 = __fmul_rn(x, invLevels);
        const float e = __fmaf_rn(-q0, levels, x);
        const float q1 = __fmaf_rn(e, invLevels, q0);
        const float ax = fabsf(x);
        if ((ax > 1e-20f && ax < 1e20f) || __float_as_uint(x) == 0u) {
            return q1;
        }
    }
    return __fdiv_rn(x, levels);
}

__device__ __forceinline__ float uniformDequant(float mn, float range, uint32_t v, const float levels,
                                                const float invLevels, const int divMode)
{
    return __fadd_rn(mn, divideByLevels(__fmul_rn(range, static_cast<float>(v)), levels, invLevels, divMode));
}

struct UniformState {
    float* dst;
    float2 mm;
    uint32_t packed;  // four quantised values (VEC4) or one
    bool present;
};

// DIV >= 0: the division mode (UniformParams::divMode) as a compile-time constant for the
// plain lookup; -1: read it from the parameters.
template <bool VEC4, int EPI, bool SIDE = false, int DIV = -1>
__global__ void uniform_decode_kernel(const UniformParams p, const int32_t* __restrict__ ids, uint32_t rowBegin,
                                      uint64_t n, const OutputSpec o)
{
    const int divMode = DIV >= 0 ? DIV : p.divMode;
    const uint32_t perRow = VEC4 ? p.dim / 4 : p.dim;
    forEachElement<FLAT_BATCH, UniformState>(
        n, perRow,
        [&](uint64_t i, uint32_t c) {
            UniformState st;
            const int64_t row = ids ? static_cast<int64_t>(__ldg(ids + i)) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
            st.present = row >= 0 && row < static_cast<int64_t>(p.rows);
            st.dst = o.out + i * o.rowStride + o.colOffset + (VEC4 ? 4 * c : c);
            if (VEC4 && SIDE) {
                st.dst = reinterpret_cast<float*>(redirectUnit(o, reinterpret_cast<float4*>(st.dst), i, c, perRow - 1));
            }
            st.mm = make_float2(0.f, 0.f);
            st.packed = 0;
            if (st.present) {
                const uint8_t* rec = p.records + static_cast<uint64_t>(row) * p.stride;
                st.packed = VEC4 ? __ldg(reinterpret_cast<const uint32_t*>(rec + 8) + c) : static_cast<uint32_t>(__ldg(rec + 8 + c));
                st.mm = __ldg(reinterpret_cast<const float2*>(rec));
            }
            return st;
        },
        [&](const UniformState& st) {
            const float range = __fsub_rn(st.mm.y, st.mm.x);
            if (VEC4) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                // levels == 255: the constant-divisor sequence is verified for x = +0 and
                // 1e-20 < |x| < 1e20 (divideByLevels).  x = RN(range * k), k = 0..255, stays inside
                // that set for the whole row when range is +0 or 1e-20 < range < 1e20 / 256
                // (rounding is monotone: range <= x <= RN(255 * range) < 1e20 for k >= 1, +0 for
                // k = 0), so the check is made once per unit instead of once per element.
                // (Power-of-two levels, mode 1, need no check at all: x * 2^-k is exact division.)
                const bool wholeRow = divMode == 1 || (divMode == 2 && (__float_as_uint(range) == 0u ||
                                                                             (range > 1e-20f && range < 3.9e17f)));
                if (st.present && wholeRow) {
                    auto one = [&](uint32_t k) {
                        const float x = __fmul_rn(range, static_cast<float>(k));
                        const float q0 = __fmul_rn(x, p.invLevels);
                        if (divMode == 1) {
                            return __fadd_rn(st.mm.x, q0);
                        }
                        const float e = __fmaf_rn(-q0, p.levels, x);
                        return __fadd_rn(st.mm.x, __fmaf_rn(e, p.invLevels, q0));
                    };
                    v.x = one(st.packed & 0xff);
                    v.y = one((st.packed >> 8) & 0xff);
                    v.z = one((st.packed >> 16) & 0xff);
                    v.w = one(st.packed >> 24);
                } else if (st.present) {
                    v.x = uniformDequant(st.mm.x, range, st.packed & 0xff, p.levels, p.invLevels, divMode);
                    v.y = uniformDequant(st.mm.x, range, (st.packed >> 8) & 0xff, p.levels, p.invLevels, divMode);
                    v.z = uniformDequant(st.mm.x, range, (st.packed >> 16) & 0xff, p.levels, p.invLevels, divMode);
                    v.w = uniformDequant(st.mm.x, range, st.packed >> 24, p.levels, p.invLevels, divMode);
                }
                float4* dst = reinterpret_cast<float4*>(st.dst);
                if (EPI != EPI_WRITE) {
                    const float4 old = *dst;
                    v.x = applyEpilogue(v.x, old.x, EPI, o.divisor);
                    v.y = applyEpilogue(v.y, old.y, EPI, o.divisor);
                    v.z = applyEpilogue(v.z, old.z, EPI, o.divisor);
                    v.w = applyEpilogue(v.w, old.w, EPI, o.divisor);
                }
                __stcs(dst, v);
            } else {
                float v = st.present ? uniformDequant(st.mm.x, range, st.packed, p.levels, p.invLevels, divMode) : 0.f;
                if (EPI != EPI_WRITE) {
                    v = applyEpilogue(v, *st.dst, EPI, o.divisor);
                }
                __stcs(st.dst, v);
            }
        });
}

// K2, plain lookup into a dense (n x dim) result (rowStride == dim, no column offset, dim a
// multiple of 4 up to 8188): the form Reader.batch_embedding / decode_range launch.
// ncu r01 of the element-walk kernel above: 95 thread instructions per 16 B unit, 69 % of the
// issue slots at 8 bits, and with random ids every one of a row's 75 units walked the dependent
// chain id -> record.  Here a WARP owns a group of 32 consecutive output rows; the lane of a
// row's number turns its id into the record's word offset a group AHEAD (ids are loaded two
// groups ahead) and leaves it in the warp's double-buffered shared table.  A unit then is
// multiply-high -> LDS -> (min, max) + payload word of the SAME record, issued together ->
// dequant -> one 128-bit streaming store at a LINEAR address (512 contiguous bytes per warp
// store).  No block barrier: warps run their own pipelines, groups are drawn from a ticket
// counter.  Measured on the way (B200, 1 M random ids / 2 M-row full vocabulary, 8 bit):
//   element walk, separate (min, max) array   0.351 / 0.575 ms
//   256-thread groups, __syncthreads          0.338 / 0.561 ms
//   warp groups, static stride                0.296 / 0.532 ms
//   + ticket counter                          0.297 / 0.509 ms
//   + records, (min, max) read a group ahead  0.300 / 0.562 ms  (the early 8 B read opens the
//     record's DRAM page a second time; reading it WITH the payload is what the records are for)
// Same arithmetic, same order as uniform_decode_kernel (bit-identical results).
#ifndef MEMB_UNIFORM_ROWS_BATCH
#define MEMB_UNIFORM_ROWS_BATCH 5
#endif
constexpr int UNIFORM_ROWS_BATCH = MEMB_UNIFORM_ROWS_BATCH;  // units in flight per lane: 15 passes of 5 x 32 cover 32 rows x 75 units
constexpr int UNIFORM_ROWS_WARPS = 8;
#ifndef MEMB_UNIFORM_ROWS_MINBLOCKS
#define MEMB_UNIFORM_ROWS_MINBLOCKS 4
#endif

template <int DIV>
__global__ void __launch_bounds__(UNIFORM_ROWS_WARPS * 32, MEMB_UNIFORM_ROWS_MINBLOCKS)
uniform_rows_kernel(const UniformParams p, const int32_t* __restrict__ ids, uint32_t rowBegin, uint64_t n,
                    float* __restrict__ out, uint32_t* __restrict__ groupCounter)
{
    // first word of the row's record; 0xffffffff: no such row (zeros)
    __shared__ uint32_t infoAll[UNIFORM_ROWS_WARPS][2][32];
    const uint32_t perRow = p.dim / 4;
    const uint32_t inv = 0xffffffffu / perRow + 1;  // floor(t / perRow) == umulhi(t, inv) for t < 2^21
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t(*info)[32] = infoAll[threadIdx.x >> 5];
    const uint64_t groups = (n + 31) / 32;
    const uint64_t step = static_cast<uint64_t>(gridDim.x) * UNIFORM_ROWS_WARPS;
    const uint32_t* __restrict__ words = reinterpret_cast<const uint32_t*>(p.records);
    const uint32_t strideWords = p.stride / 4;

    auto loadId = [&](uint64_t g) -> int64_t {
        const uint64_t i = g * 32 + lane;
        if (g >= groups || i >= n) {
            return -1;
        }
        return ids ? static_cast<int64_t>(__ldg(ids + i)) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
    };
    auto recordWord = [&](int64_t row) {  // records hold < 2^32 words (loader checks)
        return row >= 0 && row < static_cast<int64_t>(p.rows) ? static_cast<uint32_t>(row) * strideWords : 0xffffffffu;
    };

    // group queue of this warp: g (units being written), gNext (ids loaded), gAfter (ids
    // loading); the first three are static, later ones come from a ticket counter
    // (groupCounter, zeroed before the launch) drawn one group before it is needed, exactly as
    // the trained kernel draws its tiles -- with a static stride the last of the ~7 (random
    // 1 M) or ~13 (2 M rows) groups per warp ran on a half-empty GPU.
    uint64_t g = static_cast<uint64_t>(blockIdx.x) * UNIFORM_ROWS_WARPS + (threadIdx.x >> 5);
    if (g >= groups) {
        return;
    }
    uint64_t gNext = g + step, gAfter = gNext + step;
    uint32_t ticket = 0, ticketLane = 0;
    auto drawTicket = [&]() {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "elect.sync %1|p, 0xffffffff;\n"
            "@p atom.global.add.u32 %0, [%2], 1;\n"
            "}\n"
            : "+r"(ticket), "=r"(ticketLane)
            : "l"(groupCounter)
            : "memory");
    };
    drawTicket();
    info[0][lane] = recordWord(loadId(g));
    int64_t rowNext = loadId(gNext);
    __syncwarp();
    uint32_t buf = 0;
    for (; g < groups; buf ^= 1u) {
        const uint64_t gAfter2 = 3 * step + shflFrom(ticket, ticketLane);
        drawTicket();
        const int64_t rowAfter = loadId(gAfter);  // consumed a group later
        const uint64_t rowBase = g * 32;
        const uint32_t rowsHere = static_cast<uint32_t>(min(static_cast<uint64_t>(32), n - rowBase));
        const uint32_t elems = rowsHere * perRow;
        float4* __restrict__ dst = reinterpret_cast<float4*>(out) + rowBase * perRow;
        for (uint32_t t0 = lane; t0 < elems; t0 += UNIFORM_ROWS_BATCH * 32) {
            float2 mm[UNIFORM_ROWS_BATCH];
            uint32_t packed[UNIFORM_ROWS_BATCH];
            bool present[UNIFORM_ROWS_BATCH];
#pragma unroll
            for (int k = 0; k < UNIFORM_ROWS_BATCH; ++k) {
                const uint32_t t = t0 + k * 32;
                packed[k] = 0;
                mm[k] = make_float2(0.f, 0.f);
                present[k] = false;
                if (t < elems) {
                    const uint32_t r = __umulhi(t, inv);
                    const uint32_t w = info[buf][r];
                    present[k] = w != 0xffffffffu;
                    if (present[k]) {
                        mm[k] = __ldg(reinterpret_cast<const float2*>(words + w));
                        packed[k] = __ldg(words + w + 2 + (t - r * perRow));
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < UNIFORM_ROWS_BATCH; ++k) {
                const uint32_t t = t0 + k * 32;
                if (t < elems) {
                    const float mn = mm[k].x, range = __fsub_rn(mm[k].y, mm[k].x);
                    const uint32_t q = packed[k];
                    // the gate of uniform_decode_kernel
                    const bool wholeRow = DIV == 1 || (DIV == 2 && (__float_as_uint(range) == 0u || (range > 1e-20f && range < 3.9e17f)));
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (present[k] && wholeRow) {
                        auto one = [&](uint32_t kk) {
                            const float x = __fmul_rn(range, static_cast<float>(kk));
                            const float q0 = __fmul_rn(x, p.invLevels);
                            if (DIV == 1) {
                                return __fadd_rn(mn, q0);
                            }
                            const float e = __fmaf_rn(-q0, p.levels, x);
                            return __fadd_rn(mn, __fmaf_rn(e, p.invLevels, q0));
                        };
                        v.x = one(q & 0xff);
                        v.y = one((q >> 8) & 0xff);
                        v.z = one((q >> 16) & 0xff);
                        v.w = one(q >> 24);
                    } else if (present[k]) {
                        v.x = uniformDequant(mn, range, q & 0xff, p.levels, p.invLevels, DIV);
                        v.y = uniformDequant(mn, range, (q >> 8) & 0xff, p.levels, p.invLevels, DIV);
                        v.z = uniformDequant(mn, range, (q >> 16) & 0xff, p.levels, p.invLevels, DIV);
                        v.w = uniformDequant(mn, range, q >> 24, p.levels, p.invLevels, DIV);
                    }
                    __stcs(dst + t, v);
                }
            }
        }
        // info[buf ^ 1] was last read a group ago, before the previous __syncwarp
        info[buf ^ 1u][lane] = recordWord(rowNext);
        rowNext = rowAfter;
        g = gNext;
        gNext = gAfter;
        gAfter = gAfter2;
        __syncwarp();
    }
}

// --------------------------------------------------------------- K3 full ----
struct FullParams {
    const float* values;  // rows x dim
    uint32_t rows;
    uint32_t dim;
};

struct FullState {
    float* dst;
    float4 v;  // VEC4: four values; else .x
};

// KEEP: the loads ask L2 to keep what they touch (evict_last) -- the copy-out of the distinct-tuples path, whose
// source is far larger than L2 but whose frequent rows carry most of the reads while the output streams through.
template <bool VEC4, int EPI, bool SIDE = false, bool KEEP = false>
__global__ void full_gather_kernel(const FullParams p, const int32_t* __restrict__ ids, uint32_t rowBegin,
                                   uint64_t n, const OutputSpec o)
{
    const uint32_t perRow = VEC4 ? p.dim / 4 : p.dim;
    uint64_t keep = 0;
    if (KEEP) {
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
    }
    forEachElement<FLAT_BATCH, FullState>(
        n, perRow,
        [&](uint64_t i, uint32_t c) {
            FullState st;
            const int64_t row = ids ? static_cast<int64_t>(__ldg(ids + i)) : static_cast<int64_t>(rowBegin) + static_cast<int64_t>(i);
            const bool present = row >= 0 && row < static_cast<int64_t>(p.rows);
            st.dst = o.out + i * o.rowStride + o.colOffset + (VEC4 ? 4 * c : c);
            if (VEC4 && SIDE) {
                st.dst = reinterpret_cast<float*>(redirectUnit(o, reinterpret_cast<float4*>(st.dst), i, c, perRow - 1));
            }
            st.v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (present) {
                const float* src = p.values + static_cast<uint64_t>(row) * p.dim;
                if (VEC4 && KEEP) {
                    asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
                                 : "=f"(st.v.x), "=f"(st.v.y), "=f"(st.v.z), "=f"(st.v.w)
                                 : "l"(reinterpret_cast<const float4*>(src) + c), "l"(keep));
                } else if (VEC4) {
                    st.v = __ldg(reinterpret_cast<const float4*>(src) + c);
                } else {
                    st.v.x = __ldg(src + c);
                }
            }
            return st;
        },
        [&](const FullState& st) {
            if (VEC4) {
                float4 v = st.v;
                float4* dst = reinterpret_cast<float4*>(st.dst);
                if (EPI != EPI_WRITE) {
                    const float4 old = *dst;
                    v.x = applyEpilogue(v.x, old.x, EPI, o.divisor);
                    v.y = applyEpilogue(v.y, old.y, EPI, o.divisor);
                    v.z = applyEpilogue(v.z, old.z, EPI, o.divisor);
                    v.w = applyEpilogue(v.w, old.w, EPI, o.divisor);
                }
                __stcs(dst, v);
            } else {
                float v = st.v.x;
                if (EPI != EPI_WRITE) {
                    v = applyEpilogue(v, *st.dst, EPI, o.divisor);
                }
                __stcs(st.dst, v);
            }
        });
}

// ------------------------------------------------------ repeated queries ----
// Text is Zipf-distributed: a batch of 4 M words over a 2.2 M-word vocabulary (BASELINE configs[4]) holds
// ~0.7 M DISTINCT words.  A ReadersUnion lookup of such a batch decodes every distinct id tuple once into a
// temporary and copies the rows out (full_gather_kernel: the frequent rows come from L2), instead of running
// the Huffman chain 4 M x R times.  The result is the same bytes: the merge of a tuple does not depend on where
// in the batch it stands.  The reference has nothing like it (it decodes every word: src/reader.cpp:41-47).
//
// dedup_claim_kernel: open-addressing table over the batch's id tuples (slot = index of the first query that
// claimed it, -1 free).  rep[i] = the query whose tuple equals query i's and owns the slot; an owner also takes
// the next compact index (warp-aggregated counter) and leaves its tuple in distinctIds.  A plain load in front of
// the CAS keeps the 260 k copies of the most frequent word from queueing on one atomic.
// A slot changes once (free -> claimed), so a claimed value seen through L1 is final and a stale "free" only costs
// the CAS that tells the truth.  Through L1 on purpose: the slot of the batch's most frequent word is read by
// 260 k queries, and uncached (volatile) loads of ONE sector queue on its L2 slice.
__device__ __forceinline__ int32_t ldCachedI32(const int32_t* p)
{
    int32_t v;
    asm volatile("ld.global.ca.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ldCachedU64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.global.ca.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
// Compact index of an owner: ONE atomicAdd on the batch's counter per BLOCK (a warp-aggregated atomic per warp was
// 125 k atomics on one address for 4 M Zipf words -- 0.19 of the claim kernel's 0.21 ms, ncu launch list r02).
// Every thread of the block calls it (two __syncthreads); returns the index for owners, junk otherwise.
constexpr uint32_t DEDUP_BLOCK = 512;
__device__ __forceinline__ uint32_t claimCompactIndex(bool owner, uint32_t* count)
{
    __shared__ uint32_t warpBase[DEDUP_BLOCK / WARP];
    __shared__ uint32_t blockBase;
    const uint32_t lane = threadIdx.x % WARP, warp = threadIdx.x / WARP;
    const uint32_t owners = __ballot_sync(0xffffffffu, owner);
    if (lane == 0) {
        warpBase[warp] = __popc(owners);
    }
    __syncthreads();
    if (warp == 0) {
        const uint32_t mine = lane < blockDim.x / WARP ? warpBase[lane] : 0u;
        uint32_t incl = mine;
#pragma unroll
        for (uint32_t d = 1; d < WARP; d *= 2) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, incl, d);
            incl += lane >= d ? up : 0u;
        }
        if (lane < blockDim.x / WARP) {
            warpBase[lane] = incl - mine;
        }
        if (lane == WARP - 1 && incl) {
            blockBase = atomicAdd(count, incl);
        }
    }
    __syncthreads();
    return blockBase + warpBase[warp] + __popc(owners & ((1u << lane) - 1u));
}
__device__ __forceinline__ uint64_t tupleHash(const int32_t* __restrict__ ids, uint64_t idStride, uint32_t R, uint64_t i)
{
    uint64_t h = 0x9e3779b97f4a7c15ull;
    for (uint32_t k = 0; k < R; ++k) {
        h = mix64(h ^ static_cast<uint32_t>(__ldg(ids + k * idStride + i)));
    }
    return h;
}
__global__ void dedup_claim_kernel(const int32_t* __restrict__ ids, uint64_t idStride, uint32_t R, uint32_t n,
                                   int32_t* slots, uint32_t slotMask, int32_t* __restrict__ rep,
                                   int32_t* __restrict__ compact, int32_t* __restrict__ distinctIds,
                                   uint64_t distinctStride, uint32_t* count)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool owner = false;
    if (i < n) {
        uint32_t slot = static_cast<uint32_t>(tupleHash(ids, idStride, R, i)) & slotMask;
        int32_t mine = static_cast<int32_t>(i);
        for (;;) {
            int32_t cur = ldCachedI32(slots + slot);
            if (cur < 0) {
                cur = atomicCAS(slots + slot, -1, static_cast<int32_t>(i));
                if (cur < 0) {
                    owner = true;
                    break;
                }
            }
            bool same = true;
            for (uint32_t k = 0; k < R; ++k) {
                same = same && __ldg(ids + k * idStride + cur) == __ldg(ids + k * idStride + i);
            }
            if (same) {
                mine = cur;
                break;
            }
            slot = (slot + 1) & slotMask;
        }
        rep[i] = mine;
    }
    const uint32_t c = claimCompactIndex(owner, count);
    if (owner) {
        compact[i] = static_cast<int32_t>(c);
        for (uint32_t k = 0; k < R; ++k) {
            distinctIds[k * distinctStride + c] = __ldg(ids + k * idStride + i);
        }
    }
}
// One or two readers (every ReadersUnion of the benchmark): the tuple itself is the 64-bit key of the slot, so a
// probe is ONE random 8-byte access instead of slot + the owner's ids (r02, 4 M Zipf words: 0.22 ms for the
// generic kernel).  Ids are normalised first (anything outside [0, rows) is "missing" = -1: the same zero row),
// which also keeps the key (id + 1 per half) away from the all-ones EMPTY.  slotOf[i] = the tuple's slot, with
// bit 31 set for the query that owns it; compactOfSlot[slot] = its compact index, read by dedup_follow64_kernel.
__global__ void dedup_claim64_kernel(const int32_t* __restrict__ ids, uint64_t idStride, uint32_t R, uint32_t n,
                                     uint32_t rows0, uint32_t rows1, unsigned long long* slots, uint32_t slotMask,
                                     uint32_t* __restrict__ slotOf, int32_t* __restrict__ compactOfSlot,
                                     int32_t* __restrict__ compact, int32_t* __restrict__ distinctIds,
                                     uint64_t distinctStride, uint32_t* count)
{
    constexpr unsigned long long EMPTY = ~0ull;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool owner = false;
    uint32_t slot = 0;
    int32_t id0 = -1, id1 = -1;
    if (i < n) {
        id0 = __ldg(ids + i);
        id0 = static_cast<uint32_t>(id0) < rows0 ? id0 : -1;
        if (R > 1) {
            id1 = __ldg(ids + idStride + i);
            id1 = static_cast<uint32_t>(id1) < rows1 ? id1 : -1;
        }
        const unsigned long long key = (static_cast<unsigned long long>(static_cast<uint32_t>(id1 + 1)) << 32) |
                                       static_cast<uint32_t>(id0 + 1);
        slot = static_cast<uint32_t>(mix64(key)) & slotMask;
        for (;;) {
            unsigned long long cur = ldCachedU64(slots + slot);
            if (cur == EMPTY) {
                cur = atomicCAS(slots + slot, EMPTY, key);
                if (cur == EMPTY) {
                    owner = true;
                    break;
                }
            }
            if (cur == key) {
                break;
            }
            slot = (slot + 1) & slotMask;
        }
    }
    const uint32_t c = claimCompactIndex(owner, count);
    if (i < n) {
        slotOf[i] = slot | (owner ? 0x80000000u : 0u);
        if (owner) {
            compactOfSlot[slot] = static_cast<int32_t>(c);
            compact[i] = static_cast<int32_t>(c);
            distinctIds[c] = id0;
            if (R > 1) {
                distinctIds[distinctStride + c] = id1;
            }
        }
    }
}
__global__ void dedup_follow64_kernel(const uint32_t* __restrict__ slotOf, const int32_t* __restrict__ compactOfSlot,
                                      int32_t* __restrict__ compact, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const uint32_t s = slotOf[i];
        if (!(s & 0x80000000u)) {
            compact[i] = compactOfSlot[s];
        }
    }
}
// every other query takes the compact index of the query that stands for it
__global__ void dedup_follow_kernel(const int32_t* __restrict__ rep, int32_t* compact, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        const int32_t r = rep[i];
        if (r != static_cast<int32_t>(i)) {
            compact[i] = compact[r];
        }
    }
}

// The copy-out of the distinct-tuples path for the average of two trained readers with small codebooks: the temporary
// holds the distinct tuples' SYMBOL words of both readers (two SYM = 1 launches: dim bytes per row and reader instead
// of 4 * dim of fp32), and this kernel turns them into rows on the way out -- out[i] = T[symA[index[i]]][symB[index[i]]]
// with the per-bank pair table of the merged flush (same arithmetic: bit-identical).  Half the bytes to read per
// output row, twice the rows of the temporary in L2, and the distinct tuples never go through the fp32 flush.
struct PairGatherState {
    float4* dst;
    uint32_t a, b;
};
__global__ void __launch_bounds__(256) pair_gather_kernel(const uint32_t* __restrict__ symA, const uint32_t* __restrict__ symB,
                                                          const int32_t* __restrict__ index, uint64_t n, uint32_t words,
                                                          const float* __restrict__ codebook0, uint32_t size0, uint32_t fieldShift0,
                                                          const float* __restrict__ codebook1, uint32_t size1, uint32_t fieldShift1,
                                                          float divisor, float4* __restrict__ out)
{
    extern __shared__ __align__(128) unsigned char pairSmem[];
    float* table = reinterpret_cast<float*>(pairSmem);
    const uint32_t stride = size1 + 1;
    const uint32_t entries = (size0 + 1) * stride;
    for (uint32_t i = threadIdx.x; i < entries; i += blockDim.x) {
        const uint32_t s0 = i / stride, s1 = i % stride;
        const float c0 = s0 < size0 ? codebook0[s0] : 0.f;
        const float c1 = s1 < size1 ? codebook1[s1] : 0.f;
        const float t = applyEpilogue(c1, c0, EPI_FINALIZE, divisor);
        float4* dst = reinterpret_cast<float4*>(table + i * WARP);
#pragma unroll
        for (uint32_t q = 0; q < WARP / 4; ++q) {
            dst[q] = make_float4(t, t, t, t);
        }
    }
    __syncthreads();
    const uint32_t pairBase = smemAddr(table) + (threadIdx.x % WARP) * 4u;  // lane l reads bank l
    // a field is symbol << (2 - fieldShift): byte offset of entry (s0, s1) = (s0 * stride + s1) * 128
    const uint32_t m0 = (stride * WARP) << fieldShift0, m1 = WARP << fieldShift1;
    forEachElement<FLAT_BATCH, PairGatherState>(
        n, words,
        [&](uint64_t i, uint32_t c) {
            PairGatherState st;
            const uint64_t row = static_cast<uint64_t>(__ldg(index + i)) * words + c;
            st.a = __ldg(symA + row);
            st.b = __ldg(symB + row);
            st.dst = out + i * words + c;
            return st;
        },
        [&](const PairGatherState& st) { __stcs(st.dst, gatherPairsWide(pairBase, st.a, st.b, m0, m1)); });
}

}  // namespace kernels
}  // namespace memb_b200
