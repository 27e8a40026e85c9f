// K1 (second design) and K4 in ONE launch: `trained_rows_kernel`.
//
// Decodes whole rows of 1..4 trained readers per output row and merges them on the way out
// (python/memb/readers_union.py:16-18,30-32), replacing
//   src/trained_compression.cpp:129-136, src/huffman_table_decoder.h:102-118,
//   src/bit_stream_reader.h:16-31 (and np.mean / np.concatenate for a union).
//
// What ncu said about the first design (trained_decode_kernel, kernels.cuh; profiles/r01_k1_*):
// bound by the shared-memory pipe (945 shared-memory instructions per 32-row tile, one LUT
// load per symbol on the critical path) and by the DRAM efficiency of a flush that writes 32
// rows 128 B at a time.  This kernel changes both:
//   * MULTI-SYMBOL LUT: an entry decodes up to three symbols that fit the 9-bit window
//     (4-bit table: 2.23 symbols per lookup, 300 -> ~135 LUT loads and half the window
//     reloads per row).  Codes longer than the window go through a SECOND-LEVEL table in
//     shared memory indexed by the following bits -- the scheme of
//     src/huffman_table_decoder.h:58-93 -- instead of a bit-serial escape.
//   * Lanes run free (a lane decodes its whole row at its own symbol rate) into a WHOLE-ROW
//     symbol tile, and the warp then writes the tile's 32 rows as ONE contiguous run
//     (38.4 KB at dim 300): 512 contiguous bytes per store instruction.
//   * Rows live in fixed-stride RECORDS (no offset table: one dependent HBM round trip and a
//     64 B over-fetch per row less), loaded a tile ahead.
//   * LUT images (bank-replicated) are built once at load; a CTA pulls them with
//     cp.async.bulk + mbarrier instead of re-replicating them in every launch.
//   * R readers: the warp decodes reader 0, 1, .. of its 32 output rows back to back into R
//     symbol tiles and flushes once: average = centroids of the R codebooks added in reader
//     order, then / R; concatenate = the union row written as one contiguous run.  No symbol
//     words, no read-modify-write of the output, no stitch pass.
#pragma once

#include "kernels.cuh"

namespace memb_b200 {
namespace kernels {

constexpr uint32_t ROWS_MAX_READERS = 4;
constexpr uint32_t ROWS_MAX_WARPS = 12;
constexpr uint32_t ROWS_PREFETCH_UNITS = 12;  // 16 B units of the next row held in registers
constexpr uint32_t ROWS_PAD_BITS = 7;         // stream bit k of a record = bit k + 7 (LSB-first): LUT index = window & mask
constexpr uint32_t ROWS_MAX_REGS = 168;
constexpr uint32_t ROWS_LUT2_MAX_ENTRIES = 2048;
enum RowsMode : int { ROWS_CONCAT = 0, ROWS_AVERAGE = 1 };  // one reader: ROWS_CONCAT

// LUT entry (first and second level):
//   bits 0..4   stream bits the entry consumes
//   bits 5..6   number of symbols (1..3); 0 = escape: bits 8..11 = extra bits to index the
//               second level with, bits 12..31 = first entry of the sub-table
//   bits 8..31  the symbols' fields, first symbol in bits 8..15 (unused ones are 0)
struct RowsReader {
    const RowUnit* blob;       // records: row r at blob + r * recUnits
    const uint32_t* lutImage;  // (1 << lutBits) x 32 bank replicas, then lut2, then 256 codebook floats (one bulk copy each)
    uint32_t recUnits;
    uint32_t rows;
    uint32_t dim;
    uint32_t lutBits;
    uint32_t lut2Entries;   // padded to a multiple of 4
    uint32_t zeroWord;      // symbol field of the codebook's +0.0 slot, four times
    uint32_t fieldShift;    // symbol field << fieldShift = byte offset into the codebook (0 or 2)
    // shared-window addresses (host computed, see rowsLayout)
    uint32_t sLut, sLut2, sCodebook;
    uint32_t tileStride;    // bytes between rows of the symbol tile (odd number of words)
    uint32_t tileOffset;    // this reader's tile within a warp's tile block
    uint32_t colUnit;       // first 16 B unit of this reader's columns in the output row
};

struct RowsParams {
    RowsReader rd[ROWS_MAX_READERS];
    uint32_t sMbar;
    uint32_t sCols[ROWS_MAX_WARPS];   // per warp: staged records, word w of lane l at + w * 128 + l * 4
    uint32_t sTiles[ROWS_MAX_WARPS];  // per warp: symbol tiles
    uint32_t colWords;                // words reserved per staged record (max recUnits * 4 + 4)
    uint32_t outUnits;                // 16 B units per output row
    uint32_t smemEnd;                 // first shared-window byte not used
    float divisor;
};

struct RowsLane {
    uint32_t lo, hi;     // window: bit 0 of lo = bit `consumed` of the record
    uint32_t consumed;
    uint32_t accLo, accHi, fill8;  // symbol fields not yet stored: fill8 / 8 of them, first in bits 0..7 of accLo
    uint32_t mul;        // 1 << fill8
    uint32_t tilePtr;    // next word of the lane's tile row
};

__device__ __forceinline__ void rowsReload(RowsLane& d, uint32_t colLane, uint32_t lastWindow)
{
    // a lane that has finished its row keeps decoding whatever follows; it must not leave its columns
    const uint32_t a = colLane + min((d.consumed << 2) & 0x3ff80u, lastWindow);
    const uint32_t w0 = ldsU32At<0>(a), w1 = ldsU32At<128>(a), w2 = ldsU32At<256>(a);
    d.lo = __funnelshift_r(w0, w1, d.consumed);
    d.hi = __funnelshift_r(w1, w2, d.consumed);
}

__device__ __forceinline__ void stsU32If(bool pred, uint32_t addr, uint32_t v)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.u32 p, %2, 0;\n"
        "@p st.shared.u32 [%0], %1;\n"
        "}\n" ::"r"(addr),
        "r"(v), "r"(static_cast<uint32_t>(pred))
        : "memory");
}

// One LUT lookup: up to three symbols leave the window and join the pending word; a full
// word goes to the tile row (until the row is complete).  Straight-line: a divergent branch
// around the store cost a BSSY / BRA / BSYNC per lookup (first version: 48 K cycles per tile).
template <bool ESC>
__device__ __forceinline__ void rowsStep(RowsLane& d, uint32_t lutLane, uint32_t lutMask, uint32_t lut2, uint32_t padL,
                                         uint32_t tileEnd)
{
    uint32_t e = ldsU32((d.lo & lutMask) | lutLane);
    if (ESC) {
        if ((e & 0x60u) == 0u) {
            const uint32_t extra = (e >> 8) & 15u;
            const uint32_t idx = (e >> 12) + ((d.lo >> padL) & ((1u << extra) - 1u));
            e = ldsU32(lut2 + idx * 4u);
        }
    }
    d.lo = __funnelshift_r(d.lo, d.hi, e);  // shift = e & 31
    d.hi = __funnelshift_r(d.hi, 0u, e);
    d.consumed += e & 31u;
    // the (up to three) symbol fields join the pending 64-bit word at byte fill8 / 8.  The bits
    // never overlap, so "or in, shifted" is a multiply-add by 1 << fill8: two IMADs on the FMA
    // pipe instead of two shifts and two ORs on the ALU pipe, which ncu showed to be the busy one
    // (62 % with math_pipe_throttle stalls, r02)
    const uint32_t syms = e >> 8;
    d.accLo = syms * d.mul + d.accLo;
    d.accHi = __umulhi(syms, d.mul) + d.accHi;
    const uint32_t fill = d.fill8 + ((e >> 2) & 0x18u);  // <= 24 + 24
    asm volatile(
        "{\n"
        ".reg .pred full, store;\n"
        "setp.ge.u32 full, %3, 32;\n"
        "setp.lt.and.u32 store, %2, %4, full;\n"
        "@store st.shared.u32 [%2], %0;\n"
        "@full mov.u32 %0, %1;\n"
        "@full mov.u32 %1, 0;\n"
        "@full add.u32 %2, %2, 4;\n"
        "}\n"
        : "+r"(d.accLo), "+r"(d.accHi), "+r"(d.tilePtr)
        : "r"(fill), "r"(tileEnd)
        : "memory");
    d.fill8 = fill & 31u;
    d.mul = 1u << d.fill8;
}

// Cursor of a lane over the tile's output units: unit (rr, seg, c) = 16 B unit c of reader
// seg's columns in tile row rr.  FAST: every reader has >= 32 units, so a step of 32 units
// wraps at most once.
template <int R, bool FAST>
struct RowsCursor {
    uint32_t rr, seg, c;
    __device__ __forceinline__ void wrap(const RowsParams& P)
    {
        const uint32_t U = P.rd[R == 1 ? 0 : seg].dim >> 2;
        if (c >= U) {
            c -= U;
            if (R == 1 || ++seg == R) {
                seg = 0;
                ++rr;
            }
        }
    }
    __device__ __forceinline__ void init(const RowsParams& P, uint32_t lane)
    {
        rr = seg = 0;
        c = lane;
        while (c >= (P.rd[R == 1 ? 0 : seg].dim >> 2)) {
            wrap(P);
        }
    }
    __device__ __forceinline__ void advance(const RowsParams& P)
    {
        c += WARP;
        if (FAST) {
            wrap(P);
        } else {
            while (c >= (P.rd[R == 1 ? 0 : seg].dim >> 2)) {
                wrap(P);
            }
        }
    }
};

// concatenate (and the single reader): the tile's rows as one contiguous run of the output.
// Every lane owns units lane, lane + 32, ...: at least units / 32 of them, which go in
// straight-line batches (all symbol words, then all centroids, then the stores); the rest
// of a ragged tile one by one.
template <int R, bool SMALLCB, bool FAST>
__device__ __forceinline__ void rowsFlushConcat(const RowsParams& P, uint32_t tiles0, uint32_t units, float4* dst,
                                                uint32_t lane)
{
    constexpr int BATCH = 5;
    RowsCursor<R, FAST> cur;
    cur.init(P, lane);
    uint32_t u = lane;
    for (uint32_t b = (units / WARP) / BATCH; b; --b) {
        uint32_t s4[BATCH], cbs[BATCH], fs[BATCH];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            const RowsReader& rd = P.rd[R == 1 ? 0 : cur.seg];
            cbs[k] = rd.sCodebook;
            fs[k] = rd.fieldShift;
            s4[k] = ldsU32(tiles0 + rd.tileOffset + cur.rr * rd.tileStride + cur.c * 4u);
            cur.advance(P);
        }
        float4 v[BATCH];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            v[k] = SMALLCB ? gatherCentroids<true>(cbs[k], s4[k]) : gatherCentroidsShift(cbs[k], s4[k], fs[k]);
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            stcsF4(dst + u + k * WARP, v[k]);
        }
        u += BATCH * WARP;
    }
    for (; u < units; u += WARP) {
        const RowsReader& rd = P.rd[R == 1 ? 0 : cur.seg];
        const uint32_t s4 = ldsU32(tiles0 + rd.tileOffset + cur.rr * rd.tileStride + cur.c * 4u);
        stcsF4(dst + u, SMALLCB ? gatherCentroids<true>(rd.sCodebook, s4) : gatherCentroidsShift(rd.sCodebook, s4, rd.fieldShift));
        cur.advance(P);
    }
}

// average: centroids of the R codebooks added in reader order, then / R (np.mean(axis=0))
template <int R, bool SMALLCB>
__device__ __forceinline__ float4 rowsAverageUnit(const RowsParams& P, const uint32_t (&s4)[R])
{
    float4 sum;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const RowsReader& rd = P.rd[r];
        const float4 v = SMALLCB ? gatherCentroids<true>(rd.sCodebook, s4[r]) : gatherCentroidsShift(rd.sCodebook, s4[r], rd.fieldShift);
        if (r == 0) {
            sum = v;
        } else {
            sum.x = __fadd_rn(sum.x, v.x);
            sum.y = __fadd_rn(sum.y, v.y);
            sum.z = __fadd_rn(sum.z, v.z);
            sum.w = __fadd_rn(sum.w, v.w);
        }
    }
    sum.x = divideByReaders(sum.x, P.divisor);
    sum.y = divideByReaders(sum.y, P.divisor);
    sum.z = divideByReaders(sum.z, P.divisor);
    sum.w = divideByReaders(sum.w, P.divisor);
    return sum;
}

template <int R, bool SMALLCB, bool FAST>
__device__ __forceinline__ void rowsFlushAverage(const RowsParams& P, uint32_t tiles0, uint32_t units, float4* dst,
                                                 uint32_t lane)
{
    constexpr int BATCH = 3;
    RowsCursor<1, FAST> cur;  // all readers have the same dim
    cur.init(P, lane);
    uint32_t u = lane;
    for (uint32_t b = (units / WARP) / BATCH; b; --b) {
        uint32_t s4[BATCH][R];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                s4[k][r] = ldsU32(tiles0 + P.rd[r].tileOffset + cur.rr * P.rd[r].tileStride + cur.c * 4u);
            }
            cur.advance(P);
        }
        float4 v[BATCH];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            v[k] = rowsAverageUnit<R, SMALLCB>(P, s4[k]);
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
            stcsF4(dst + u + k * WARP, v[k]);
        }
        u += BATCH * WARP;
    }
    for (; u < units; u += WARP) {
        uint32_t s4[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            s4[r] = ldsU32(tiles0 + P.rd[r].tileOffset + cur.rr * P.rd[r].tileStride + cur.c * 4u);
        }
        stcsF4(dst + u, rowsAverageUnit<R, SMALLCB>(P, s4));
        cur.advance(P);
    }
}

// G lookups per window reload: (G - 1) * maxConsumed + maxPeek <= 64 - ROWS_PAD_BITS.
// SMALLCB: every reader's symbol field already is the codebook byte offset (gatherCentroids).
template <int R, int MODE, int G, bool ESC, bool SMALLCB>
__global__ void __maxnreg__(ROWS_MAX_REGS)
trained_rows_kernel(const RowsParams P, const int32_t* __restrict__ ids, uint64_t idStride, uint32_t rowBegin, uint64_t n,
                    float* __restrict__ out, uint32_t* __restrict__ tileCounter)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t lane = threadIdx.x % WARP;
    const uint32_t warp = threadIdx.x / WARP;
    const uint32_t warpsPerCta = blockDim.x / WARP;
    {
        uint32_t dynBytes;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dynBytes));
        if (P.smemEnd - smemAddr(smem) > dynBytes) {
            __trap();  // the host laid the shared window out for another base
        }
    }
    // ---- tables: bulk copies of the images built at load ----
    if (threadIdx.x == 0) {
        mbarInit(P.sMbar, 1);
        uint32_t total = 0;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            total += (128u << P.rd[r].lutBits) + P.rd[r].lut2Entries * 4u + 1024u;
        }
        mbarExpectTx(P.sMbar, total);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const RowsReader& rd = P.rd[r];
            const uint32_t lutBytes = 128u << rd.lutBits;
            const unsigned char* src = reinterpret_cast<const unsigned char*>(rd.lutImage);
            for (uint32_t off = 0; off < lutBytes; off += 16384u) {
                bulkLoad(rd.sLut + off, src + off, min(16384u, lutBytes - off), P.sMbar);
            }
            if (rd.lut2Entries) {
                bulkLoad(rd.sLut2, src + lutBytes, rd.lut2Entries * 4u, P.sMbar);
            }
            bulkLoad(rd.sCodebook, src + lutBytes + rd.lut2Entries * 4u, 1024u, P.sMbar);
        }
    }
    __syncthreads();  // the barrier is initialised before anybody waits on it

    const uint32_t colLane = P.sCols[warp] + lane * 4u;
    const uint32_t tiles0 = P.sTiles[warp];
    const uint32_t lastWindow = (P.colWords - 3u) * 128u;
    const uint32_t numTiles = static_cast<uint32_t>((n + WARP - 1) / WARP);
    const uint32_t totalWarps = gridDim.x * warpsPerCta;
    bool fastCursor = true;  // every reader's columns span >= 32 units: a 32-unit step wraps at most once
#pragma unroll
    for (int r = 0; r < R; ++r) {
        fastCursor = fastCursor && P.rd[r].dim >= 4 * WARP;
    }

    auto fetchIds = [&](uint32_t tt, uint32_t (&dst)[R]) {
        const uint64_t i = static_cast<uint64_t>(tt) * WARP + lane;
        const bool has = tt < numTiles && i < n;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            // 0xffffffff (no row) also stands for negative and out-of-range ids
            dst[r] = !has ? 0xffffffffu
                          : (ids ? static_cast<uint32_t>(__ldg(ids + r * idStride + i)) : rowBegin + static_cast<uint32_t>(i));
        }
    };
    RowUnit pre[ROWS_PREFETCH_UNITS];
    auto prefetch = [&](const RowsReader& rd, uint32_t row) {
        const RowUnit* src = rd.blob + static_cast<uint64_t>(row < rd.rows ? row : 0u) * rd.recUnits;
#pragma unroll
        for (uint32_t k = 0; k < ROWS_PREFETCH_UNITS; ++k) {
            if (k < rd.recUnits) {
                pre[k] = loadRowUnit(src + k);
            }
        }
    };

    uint32_t t = blockIdx.x * warpsPerCta + warp;
    uint32_t t1 = t + totalWarps, t2 = t1 + totalWarps;
    uint32_t ticket = 0, ticketLane = 0;
    auto drawTicket = [&]() {
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
    uint32_t idCur[R], idNext[R], idAfter[R];
    fetchIds(t, idCur);
    fetchIds(t1, idNext);
    prefetch(P.rd[0], idCur[0]);
    mbarWait(P.sMbar, 0);

    while (t < numTiles) {
        const uint32_t t3 = 3u * totalWarps + shflFrom(ticket, ticketLane);
        drawTicket();
        const uint64_t row0 = static_cast<uint64_t>(t) * WARP;
        const uint32_t rowsHere = static_cast<uint32_t>(min(static_cast<uint64_t>(WARP), n - row0));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const RowsReader& rd = P.rd[r];
            // ---- stage: own record -> own bank ----
#pragma unroll
            for (uint32_t k = 0; k < ROWS_PREFETCH_UNITS; ++k) {
                if (k < rd.recUnits) {
                    stageRowUnit(colLane + k * (TRAINED_UNIT_WORDS * 128), pre[k]);
                }
            }
            if (rd.recUnits > ROWS_PREFETCH_UNITS) {
                // records longer than the prefetch window: the rest comes straight from HBM
                const uint32_t row = idCur[r];
                const RowUnit* src = rd.blob + static_cast<uint64_t>(row < rd.rows ? row : 0u) * rd.recUnits;
                for (uint32_t j = ROWS_PREFETCH_UNITS; j < rd.recUnits; j += 4) {
                    RowUnit c[4];
#pragma unroll
                    for (uint32_t k = 0; k < 4; ++k) {
                        if (j + k < rd.recUnits) {
                            c[k] = loadRowUnit(src + j + k);
                        }
                    }
#pragma unroll
                    for (uint32_t k = 0; k < 4; ++k) {
                        if (j + k < rd.recUnits) {
                            stageRowUnit(colLane + (j + k) * (TRAINED_UNIT_WORDS * 128), c[k]);
                        }
                    }
                }
            }
            // ---- next record into registers; ids of the tile after next ----
            if (r + 1 < R) {
                prefetch(P.rd[(r + 1) % R], idCur[(r + 1) % R]);
            } else {
                prefetch(P.rd[0], idNext[0]);
            }
            if (r == 0) {
                fetchIds(t2, idAfter);
            }
            // ---- decode the whole row into the lane's tile row ----
            const bool present = idCur[r] < rd.rows;
            RowsLane d;
            d.consumed = 0;
            d.accLo = d.accHi = d.fill8 = 0;
            d.mul = 1;
            const uint32_t tileRow = tiles0 + rd.tileOffset + lane * rd.tileStride;
            // a lane without a row (missing word) decodes nothing and fills its tile row with the
            // symbol of the codebook's +0.0 slot: the flush needs no per-row predicate
            const uint32_t tileEnd = tileRow + (present ? rd.dim : 0u);  // dim symbols = dim bytes
            d.tilePtr = tileRow;
            uint32_t lutLane = rd.sLut | (lane * 4u);
            uint32_t lutMask = ((1u << rd.lutBits) - 1u) << ROWS_PAD_BITS;
            asm volatile("" : "+r"(lutLane), "+r"(lutMask));  // one register each: the lookup address is ONE LOP3
            const uint32_t padL = ROWS_PAD_BITS + rd.lutBits;
            __syncwarp();
            do {
                rowsReload(d, colLane, lastWindow);
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    rowsStep<ESC>(d, lutLane, lutMask, rd.sLut2, padL, tileEnd);
                }
            } while (__any_sync(0xffffffffu, d.tilePtr < tileEnd));
            if (!present) {
#pragma unroll 1
                for (uint32_t w = 0; w < rd.dim; w += 4) {
                    stsU32(tileRow + w, rd.zeroWord);
                }
            }
            __syncwarp();
        }
        // ---- flush: the tile's rows as one contiguous run of the output ----
        {
            const uint32_t units = rowsHere * P.outUnits;
            float4* const dst = reinterpret_cast<float4*>(out) + row0 * P.outUnits;
            if (MODE == ROWS_AVERAGE) {
                if (fastCursor) {
                    rowsFlushAverage<R, SMALLCB, true>(P, tiles0, units, dst, lane);
                } else {
                    rowsFlushAverage<R, SMALLCB, false>(P, tiles0, units, dst, lane);
                }
            } else if (fastCursor) {
                rowsFlushConcat<R, SMALLCB, true>(P, tiles0, units, dst, lane);
            } else {
                rowsFlushConcat<R, SMALLCB, false>(P, tiles0, units, dst, lane);
            }
        }
        __syncwarp();
#pragma unroll
        for (int r = 0; r < R; ++r) {
            idCur[r] = idNext[r];
            idNext[r] = idAfter[r];
        }
        t = t1;
        t1 = t2;
        t2 = t3;
    }
}

}  // namespace kernels
}  // namespace memb_b200
