// Host side of the C ABI (include/memb_b200.h): file -> HBM loader, kernel
// launchers, ReadersUnion fusion, pinned-memory helpers and the Builder glue.
// The only compute outside this translation unit's kernels is the one-time host
// parse; there is deliberately no CPU decode path in the product.
#include "../../include/memb_b200.h"

#include "builder.h"
#include "fb_format.h"
#include "kernels.cuh"
#include "trained_rows.cuh"
#include "memb_file.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

using namespace memb_b200;
namespace K = memb_b200::kernels;

namespace {

thread_local std::string g_lastError;
std::atomic<uint64_t> g_launches{0};
std::atomic<int64_t> g_optLutReplicas{0};
std::atomic<int64_t> g_optWarps{0};
std::atomic<int64_t> g_optGeneric{0};
std::atomic<int64_t> g_optPassRows{0};
std::atomic<int64_t> g_optSymAverage{1};
std::atomic<int64_t> g_optPairTable{3};     // average through symbol words: 0 two gathers, 1 pair table (small codebooks), 2 + predicate-free middle-chunk flush, 3 + one table copy per bank
std::atomic<int64_t> g_optUniformRows{1};
std::atomic<int64_t> g_optDedupKeys{1};     // distinct path, one or two readers: 1 the tuple is the slot's 64-bit key, 0 the generic claim (slots hold query indices)
std::atomic<int64_t> g_optDedupSymbols{1};  // distinct path, average of two trained readers with small codebooks: the temporary holds symbol words, pair_gather_kernel merges on the way out
std::atomic<int64_t> g_optDedupKeep{1};     // copy-out of the distinct path: loads carry an L2 evict_last hint
std::atomic<int64_t> g_optDedup{1};         // union lookups of batches that repeat their words: 0 decode every query, 1 decode distinct id tuples once when a sample says the batch repeats (default), 2 always take the distinct path (tests)
std::atomic<int64_t> g_optTrainedRows{0};   // trained_rows_kernel: 0 never (default: measured slower than the ring kernel), 1 unions of trained readers in ONE launch, 2 single lookups too
std::atomic<int64_t> g_optRowsWarps{0};     // cap on resident warps of trained_rows_kernel (0 = as many as fit)
std::atomic<int64_t> g_optTma{1};           // ring kernel bulk-async staging: 0 off, 1 range decodes (one copy per tile) and long records, 2 one copy per row everywhere
std::atomic<int64_t> g_optRecordAlign{1};   // fixed-stride records: stride a multiple of this many 16 B units (experiment: 2 = whole 32 B sectors)
std::atomic<int64_t> g_optFixedRecords{1};  // 0: keep the offset table even when fixed-stride records would be cheap  // dense uniform lookups: 0 element walk, 1 uniform_rows_kernel unless levels is a power of two, 2 always

struct StatusError {
    memb_b200_status status;
    std::string message;
};

[[noreturn]] void fail(memb_b200_status status, const std::string& message)
{
    throw StatusError{status, message};
}

void cudaCheck(cudaError_t err, const char* what)
{
    if (err != cudaSuccess) {
        fail(MEMB_B200_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(err));
    }
}
#define CUDA_CHECK(expr) cudaCheck((expr), #expr)

// Runs `body`, mapping exceptions to status codes + last_error.
template <class Fn>
memb_b200_status guarded(Fn body)
{
    try {
        body();
        return MEMB_B200_OK;
    } catch (const StatusError& e) {
        g_lastError = e.message;
        return e.status;
    } catch (const fb::FormatError& e) {
        g_lastError = e.what();
        return MEMB_B200_ERR_FORMAT;
    } catch (const std::bad_alloc&) {
        g_lastError = "out of host memory";
        return MEMB_B200_ERR_IO;
    } catch (const std::exception& e) {
        g_lastError = e.what();
        return MEMB_B200_ERR_IO;
    }
}

struct DeviceGuard {
    int previous = -1;
    explicit DeviceGuard(int device)
    {
        CUDA_CHECK(cudaGetDevice(&previous));
        if (previous != device) {
            CUDA_CHECK(cudaSetDevice(device));
        } else {
            previous = -1;
        }
    }
    ~DeviceGuard()
    {
        if (previous >= 0) {
            cudaSetDevice(previous);
        }
    }
};

double msSince(std::chrono::steady_clock::time_point t0)
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

bool deviceIsBlackwell(int device)
{
    int major = 0;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device) != cudaSuccess) {
        return false;
    }
    return major == 10;  // the fatbin holds sm_100a SASS only
}

// Growable device buffer.
struct Scratch {
    void* ptr = nullptr;
    size_t bytes = 0;
    void* ensure(size_t need)
    {
        if (need > bytes) {
            if (ptr) {
                CUDA_CHECK(cudaFree(ptr));
                ptr = nullptr;
                bytes = 0;
            }
            const size_t rounded = (need + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
            CUDA_CHECK(cudaMalloc(&ptr, rounded));
            bytes = rounded;
        }
        return ptr;
    }
    void release()
    {
        if (ptr) {
            cudaFree(ptr);
        }
        ptr = nullptr;
        bytes = 0;
    }
};

template <class T>
T* deviceUpload(const T* host, size_t count, uint64_t& accounted)
{
    T* d = nullptr;
    const size_t bytes = std::max<size_t>(count * sizeof(T), 16);
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&d), bytes));
    if (count) {
        CUDA_CHECK(cudaMemcpy(d, host, count * sizeof(T), cudaMemcpyHostToDevice));
    }
    accounted += bytes;
    return d;
}

// ---- words -> rows without any allocation (memb_b200_lookup_words_noalloc) ----
// The resolved row ids of one call.  In HOST code this name is only a stand-in for the symbol, never a
// device address: the address is looked up once per device (with that device current) below.
static __device__ __align__(256) int32_t g_noallocIds[MEMB_B200_NOALLOC_MAX_WORDS];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noallocMutex;
int32_t* g_noallocIdsOnDevice[NOALLOC_MAX_DEVICES] = {};

// Device address of g_noallocIds on `device`, which must be the CURRENT device.  Called at set-up
// (memb_b200_open) and, should that have failed, again by the entry point.  Throws on a CUDA error.
int32_t* noallocIdsFor(int device)
{
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) {
        fail(MEMB_B200_ERR_ARGUMENT, "lookup_words_noalloc: device ordinal out of range");
    }
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocIdsOnDevice[device]) {
        void* ptr = nullptr;
        CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_noallocIds));
        g_noallocIdsOnDevice[device] = static_cast<int32_t*>(ptr);
    }
    return g_noallocIdsOnDevice[device];
}

}  // namespace

struct memb_b200_reader {
    int device = 0;
    int numSMs = 148;
    size_t smemOptin = 0;
    uint32_t smemBase = 0;  // first dynamic shared byte of a kernel (shared-window address)
    std::unique_ptr<MembFile> file;
    memb_b200_load_stats stats{};
    cudaStream_t stream = nullptr;
    // host outputs: the D2H copy of pass k runs on copyStream while pass k + 1 is decoded on `stream`
    cudaStream_t copyStream = nullptr;
    cudaEvent_t passDecoded[2] = {nullptr, nullptr}, passCopied[2] = {nullptr, nullptr};
    std::mutex mutex;  // serialises use of the scratch buffers

    // vocabulary
    char* dWords = nullptr;
    uint32_t* dWordOffsets = nullptr;
    unsigned long long* dSlots = nullptr;
    uint32_t slotMask = 0;

    // storage-specific
    K::TrainedParams trained{};
    // trained_rows_kernel: LUT images of two widths ([0]: one reader, [1]: unions, half the shared memory)
    struct RowsFlavour {
        uint32_t lutBits = 0;
        uint32_t lut2Entries = 0;
        uint32_t* dImage = nullptr;
        bool escape = false;
    };
    RowsFlavour rowsFlavour[2];
    bool trainedTma = false;    // fixed-stride records: the ring kernel may stage rows with cp.async.bulk
    bool rowsOk = false;        // the table can go through trained_rows_kernel
    uint32_t rowsFieldShift = 0, rowsZeroWord = 0;
    uint32_t lut2Entries = 0;   // second-level entries of the ring kernel's LUT
    uint32_t trainedWarps = 0;  // 0 => generic kernel
    uint32_t trainedG = 4;
    bool trainedEscape = false;
    K::UniformParams uniform{};
    K::FullParams full{};
    std::vector<void*> owned;

    Scratch scratchIds, scratchOut, scratchQuery, scratchOffsets;
    // dynamic tile scheduling of the trained kernel: one counter per launch in flight
    // (a ring: a slot is reused after TILE_COUNTERS launches -- far more than can be in flight on a device)
    static constexpr uint32_t TILE_COUNTERS = 4096;
    uint32_t* dTileCounters = nullptr;
    std::atomic<uint32_t> launchSeq{0};

    ~memb_b200_reader()
    {
        for (void* p : owned) {
            cudaFree(p);
        }
        if (dWords) cudaFree(dWords);
        if (dWordOffsets) cudaFree(dWordOffsets);
        if (dSlots) cudaFree(dSlots);
        if (dTileCounters) cudaFree(dTileCounters);
        scratchIds.release();
        scratchOut.release();
        scratchQuery.release();
        scratchOffsets.release();
        for (int k = 0; k < 2; ++k) {
            if (passDecoded[k]) cudaEventDestroy(passDecoded[k]);
            if (passCopied[k]) cudaEventDestroy(passCopied[k]);
        }
        if (copyStream) cudaStreamDestroy(copyStream);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

uint32_t reverseBits(uint32_t v, uint32_t width)
{
    uint32_t r = 0;
    for (uint32_t b = 0; b < width; ++b) {
        r |= ((v >> b) & 1u) << (width - 1 - b);
    }
    return r;
}

// MSB-first direct table over L bits (the table of src/huffman_table_decoder.h:44-56):
// direct[v] = index into code.keys of the code that prefixes the L-bit value v, or -1 when
// that code is longer than L bits.
std::vector<int> directTable(const PrefixCodeTable& code, uint32_t L)
{
    std::vector<int> direct(size_t(1) << L, -1);
    for (size_t i = 0; i < code.keys.size(); ++i) {
        const uint32_t len = code.lengths[i];
        if (len <= L) {
            const uint32_t first = static_cast<uint32_t>(code.codes[i]) << (L - len);
            for (uint32_t k = 0; k < (1u << (L - len)); ++k) {
                direct[first + k] = static_cast<int>(i);
            }
        }
    }
    return direct;
}

// Second level for codes longer than L bits, the scheme of src/huffman_table_decoder.h:58-93:
// one sub-table per distinct L-bit prefix of such codes, indexed by the NEXT `extra` stream
// bits (first bit in bit 0, like the decoder's window), extra = longest code under the
// prefix - L.  entryOf(i) is the table entry of code i.  Returns, per MSB-first prefix, the
// escape description (base << 4 | extra), 0 for prefixes that need none.
template <class EntryOf>
std::vector<uint32_t> buildSecondLevel(const PrefixCodeTable& code, uint32_t L, EntryOf entryOf,
                                       std::vector<uint32_t>& lut2)
{
    std::vector<uint32_t> extraOf(size_t(1) << L, 0), escape(size_t(1) << L, 0);
    for (size_t i = 0; i < code.keys.size(); ++i) {
        const uint32_t len = code.lengths[i];
        if (len > L) {
            const uint32_t prefix = code.codes[i] >> (len - L);
            extraOf[prefix] = std::max(extraOf[prefix], len - L);
        }
    }
    lut2.clear();
    for (uint32_t prefix = 0; prefix < extraOf.size(); ++prefix) {
        if (extraOf[prefix]) {
            escape[prefix] = (static_cast<uint32_t>(lut2.size()) << 4) | extraOf[prefix];
            lut2.resize(lut2.size() + (size_t(1) << extraOf[prefix]), 0);
        }
    }
    for (size_t i = 0; i < code.keys.size(); ++i) {
        const uint32_t len = code.lengths[i];
        if (len > L) {
            const uint32_t prefix = code.codes[i] >> (len - L);
            const uint32_t extra = escape[prefix] & 15u, base = escape[prefix] >> 4;
            const uint32_t rest = len - L;
            const uint32_t tail = code.codes[i] & ((1u << rest) - 1u);
            for (uint32_t k = 0; k < (1u << (extra - rest)); ++k) {
                lut2[base + reverseBits((tail << (extra - rest)) | k, extra)] = entryOf(i);
            }
        }
    }
    return escape;
}

// LUT of the ring kernel (trained_decode_kernel) over the next `lutBits` bits of the stream.
// The decode kernels hold their window LSB-first, so entry i is the code whose first bit is
// bit 0 of i: the canonical (MSB-first) prefix with its bits reversed.
// entry = (key << symShift) << 16 | length; a code longer than the LUT is wide:
// 0x8000 | extra | base << 16 -> lut2[base + next `extra` bits] holds the entry.
std::vector<uint32_t> buildLut(const PrefixCodeTable& code, uint32_t lutBits, uint32_t symShift,
                               std::vector<uint32_t>& lut2)
{
    const auto entryOf = [&](size_t i) {
        return ((static_cast<uint32_t>(code.keys[i]) << symShift) << 16) | code.lengths[i];
    };
    const std::vector<int> direct = directTable(code, lutBits);
    const std::vector<uint32_t> escape = buildSecondLevel(code, lutBits, entryOf, lut2);
    std::vector<uint32_t> lut(size_t(1) << lutBits, 0);
    for (uint32_t v = 0; v < lut.size(); ++v) {
        lut[reverseBits(v, lutBits)] =
            direct[v] >= 0 ? entryOf(direct[v]) : (0x8000u | (escape[v] & 15u) | ((escape[v] >> 4) << 16));
    }
    return lut;
}

// Multi-symbol LUT of trained_rows_kernel (entry format: trained_rows.cuh): entry i decodes,
// greedily, up to three whole codes out of the L stream bits i (first bit = bit 0).
std::vector<uint32_t> buildRowsLut(const PrefixCodeTable& code, uint32_t L, uint32_t symShift,
                                   std::vector<uint32_t>& lut2)
{
    const std::vector<int> direct = directTable(code, L);
    const std::vector<uint32_t> escape = buildSecondLevel(
        code, L,
        [&](size_t i) { return ((static_cast<uint32_t>(code.keys[i]) << symShift) << 8) | (1u << 5) | code.lengths[i]; }, lut2);
    std::vector<uint32_t> lut(size_t(1) << L, 0);
    const uint32_t mask = (1u << L) - 1u;
    for (uint32_t v = 0; v <= mask; ++v) {
        uint32_t pos = 0, count = 0, syms = 0;
        while (count < 3) {
            const int k = direct[(v << pos) & mask];  // the remaining bits, left-aligned, zero-filled
            if (k < 0 || pos + code.lengths[k] > L) {
                break;
            }
            syms |= (static_cast<uint32_t>(code.keys[k]) << symShift) << (8 * count);
            pos += code.lengths[k];
            ++count;
        }
        lut[reverseBits(v, L)] = count ? (syms << 8) | (count << 5) | pos : ((escape[v] >> 4) << 12) | ((escape[v] & 15u) << 8);
    }
    return lut;
}

K::CanonicalCode buildCanonical(const PrefixCodeTable& code)
{
    K::CanonicalCode c;
    std::memset(&c, 0, sizeof(c));
    c.minLen = code.minLen;
    c.maxLen = code.maxLen;
    for (size_t i = 0; i < code.keys.size(); ++i) {
        const uint32_t len = code.lengths[i];
        if (c.count[len] == 0) {
            c.firstCode[len] = code.codes[i];
            c.firstIndex[len] = static_cast<uint16_t>(i);
        }
        c.count[len] += 1;
        c.keys[i] = code.keys[i];
    }
    return c;
}

void buildVocabulary(memb_b200_reader& r)
{
    const MembFile& f = *r.file;
    r.dWords = deviceUpload(f.words, f.wordsBytes, r.stats.device_bytes);
    r.dWordOffsets = deviceUpload(f.wordOffsets.data(), f.wordOffsets.size(), r.stats.device_bytes);
    uint64_t slots = 16;
    while (slots < 2 * f.rows) {
        slots <<= 1;
    }
    r.slotMask = static_cast<uint32_t>(slots - 1);
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&r.dSlots), slots * 8));
    r.stats.device_bytes += slots * 8;
    CUDA_CHECK(cudaMemsetAsync(r.dSlots, 0xff, slots * 8, r.stream));
    if (f.rows) {
        const uint32_t rows = static_cast<uint32_t>(f.rows);
        K::hash_build_kernel<<<(rows + 255) / 256, 256, 0, r.stream>>>(r.dWords, r.dWordOffsets, rows, r.dSlots,
                                                                     r.slotMask);
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
    }
}

// First dynamic shared byte of a kernel without static shared memory on this device.
uint32_t probeSmemBase(cudaStream_t stream)
{
    uint32_t* dBase = nullptr;
    uint32_t base = 0;
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&dBase), sizeof(uint32_t)));
    K::smem_base_probe_kernel<<<1, 1, 1024, stream>>>(dBase);
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
    CUDA_CHECK(cudaMemcpyAsync(&base, dBase, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
    CUDA_CHECK(cudaStreamSynchronize(stream));
    CUDA_CHECK(cudaFree(dBase));
    return base;
}

void loadTrained(memb_b200_reader& r)
{
    MembFile& f = *r.file;
    const uint32_t rows = static_cast<uint32_t>(f.rows);
    const auto t0 = std::chrono::steady_clock::now();

    uint64_t tempBytes = 0;
    uint8_t* dPacked = deviceUpload(f.packedValues, f.packedSize, tempBytes);
    uint32_t* dValueOffsets = deviceUpload(f.valueOffsets.data(), f.valueOffsets.size(), tempBytes);
    const K::CanonicalCode canonical = buildCanonical(f.code);
    K::CanonicalCode* dCode = deviceUpload(&canonical, 1, r.stats.device_bytes);
    r.owned.push_back(dCode);
    float* dCodebook = deviceUpload(f.centroids.data(), f.centroids.size(), r.stats.device_bytes);
    r.owned.push_back(dCodebook);
    r.stats.upload_ms += msSince(t0);

    const auto t1 = std::chrono::steady_clock::now();
    // decode tables first: the LUT flavour fixes how the rows are re-packed.  Codes of up to 9 bits:
    // one LUT copy per shared-memory bank (conflict-free).  Longer codes: ONE copy over up to 12
    // bits -- measured r02 on 2 M x 300 tables, 1 M random ids: bank conflicts on the single copy
    // cost far less than a second-level lookup in the per-symbol chain (5 bit, maxLen 10: 0.326
    // vs 0.518 ms; 8 bit, maxLen 12: 0.426 vs 0.586 ms).  Codes longer than the LUT (13..16 bits,
    // no Builder emits them) go through a second-level table (src/huffman_table_decoder.h:58-93).
    const uint32_t maxLen = f.code.maxLen;
    const int64_t replicasOpt = g_optLutReplicas.load();
    uint32_t replicas, lutBits, padBits;
    if (replicasOpt == 32 || (replicasOpt == 0 && maxLen <= 9)) {
        replicas = 32;  // entry i of lane l at i*128 + l*4
        lutBits = std::max(1u, std::min(maxLen, 9u));
        padBits = K::ROWS_PAD_BITS;
    } else {
        replicas = 1;
        lutBits = std::max(1u, std::min(maxLen, 12u));
        padBits = 2;
    }
    const bool escape = maxLen > lutBits;
    const uint32_t codebookSize = static_cast<uint32_t>(f.centroids.size());
    // <= 64 centroids: the symbol field carries sym * 4, the byte offset into the codebook
    const uint32_t symShift = (!escape && codebookSize <= 64) ? 2u : 0u;

    // exact row lengths -> 16 B-aligned re-pack
    uint32_t* dRowBytes = nullptr;
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&dRowBytes), std::max<size_t>(size_t(rows) * 4, 16)));
    if (rows) {
        K::measure_rows_kernel<<<(rows + 127) / 128, 128, 0, r.stream>>>(dPacked, f.packedSize, dValueOffsets, rows,
                                                                       f.dim, dCode, dRowBytes);
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
    }
    std::vector<uint32_t> rowBytes(rows);
    CUDA_CHECK(cudaMemcpyAsync(rowBytes.data(), dRowBytes, size_t(rows) * 4, cudaMemcpyDeviceToHost, r.stream));
    CUDA_CHECK(cudaStreamSynchronize(r.stream));
    // padBits (< 8) extra bits in front of the row: one more byte
    const auto unitsOf = [](uint32_t bytes) {
        return bytes ? (bytes + 1 + K::TRAINED_UNIT_BYTES - 1) / K::TRAINED_UNIT_BYTES : 0u;
    };
    uint64_t sumUnits = 0, payload = 0;
    uint32_t maxRowUnits = 0;
    for (uint32_t i = 0; i < rows; ++i) {
        sumUnits += unitsOf(rowBytes[i]);
        payload += rowBytes[i];
        maxRowUnits = std::max(maxRowUnits, unitsOf(rowBytes[i]));
    }
    r.stats.payload_bytes = payload;
    // Fixed-stride RECORDS when the longest row is close to the average one (rows of a
    // k-means + Huffman table are: 131 +- 2 B at 4 bits): row r at r * recUnits, no offset
    // table -- a lookup is id -> record instead of id -> offset -> row, and the 4 B gather
    // that cost a 64 B DRAM access per row is gone.  Otherwise rows back to back + offsets.
    const uint32_t recAlign = static_cast<uint32_t>(g_optRecordAlign.load());
    const uint32_t recUnits = (std::max(maxRowUnits, 1u) + recAlign - 1) / recAlign * recAlign;
    const bool fixed = g_optFixedRecords.load() && uint64_t(rows) * recUnits <= sumUnits + sumUnits / 4 + 4096;
    std::vector<uint32_t> rowOff16;
    uint64_t units = fixed ? uint64_t(std::max(rows, 1u)) * recUnits : 0;
    if (!fixed) {
        rowOff16.resize(size_t(rows) + 1);
        for (uint32_t i = 0; i < rows; ++i) {
            rowOff16[i] = static_cast<uint32_t>(units);
            units += unitsOf(rowBytes[i]);
            if (units > 0xffffffffULL) {
                fail(MEMB_B200_ERR_FORMAT, "memb file: re-packed rows exceed 64 GiB");
            }
        }
        rowOff16[rows] = static_cast<uint32_t>(units);
    } else if (units > 0xffffffffULL) {
        fail(MEMB_B200_ERR_FORMAT, "memb file: re-packed rows exceed 64 GiB");
    }
    uint32_t* dRowOff16 = nullptr;
    if (!fixed) {
        dRowOff16 = deviceUpload(rowOff16.data(), rowOff16.size(), r.stats.device_bytes);
        r.owned.push_back(dRowOff16);
    }
    K::RowUnit* dBlob = nullptr;
    // one spare unit so the decoder's 96-bit window never leaves the allocation
    const size_t blobBytes = (units + 1) * K::TRAINED_UNIT_BYTES;
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&dBlob), blobBytes));
    r.stats.device_bytes += blobBytes;
    r.owned.push_back(dBlob);
    if (rows == 0) {
        // a table without rows still has record 0: lanes without a row decode it
        CUDA_CHECK(cudaMemsetAsync(dBlob, 0, blobBytes, r.stream));
    }
    CUDA_CHECK(cudaMemsetAsync(reinterpret_cast<unsigned char*>(dBlob) + units * K::TRAINED_UNIT_BYTES, 0,
                               K::TRAINED_UNIT_BYTES, r.stream));
    if (rows) {
        const uint64_t threads = uint64_t(rows) * 32;
        K::repack_rows_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, r.stream>>>(
            dPacked, f.packedSize, dValueOffsets, dRowBytes, dRowOff16, fixed ? recUnits : 0u, rows, padBits,
            reinterpret_cast<uint32_t*>(dBlob));
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
    }

    // the same records interleaved per 32 rows: a range decode stages a tile with one bulk copy
    K::RowUnit* dBlobTiled = nullptr;
    if (fixed && rows && g_optTma.load()) {
        const uint64_t groups = (uint64_t(rows) + 31) / 32;
        const uint64_t tiledWords = groups * 32 * recUnits * K::TRAINED_UNIT_WORDS;
        CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&dBlobTiled), tiledWords * 4));
        r.stats.device_bytes += tiledWords * 4;
        r.owned.push_back(dBlobTiled);
        K::interleave_records_kernel<<<static_cast<unsigned>(std::min<uint64_t>((tiledWords + 255) / 256, uint64_t(r.numSMs) * 32)),
                                       256, 0, r.stream>>>(reinterpret_cast<const uint32_t*>(dBlob), rows,
                                                           recUnits * K::TRAINED_UNIT_WORDS,
                                                           reinterpret_cast<uint32_t*>(dBlobTiled), tiledWords);
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
    }

    std::vector<uint32_t> lut2;
    const std::vector<uint32_t> lut = buildLut(f.code, lutBits, symShift, lut2);
    if (lut2.size() > K::ROWS_LUT2_MAX_ENTRIES) {
        fail(MEMB_B200_ERR_UNSUPPORTED, "memb file: prefix code needs more second-level table entries than the kernels hold");
    }
    uint32_t* dLut = deviceUpload(lut.data(), lut.size(), r.stats.device_bytes);
    r.owned.push_back(dLut);
    uint32_t* dLut2 = deviceUpload(lut2.data(), lut2.size(), r.stats.device_bytes);
    r.owned.push_back(dLut2);

    K::TrainedParams& p = r.trained;
    p.blob = dBlob;
    p.blobTiled = dBlobTiled;
    p.rowOff16 = dRowOff16;
    p.recUnits = fixed ? recUnits : 0u;
    p.lut = dLut;
    p.lut2 = dLut2;
    p.lut2Entries = static_cast<uint32_t>(lut2.size());
    p.codebook = dCodebook;
    p.code = dCode;
    p.rows = rows;
    p.dim = f.dim;
    p.lutBits = lutBits;
    p.lutReplicas = replicas;
    p.padBits = padBits;
    p.lutMask = ((1u << lutBits) - 1) << padBits;
    p.symShift = symShift;
    p.codebookSize = codebookSize;
    p.rowWords = (fixed ? recUnits : maxRowUnits) * K::TRAINED_UNIT_WORDS + 4;  // every staged unit + the window's slack
    // slots of the 256-entry shared codebook past the last centroid hold +0.0: the row of a
    // missing word is "decoded" to that slot (symbol fields reach 63 with symShift 2, else 255)
    p.hasZeroSlot = codebookSize < (symShift == 2 ? 64u : 256u) ? 1u : 0u;
    p.zeroWord = p.hasZeroSlot ? ((codebookSize << symShift) & 0xffu) * 0x01010101u : 0u;
    r.trainedEscape = escape;
    // symbols per window reload: G * maxLen <= 64 - padBits valid bits
    const uint32_t allowed = (64 - padBits) / std::max(maxLen, 1u);
    // (re-centring the window inside the long-code branch would allow G = 4 there; measured r02 on the
    // 16-bit-code table: 0.501 ms against 0.474 ms with plain G = 3 -- the branch body is inlined 32 times)
    r.trainedG = escape ? 3 : allowed >= 16 ? 16 : allowed >= 8 ? 8 : allowed >= 7 ? 7 : allowed >= 6 ? 6 : 5;

    // ---- trained_rows_kernel: multi-symbol LUT images, 9 bits wide for one reader, 8 for unions ----
    r.rowsOk = false;
    if (fixed && padBits == K::ROWS_PAD_BITS && codebookSize <= 255 && f.dim % 4 == 0 && f.dim >= 4 &&
        f.dim <= K::TRAINED_MAX_DIM) {
        r.rowsFieldShift = codebookSize <= 63 ? 0u : 2u;  // <= 63 centroids: the field already is sym * 4
        const uint32_t rowsSymShift = 2u - r.rowsFieldShift;
        r.rowsZeroWord = ((codebookSize << rowsSymShift) & 0xffu) * 0x01010101u;
        r.rowsOk = true;
        for (int flavour = 0; flavour < 2 && r.rowsOk; ++flavour) {
            const uint32_t L = std::max(1u, std::min(maxLen * 3u, flavour == 0 ? 9u : 8u));
            std::vector<uint32_t> rowsLut2;
            const std::vector<uint32_t> rowsLut = buildRowsLut(f.code, L, rowsSymShift, rowsLut2);
            if (rowsLut2.size() > K::ROWS_LUT2_MAX_ENTRIES) {
                r.rowsOk = false;
                break;
            }
            rowsLut2.resize((rowsLut2.size() + 3) / 4 * 4, 0);
            // image: LUT with 32 bank replicas | second level | 256 codebook floats (zero padded)
            std::vector<uint32_t> image((size_t(32) << L) + rowsLut2.size() + 256, 0);
            for (size_t i = 0; i < rowsLut.size(); ++i) {
                std::fill_n(image.begin() + i * 32, 32, rowsLut[i]);
            }
            std::copy(rowsLut2.begin(), rowsLut2.end(), image.begin() + (size_t(32) << L));
            std::memcpy(image.data() + (size_t(32) << L) + rowsLut2.size(), f.centroids.data(), codebookSize * 4);
            memb_b200_reader::RowsFlavour& fl = r.rowsFlavour[flavour];
            fl.lutBits = L;
            fl.lut2Entries = static_cast<uint32_t>(rowsLut2.size());
            fl.escape = maxLen > L;
            fl.dImage = deviceUpload(image.data(), image.size(), r.stats.device_bytes);
            r.owned.push_back(fl.dImage);
        }
    }

    // how many warps (tiles in flight) fit beside the tables in the shared window
    r.smemBase = probeSmemBase(r.stream);
    const uint32_t lutBytes = K::trainedLutBytes(lutBits, replicas);
    r.trainedTma = fixed && g_optTma.load() != 0;  // (decided at load: the landing buffers cost shared memory)
    const uint32_t perWarp = K::trainedWarpBytes(p.rowWords, r.trainedTma ? K::trainedLandingBytes(p.recUnits) : 0u);
    uint32_t warps = 0;
    for (uint32_t w = 1; w <= K::TRAINED_MAX_WARPS; ++w) {
        const K::TrainedLayout lay = K::trainedLayout(r.smemBase, lutBytes, perWarp, w, p.lut2Entries);
        if (lay.end - r.smemBase > r.smemOptin) {
            break;
        }
        warps = w;
    }
    const int64_t warpsOpt = g_optWarps.load();
    if (warpsOpt > 0) {
        warps = std::min<uint32_t>(warps, static_cast<uint32_t>(warpsOpt));
    }
    if (f.dim > K::TRAINED_MAX_DIM || p.codebookSize > 256 || g_optGeneric.load()) {
        warps = 0;
    }
    r.trainedWarps = warps;

    r.stats.max_code_len = maxLen;
    r.stats.lut_bits = lutBits;
    r.stats.lut_replicas = replicas;
    r.stats.codebook_size = p.codebookSize;

    CUDA_CHECK(cudaStreamSynchronize(r.stream));
    CUDA_CHECK(cudaFree(dPacked));
    CUDA_CHECK(cudaFree(dValueOffsets));
    CUDA_CHECK(cudaFree(dRowBytes));
    r.stats.prepare_ms += msSince(t1);
    std::vector<uint32_t>().swap(f.valueOffsets);
}

void loadUniform(memb_b200_reader& r)
{
    MembFile& f = *r.file;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<float2> minmax(f.rows);
    for (uint64_t i = 0; i < f.rows; ++i) {
        minmax[i] = make_float2(f.uniformMin[i], f.uniformMax[i]);
        r.stats.payload_bytes += f.dim + 8;
    }
    // records {min, max, dim bytes, padding} (kernels.cuh: UniformParams), interleaved on the device
    size_t staged = 0;
    uint8_t* dValues = deviceUpload(f.uniformValues.data(), f.uniformValues.size(), staged);
    float2* dMinMax = deviceUpload(minmax.data(), minmax.size(), staged);
    const uint32_t recordBytes = f.dim + 8;
    const uint32_t align = recordBytes >= 256 ? 64 : 16;
    const uint32_t stride = (recordBytes + align - 1) / align * align;
    const uint64_t recordsBytes = static_cast<uint64_t>(f.rows) * stride;
    if (recordsBytes / 4 + 2 > 0xffffffffull) {
        fail(MEMB_B200_ERR_FORMAT, "uniform table too large for 32-bit record word offsets");
    }
    uint8_t* dRecords = nullptr;
    CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&dRecords), std::max<uint64_t>(recordsBytes, 16)));
    r.stats.device_bytes += std::max<uint64_t>(recordsBytes, 16);
    r.owned.push_back(dRecords);
    if (f.rows) {
        const unsigned grid = static_cast<unsigned>(std::min<uint64_t>((recordsBytes + 255) / 256, uint64_t(r.numSMs) * 64));
        CUDA_CHECK(cudaDeviceSynchronize());  // the pageable uploads above have landed
        ++g_launches;
        K::interleave_uniform_kernel<<<grid, 256, 0, r.stream>>>(dValues, dMinMax, static_cast<uint32_t>(f.rows), f.dim,
                                                                 stride, dRecords);
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaStreamSynchronize(r.stream));
    }
    CUDA_CHECK(cudaFree(dValues));
    CUDA_CHECK(cudaFree(dMinMax));
    r.uniform.records = dRecords;
    r.uniform.stride = stride;
    r.uniform.rows = static_cast<uint32_t>(f.rows);
    r.uniform.dim = f.dim;
    // quantization_levels is a uint8 promoted to float in the reference's
    // `... * floatValue / quantizationLevels` (src/uniform_compression.cpp:68-72)
    r.uniform.levels = static_cast<float>(f.quantizationLevels);
    const uint32_t q = f.quantizationLevels;
    // how `/ levels` is evaluated on the device (kernels.cuh: divideByLevels); every mode is
    // the correctly rounded quotient
    r.uniform.divMode = (q != 0 && (q & (q - 1)) == 0) ? 1 : (q == 255 ? 2 : 0);
    r.uniform.invLevels = q ? 1.0f / static_cast<float>(q) : 0.0f;
    r.stats.upload_ms += msSince(t0);
    std::vector<uint8_t>().swap(f.uniformValues);
    std::vector<float>().swap(f.uniformMin);
    std::vector<float>().swap(f.uniformMax);
}

void loadFull(memb_b200_reader& r)
{
    MembFile& f = *r.file;
    const auto t0 = std::chrono::steady_clock::now();
    float* dValues = deviceUpload(f.fullValues.data(), f.fullValues.size(), r.stats.device_bytes);
    r.owned.push_back(dValues);
    r.full.values = dValues;
    r.full.rows = static_cast<uint32_t>(f.rows);
    r.full.dim = f.dim;
    r.stats.payload_bytes = uint64_t(f.rows) * f.dim * 4;
    r.stats.upload_ms += msSince(t0);
    std::vector<float>().swap(f.fullValues);
}

}  // namespace

// ------------------------------------------------------------- launchers ----
namespace {

bool vec4Ok(uint32_t dim, const K::OutputSpec& o)
{
    return dim % 4 == 0 && o.rowStride % 4 == 0 && o.colOffset % 4 == 0 &&
           reinterpret_cast<uintptr_t>(o.out) % 16 == 0;
}

template <int G, bool VEC4, bool ESCAPE, bool SMALLCB, bool RMW, bool SIDE = false, int SYM = 0, int TMA = 0>
void launchTrainedKernel(const memb_b200_reader& r, const int32_t* ids, uint32_t rowBegin, uint64_t n,
                         const K::OutputSpec& o, int epi, cudaStream_t stream)
{
    const K::TrainedParams& p = r.trained;
    const K::TrainedLayout lay = K::trainedLayout(
        r.smemBase, K::trainedLutBytes(p.lutBits, p.lutReplicas),
        K::trainedWarpBytes(p.rowWords, TMA ? K::trainedLandingBytes(p.recUnits) : 0u), r.trainedWarps,
        SYM >= 2 ? o.pairEntries * o.pairReplicas : p.lut2Entries);
    const size_t smemBytes = lay.end - r.smemBase;
    const uint64_t tiles = (n + 31) / 32;
    const unsigned grid = static_cast<unsigned>(
        std::min<uint64_t>(static_cast<uint64_t>(r.numSMs), (tiles + r.trainedWarps - 1) / r.trainedWarps));
    const unsigned block = r.trainedWarps * 32;
    if (tiles + 4ull * grid * r.trainedWarps >= 0xffffffffull) {
        fail(MEMB_B200_ERR_ARGUMENT, "decode: too many rows in one call");
    }
    auto kernel = K::trained_decode_kernel<G, VEC4, ESCAPE, SMALLCB, RMW, SIDE, SYM, TMA>;
    {
        // once per instantiation and device: the opt-in maximum (a per-launch size could be undercut
        // by another thread's launch of the same instantiation for another reader)
        static std::atomic<uint64_t> configured{0};
        const uint64_t bit = 1ull << (r.device & 63);
        if (!(configured.load() & bit)) {
            CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(r.smemOptin)));
            configured |= bit;
        }
    }
    // a {ticket, done} pair of the ring; the kernel leaves both words zero (no memset per launch)
    // (lower half of the ring: pairs; upper half: the single-word, memset-per-launch slots of the other kernels)
    uint32_t* counter = r.dTileCounters +
                        2 * (const_cast<memb_b200_reader&>(r).launchSeq++ % (memb_b200_reader::TILE_COUNTERS / 4));
    kernel<<<grid, block, smemBytes, stream>>>(p, ids, rowBegin, n, o, epi, counter);
}

// ---- trained_rows_kernel: shared-window plan and launch ----
struct RowsPlan {
    K::RowsParams P{};
    uint32_t warps = 0;
    size_t smemBytes = 0;
    int G = 6;
    bool escape = false;
    bool smallCodebooks = true;
};

// First-fit allocator over the free fragments of the CTA's shared window.
struct SmemFragments {
    std::vector<std::pair<uint32_t, uint32_t>> free;  // [begin, end)
    uint32_t highWater = 0;
    uint32_t alloc(uint32_t bytes, uint32_t align)
    {
        for (auto& f : free) {
            const uint32_t at = (f.first + align - 1) / align * align;
            if (at + bytes <= f.second) {
                f.first = at + bytes;  // (the alignment gap is lost: blocks only grow in size order)
                highWater = std::max(highWater, at + bytes);
                return at;
            }
        }
        return 0;
    }
};

// Lays R readers out in the shared window: LUT images at addresses aligned to their own size
// (the decode ORs base, bank replica and index together), small tables and the per-warp
// blocks (staged records, symbol tiles) first-fit into the fragments around them.
// Returns false when the union cannot run as one launch (then the ring kernels take it).
bool planRows(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode, RowsPlan& plan)
{
    const int64_t use = g_optTrainedRows.load();
    if (R == 0 || R > 2 || use == 0 || (R == 1 && use < 2)) {
        return false;  // instantiated for one and two readers
    }
    const memb_b200_reader& r0 = *readers[0];
    const int flavour = R == 1 ? 0 : 1;
    K::RowsParams& P = plan.P;
    uint32_t maxLen = 0, colWords = 0, outUnits = 0, tileBytes = 0;
    for (uint32_t k = 0; k < R; ++k) {
        const memb_b200_reader& r = *readers[k];
        if (r.file->storage != STORAGE_TRAINED || !r.rowsOk || r.device != r0.device) {
            return false;
        }
        const auto& fl = r.rowsFlavour[flavour];
        K::RowsReader& rd = P.rd[k];
        rd.blob = r.trained.blob;
        rd.lutImage = fl.dImage;
        rd.recUnits = r.trained.recUnits;
        rd.rows = r.trained.rows;
        rd.dim = r.trained.dim;
        rd.lutBits = fl.lutBits;
        rd.lut2Entries = fl.lut2Entries;
        rd.zeroWord = r.rowsZeroWord;
        rd.fieldShift = r.rowsFieldShift;
        const uint32_t words = rd.dim / 4;
        rd.tileStride = (words | 1u) * 4;  // odd number of words: lanes writing the same word of their rows hit 32 banks
        rd.tileOffset = tileBytes;
        tileBytes += 32 * rd.tileStride;
        rd.colUnit = mode == MEMB_B200_CONCATENATE ? outUnits : 0;
        outUnits = mode == MEMB_B200_CONCATENATE ? outUnits + words : words;
        colWords = std::max(colWords, rd.recUnits * K::TRAINED_UNIT_WORDS + 4);
        plan.escape = plan.escape || fl.escape;
        plan.smallCodebooks = plan.smallCodebooks && rd.fieldShift == 0;
        maxLen = std::max(maxLen, r.file->code.maxLen);
    }
    // lookups per window reload: every lookup consumes (and peeks) at most max(L, maxLen) bits
    plan.G = !plan.escape ? 6 : maxLen <= 13 ? 4 : 3;
    P.colWords = colWords;
    P.outUnits = outUnits;
    P.divisor = static_cast<float>(R);

    const uint32_t base = r0.smemBase, limit = base + static_cast<uint32_t>(r0.smemOptin);
    // LUTs, largest first, back to back from the first address aligned to the largest
    std::vector<uint32_t> order(R);
    for (uint32_t k = 0; k < R; ++k) {
        order[k] = k;
    }
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return P.rd[a].lutBits > P.rd[b].lutBits; });
    const uint32_t firstBytes = 128u << P.rd[order[0]].lutBits;
    uint32_t cursor = (base + firstBytes) / firstBytes * firstBytes;
    const uint32_t lutBegin = cursor;
    for (uint32_t k : order) {
        P.rd[k].sLut = cursor;  // sizes are powers of two in descending order: still aligned
        cursor += 128u << P.rd[k].lutBits;
    }
    if (cursor > limit) {
        return false;
    }
    SmemFragments frag;
    frag.free = {{base, lutBegin}, {cursor, limit}};
    frag.highWater = cursor;
    P.sMbar = frag.alloc(8, 8);
    for (uint32_t k = 0; k < R; ++k) {
        P.rd[k].sCodebook = frag.alloc(1024, 1024);
        P.rd[k].sLut2 = P.rd[k].lut2Entries ? frag.alloc(P.rd[k].lut2Entries * 4, 16) : P.sMbar;
        if (!P.rd[k].sCodebook || !P.rd[k].sLut2) {
            return false;
        }
    }
    const int64_t cap = g_optRowsWarps.load();
    const uint32_t maxWarps = cap > 0 ? std::min<uint32_t>(K::ROWS_MAX_WARPS, static_cast<uint32_t>(cap)) : K::ROWS_MAX_WARPS;
    uint32_t warps = 0;
    // tiles first (the larger blocks), then the staging columns
    for (; warps < maxWarps; ++warps) {
        const SmemFragments saved = frag;
        const uint32_t tile = frag.alloc(tileBytes, 16);
        const uint32_t cols = tile ? frag.alloc(colWords * 128, 128) : 0;
        if (!tile || !cols) {
            frag = saved;
            break;
        }
        P.sTiles[warps] = tile;
        P.sCols[warps] = cols;
    }
    if (warps < 4 || !P.sMbar) {
        return false;
    }
    plan.warps = warps;
    P.smemEnd = frag.highWater;
    plan.smemBytes = frag.highWater - base;
    return true;
}

template <int R, int MODE, int G, bool ESC, bool SMALLCB>
void launchRowsKernel(const memb_b200_reader& r0, const RowsPlan& plan, const int32_t* ids, uint64_t idStride,
                      uint32_t rowBegin, uint64_t n, float* out, cudaStream_t stream)
{
    auto kernel = K::trained_rows_kernel<R, MODE, G, ESC, SMALLCB>;
    // once per instantiation and device: the opt-in maximum (not this launch's size, which
    // another thread's launch of the same instantiation could undercut)
    static std::atomic<uint64_t> configured{0};
    const uint64_t bit = 1ull << (r0.device & 63);
    if (!(configured.load() & bit)) {
        CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(r0.smemOptin)));
        configured |= bit;
    }
    const uint64_t tiles = (n + 31) / 32;
    const unsigned grid = static_cast<unsigned>(
        std::min<uint64_t>(static_cast<uint64_t>(r0.numSMs), (tiles + plan.warps - 1) / plan.warps));
    if (tiles + 4ull * grid * plan.warps >= 0xffffffffull) {
        fail(MEMB_B200_ERR_ARGUMENT, "decode: too many rows in one call");
    }
    uint32_t* counter = r0.dTileCounters + memb_b200_reader::TILE_COUNTERS / 2 +
                        (const_cast<memb_b200_reader&>(r0).launchSeq++ % (memb_b200_reader::TILE_COUNTERS / 2));
    CUDA_CHECK(cudaMemsetAsync(counter, 0, sizeof(uint32_t), stream));
    kernel<<<grid, plan.warps * 32, plan.smemBytes, stream>>>(plan.P, ids, idStride, rowBegin, n, out, counter);
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
}

template <int R, int MODE>
void launchRowsFlavour(const memb_b200_reader& r0, const RowsPlan& plan, const int32_t* ids, uint64_t idStride,
                       uint32_t rowBegin, uint64_t n, float* out, cudaStream_t stream)
{
    const bool s = plan.smallCodebooks;
    if (!plan.escape) {
        s ? launchRowsKernel<R, MODE, 6, false, true>(r0, plan, ids, idStride, rowBegin, n, out, stream)
          : launchRowsKernel<R, MODE, 6, false, false>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    } else if (plan.G == 4) {
        s ? launchRowsKernel<R, MODE, 4, true, true>(r0, plan, ids, idStride, rowBegin, n, out, stream)
          : launchRowsKernel<R, MODE, 4, true, false>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    } else {
        s ? launchRowsKernel<R, MODE, 3, true, true>(r0, plan, ids, idStride, rowBegin, n, out, stream)
          : launchRowsKernel<R, MODE, 3, true, false>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    }
}

// One launch for a dense (n x union dim) result of trained readers; false: not applicable.
bool tryLaunchRows(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode, const int32_t* ids,
                   uint64_t idStride, uint32_t rowBegin, uint64_t n, float* out, cudaStream_t stream)
{
    if (n == 0 || reinterpret_cast<uintptr_t>(out) % 16 != 0) {
        return false;
    }
    RowsPlan plan;
    if (!planRows(readers, R, mode, plan)) {
        return false;
    }
    const memb_b200_reader& r0 = *readers[0];
    if (R == 1) {
        launchRowsFlavour<1, K::ROWS_CONCAT>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    } else if (mode == MEMB_B200_AVERAGE) {
        launchRowsFlavour<2, K::ROWS_AVERAGE>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    } else {
        launchRowsFlavour<2, K::ROWS_CONCAT>(r0, plan, ids, idStride, rowBegin, n, out, stream);
    }
    return true;
}

// Picks the instantiation: G symbols per window reload (loadTrained); for the
// 128-bit-store kernels whether the symbol field already is a codebook offset and
// whether the launch carries an average epilogue.  Odd shapes (scalar stores) always
// run G = 5, which is valid for every non-escaping code (5 * 11 <= 57, 5 * 12 <= 62).
template <int G>
void launchTrainedVec(const memb_b200_reader& r, const int32_t* ids, uint32_t rowBegin, uint64_t n,
                      const K::OutputSpec& o, int epi, cudaStream_t stream)
{
    const bool small = r.trained.symShift == 2;
    const bool rmw = epi != K::EPI_WRITE;
    if (o.sym) {
        // average of trained readers through symbol words (OutputSpec::sym)
        if (rmw && o.pairEntries) {
            small ? launchTrainedKernel<G, true, false, true, true, false, 3>(r, ids, rowBegin, n, o, epi, stream)
                  : launchTrainedKernel<G, true, false, false, true, false, 3>(r, ids, rowBegin, n, o, epi, stream);
        } else if (rmw) {
            small ? launchTrainedKernel<G, true, false, true, true, false, 2>(r, ids, rowBegin, n, o, epi, stream)
                  : launchTrainedKernel<G, true, false, false, true, false, 2>(r, ids, rowBegin, n, o, epi, stream);
        } else {
            small ? launchTrainedKernel<G, true, false, true, false, false, 1>(r, ids, rowBegin, n, o, epi, stream)
                  : launchTrainedKernel<G, true, false, false, false, false, 1>(r, ids, rowBegin, n, o, epi, stream);
        }
        return;
    }
    if (!rmw && (o.sideHead || o.sideTail)) {
        small ? launchTrainedKernel<G, true, false, true, false, true>(r, ids, rowBegin, n, o, epi, stream)
              : launchTrainedKernel<G, true, false, false, false, true>(r, ids, rowBegin, n, o, epi, stream);
        return;
    }
    if (!rmw && r.trainedTma && !ids && rowBegin % 32 == 0 && r.trained.blobTiled && g_optTma.load()) {
        // range decode from a tile boundary: one bulk copy per tile out of the interleaved records
        small ? launchTrainedKernel<G, true, false, true, false, false, 0, 2>(r, ids, rowBegin, n, o, epi, stream)
              : launchTrainedKernel<G, true, false, false, false, false, 0, 2>(r, ids, rowBegin, n, o, epi, stream);
        return;
    }
    if (!rmw && r.trainedTma &&
        (g_optTma.load() == 2 || (g_optTma.load() == 1 && r.trained.recUnits > K::TRAINED_PREFETCH_UNITS))) {
        // one bulk copy per ROW: slower than the register path for records that fit the prefetch
        // registers (4 bit: 0.305 vs 0.296 ms per 1 M random ids), faster for longer ones, whose
        // tail the register path fetches at staging time (8 bit, 272 B records: 0.401 vs 0.440 ms)
        small ? launchTrainedKernel<G, true, false, true, false, false, 0, 1>(r, ids, rowBegin, n, o, epi, stream)
              : launchTrainedKernel<G, true, false, false, false, false, 0, 1>(r, ids, rowBegin, n, o, epi, stream);
        return;
    }
    if (small) {
        rmw ? launchTrainedKernel<G, true, false, true, true>(r, ids, rowBegin, n, o, epi, stream)
            : launchTrainedKernel<G, true, false, true, false>(r, ids, rowBegin, n, o, epi, stream);
    } else {
        rmw ? launchTrainedKernel<G, true, false, false, true>(r, ids, rowBegin, n, o, epi, stream)
            : launchTrainedKernel<G, true, false, false, false>(r, ids, rowBegin, n, o, epi, stream);
    }
}

template <bool VEC4>
void launchTrainedG(const memb_b200_reader& r, const int32_t* ids, uint32_t rowBegin, uint64_t n,
                    const K::OutputSpec& o, int epi, cudaStream_t stream)
{
    if (r.trainedEscape) {
        if (VEC4 && epi == K::EPI_WRITE && (o.sideHead || o.sideTail)) {
            launchTrainedKernel<3, VEC4, true, false, false, VEC4>(r, ids, rowBegin, n, o, epi, stream);
        } else {
            launchTrainedKernel<3, VEC4, true, false, false>(r, ids, rowBegin, n, o, epi, stream);
        }
        return;
    }
    if (!VEC4) {
        launchTrainedKernel<5, false, false, false, false>(r, ids, rowBegin, n, o, epi, stream);
        return;
    }
    switch (r.trainedG) {
    case 16:
        launchTrainedVec<16>(r, ids, rowBegin, n, o, epi, stream);
        break;
    case 8:
        launchTrainedVec<8>(r, ids, rowBegin, n, o, epi, stream);
        break;
    case 7:
        launchTrainedVec<7>(r, ids, rowBegin, n, o, epi, stream);
        break;
    case 6:
        launchTrainedVec<6>(r, ids, rowBegin, n, o, epi, stream);
        break;
    default:
        launchTrainedVec<5>(r, ids, rowBegin, n, o, epi, stream);
        break;
    }
}

// Decodes n rows (ids, or the range starting at rowBegin when ids == nullptr)
// of one reader into `o` with the given epilogue.
void launchDecode(const memb_b200_reader& r, const int32_t* ids, uint32_t rowBegin, uint64_t n,
                  const K::OutputSpec& o, int epi, cudaStream_t stream)
{
    if (n == 0) {
        return;
    }
    const uint32_t dim = r.file->dim;
    const bool vec = vec4Ok(dim, o);
    const uint64_t elems = n * (vec ? dim / 4 : dim);
    const unsigned flatGrid = static_cast<unsigned>(std::min<uint64_t>((elems + 255) / 256, uint64_t(r.numSMs) * 32));
    switch (r.file->storage) {
    case STORAGE_TRAINED:
        if (vec && epi == K::EPI_WRITE && o.rowStride == dim && o.colOffset == 0 && !o.sideHead && !o.sideTail && !o.sym) {
            // dense result of a plain lookup: whole rows, contiguous flush (trained_rows.cuh)
            memb_b200_reader* self = const_cast<memb_b200_reader*>(&r);
            if (tryLaunchRows(&self, 1, MEMB_B200_CONCATENATE, ids, 0, rowBegin, n, o.out, stream)) {
                return;
            }
        }
        if (r.trainedWarps == 0) {
            const unsigned grid = static_cast<unsigned>((n + 127) / 128);
            if (epi == K::EPI_WRITE) {
                K::trained_decode_generic_kernel<K::EPI_WRITE><<<grid, 128, 0, stream>>>(r.trained, ids, rowBegin, n, o);
            } else if (epi == K::EPI_ACCUMULATE) {
                K::trained_decode_generic_kernel<K::EPI_ACCUMULATE><<<grid, 128, 0, stream>>>(r.trained, ids, rowBegin, n, o);
            } else {
                K::trained_decode_generic_kernel<K::EPI_FINALIZE><<<grid, 128, 0, stream>>>(r.trained, ids, rowBegin, n, o);
            }
        } else if (vec) {
            launchTrainedG<true>(r, ids, rowBegin, n, o, epi, stream);
        } else {
            launchTrainedG<false>(r, ids, rowBegin, n, o, epi, stream);
        }
        break;
    case STORAGE_UNIFORM: {
        auto go = [&](auto kernel) { kernel<<<flatGrid, 256, 0, stream>>>(r.uniform, ids, rowBegin, n, o); };
        if (vec && epi == K::EPI_WRITE && (o.sideHead || o.sideTail)) {
            go(K::uniform_decode_kernel<true, K::EPI_WRITE, true>);
        } else if (vec && epi == K::EPI_WRITE && o.rowStride == dim && o.colOffset == 0 && dim >= 8 && dim <= 8188 &&
                   (g_optUniformRows.load() == 2 || (g_optUniformRows.load() == 1 && r.uniform.divMode != 1))) {
            // (power-of-two levels keep the element walk: its dequant is 3 instructions per value and on
            // the record layout it is the faster one there, 0.265 against 0.278 ms per 1 M random ids)
            // dense result: per-row work once per row, a group ahead (uniform_rows_kernel)
            auto rows = [&](auto kernel) {
                static int perSM = 0;  // same for every device of this process (one architecture)
                if (perSM == 0) {
                    int b = 0;
                    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kernel, K::UNIFORM_ROWS_WARPS * 32, 0);
                    perSM = b > 0 ? b : 4;
                }
                constexpr unsigned W = K::UNIFORM_ROWS_WARPS;  // one 32-row group per warp at a time
                const unsigned grid = static_cast<unsigned>(std::min<uint64_t>(((n + 31) / 32 + W - 1) / W, uint64_t(r.numSMs) * perSM));
                uint32_t* counter = r.dTileCounters + memb_b200_reader::TILE_COUNTERS / 2 +
                                    (const_cast<memb_b200_reader&>(r).launchSeq++ % (memb_b200_reader::TILE_COUNTERS / 2));
                CUDA_CHECK(cudaMemsetAsync(counter, 0, sizeof(uint32_t), stream));
                kernel<<<grid, W * 32, 0, stream>>>(r.uniform, ids, rowBegin, n, o.out, counter);
            };
            r.uniform.divMode == 1   ? rows(K::uniform_rows_kernel<1>)
            : r.uniform.divMode == 2 ? rows(K::uniform_rows_kernel<2>)
                                     : rows(K::uniform_rows_kernel<0>);
        } else if (vec && epi == K::EPI_WRITE) {
            r.uniform.divMode == 1   ? go(K::uniform_decode_kernel<true, K::EPI_WRITE, false, 1>)
            : r.uniform.divMode == 2 ? go(K::uniform_decode_kernel<true, K::EPI_WRITE, false, 2>)
                                     : go(K::uniform_decode_kernel<true, K::EPI_WRITE, false, 0>);
        } else if (vec) {
            epi == K::EPI_WRITE        ? go(K::uniform_decode_kernel<true, K::EPI_WRITE>)
            : epi == K::EPI_ACCUMULATE ? go(K::uniform_decode_kernel<true, K::EPI_ACCUMULATE>)
                                       : go(K::uniform_decode_kernel<true, K::EPI_FINALIZE>);
        } else {
            epi == K::EPI_WRITE        ? go(K::uniform_decode_kernel<false, K::EPI_WRITE>)
            : epi == K::EPI_ACCUMULATE ? go(K::uniform_decode_kernel<false, K::EPI_ACCUMULATE>)
                                       : go(K::uniform_decode_kernel<false, K::EPI_FINALIZE>);
        }
        break;
    }
    default: {
        auto go = [&](auto kernel) { kernel<<<flatGrid, 256, 0, stream>>>(r.full, ids, rowBegin, n, o); };
        if (vec && epi == K::EPI_WRITE && (o.sideHead || o.sideTail)) {
            go(K::full_gather_kernel<true, K::EPI_WRITE, true>);
        } else if (vec) {
            epi == K::EPI_WRITE        ? go(K::full_gather_kernel<true, K::EPI_WRITE>)
            : epi == K::EPI_ACCUMULATE ? go(K::full_gather_kernel<true, K::EPI_ACCUMULATE>)
                                       : go(K::full_gather_kernel<true, K::EPI_FINALIZE>);
        } else {
            epi == K::EPI_WRITE        ? go(K::full_gather_kernel<false, K::EPI_WRITE>)
            : epi == K::EPI_ACCUMULATE ? go(K::full_gather_kernel<false, K::EPI_ACCUMULATE>)
                                       : go(K::full_gather_kernel<false, K::EPI_FINALIZE>);
        }
        break;
    }
    }
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
}

void launchResolve(const memb_b200_reader& r, const char* dQuery, const uint64_t* dOffsets, uint64_t n,
                   int32_t* dIds, cudaStream_t stream)
{
    if (n == 0) {
        return;
    }
    K::resolve_words_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, stream>>>(
        dQuery, dOffsets, n, r.dWords, r.dWordOffsets, r.dSlots, r.slotMask, dIds);
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
}

// Rows decoded per pass when the result has to travel to a host buffer: up to 4 GiB of fp32 per pass.
// Smaller passes whose copies overlap the next pass's decode (decodeTo is double-buffered) were measured
// r02 with 128 MiB passes: e2e of 1 M rows 22.41 ms = 95.2 % of the box's pure-copy time, against 98.7 %
// with ONE 1.2 GB copy -- ten copies cost more than the 0.3 ms kernel they hide.  So: one pass whenever
// the result fits, overlapped passes beyond.
uint64_t rowsPerPass(uint64_t rowFloats)
{
    const uint64_t forced = static_cast<uint64_t>(g_optPassRows.load());  // tests: exercise the multi-pass path
    if (forced) {
        return forced;
    }
    const uint64_t rows = (uint64_t(4) << 30) / (std::max<uint64_t>(rowFloats, 1) * 4);
    return std::max<uint64_t>(rows / 32 * 32, 32);
}

// The ReadersUnion merge as a sequence of fused epilogues
// (python/memb/readers_union.py:16-18,30-32):
//   concatenate: reader r WRITEs its columns [sum dim_<r, sum dim_<=r)
//   average:     reader 0 WRITEs, readers 1..R-2 ACCUMULATE, reader R-1 FINALIZEs
// There is never a per-reader temporary.  ids: R*n reader-major (device).
// true when this reader's launch into `o` uses a kernel that honours OutputSpec::sideHead/sideTail
bool canRedirectUnits(const memb_b200_reader& r, const K::OutputSpec& o)
{
    const uint32_t dim = r.file->dim;
    if (!vec4Ok(dim, o) || dim < 8) {
        return false;
    }
    return r.file->storage != STORAGE_TRAINED || r.trainedWarps != 0;
}

void launchUnion(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode, const int32_t* dIds,
                 uint64_t idStride, uint64_t first, uint64_t n, float* dOut, uint64_t unionDim, cudaStream_t stream)
{
    // trained readers: the whole union in ONE launch (trained_rows.cuh)
    if (tryLaunchRows(readers, R, mode, dIds + first, idStride, 0, n, dOut, stream)) {
        return;
    }
    std::vector<K::OutputSpec> specs(R);
    uint32_t col = 0;
    for (uint32_t k = 0; k < R; ++k) {
        K::OutputSpec& o = specs[k];
        o.out = dOut;
        o.rowStride = unionDim;
        o.divisor = static_cast<float>(R);
        o.colOffset = 0;
        if (mode == MEMB_B200_CONCATENATE) {
            o.colOffset = col;
            col += readers[k]->file->dim;
        }
    }
    // concatenate: block boundaries in the middle of a 32 B sector (dim 300: 1200 B = 37.5
    // sectors) are written whole by a stitch pass from two side buffers (OutputSpec::sideHead)
    float4* side = nullptr;
    std::vector<uint32_t> stitched;
    if (mode == MEMB_B200_CONCATENATE && n && (unionDim * 4) % 32 == 0 &&
        reinterpret_cast<uintptr_t>(dOut) % 32 == 0) {
        for (uint32_t k = 0; k + 1 < R; ++k) {
            const uint32_t boundary = specs[k + 1].colOffset;
            if ((boundary * 4) % 32 != 0 && canRedirectUnits(*readers[k], specs[k]) &&
                canRedirectUnits(*readers[k + 1], specs[k + 1])) {
                stitched.push_back(k);
            }
        }
        if (!stitched.empty()) {
            CUDA_CHECK(cudaMallocAsync(reinterpret_cast<void**>(&side), stitched.size() * 2 * n * sizeof(float4), stream));
            for (size_t b = 0; b < stitched.size(); ++b) {
                specs[stitched[b]].sideTail = side + (2 * b) * n;
                specs[stitched[b] + 1].sideHead = side + (2 * b + 1) * n;
            }
        }
    }
    // average, first two readers trained: reader 0 leaves symbol words, reader 1 gathers from
    // them while it writes the row (OutputSpec::sym): one write of the output instead of
    // write + read + write
    uint32_t* symWordsBuf = nullptr;
    auto tiledTrained = [&](const memb_b200_reader& r, const K::OutputSpec& o) {
        return r.file->storage == STORAGE_TRAINED && r.trainedWarps != 0 && !r.trainedEscape && vec4Ok(r.file->dim, o);
    };
    if (mode == MEMB_B200_AVERAGE && n && g_optSymAverage.load() && tiledTrained(*readers[0], specs[0]) &&
        tiledTrained(*readers[1], specs[1]) && readers[0]->trained.hasZeroSlot) {
        const uint32_t words = readers[0]->file->dim / 4;
        CUDA_CHECK(cudaMallocAsync(reinterpret_cast<void**>(&symWordsBuf), n * words * sizeof(uint32_t), stream));
        for (uint32_t k = 0; k < 2; ++k) {
            specs[k].sym = symWordsBuf;
            specs[k].symWords = words;
        }
        specs[1].symCodebook = readers[0]->trained.codebook;
        specs[1].symCodebookSize = readers[0]->trained.codebookSize;
        specs[1].symFieldShift = 2u - readers[0]->trained.symShift;
        // small codebooks: T[s0][s1] for every symbol pair (+0.0 slots of missing words included) instead of
        // two gathers + add + scale per element -- when the table fits beside this reader's warps
        const memb_b200_reader& r1 = *readers[1];
        uint32_t log2Stride = 0;
        while ((1u << log2Stride) < r1.trained.codebookSize + 1) {
            ++log2Stride;
        }
        const int64_t flavour = g_optPairTable.load();
        auto fits = [&](uint32_t tableWords) {
            const K::TrainedLayout lay = K::trainedLayout(
                r1.smemBase, K::trainedLutBytes(r1.trained.lutBits, r1.trained.lutReplicas),
                K::trainedWarpBytes(r1.trained.rowWords, 0u), r1.trainedWarps, tableWords);
            return lay.end - r1.smemBase <= r1.smemOptin;
        };
        // one copy of the table per bank (flavour 3): exact row stride, every byte offset within 16 bits
        const uint32_t exact = (readers[0]->trained.codebookSize + 1) * (r1.trained.codebookSize + 1);
        const uint32_t entries = (readers[0]->trained.codebookSize + 1) << log2Stride;
        if (flavour == 3 && exact * 128u <= 65536u && fits(exact * 32u)) {
            specs[1].pairEntries = exact;
            specs[1].pairStride = r1.trained.codebookSize + 1;
            specs[1].pairReplicas = 32;
            specs[1].straightPairs = 1u;
        } else if (flavour && entries <= 4096 && fits(entries)) {
            specs[1].pairEntries = entries;
            specs[1].pairStride = 1u << log2Stride;
            specs[1].pairReplicas = 1;
            specs[1].straightPairs = flavour >= 2 ? 1u : 0u;
        }
    }
    for (uint32_t k = 0; k < R; ++k) {
        int epi = K::EPI_WRITE;
        if (mode != MEMB_B200_CONCATENATE) {
            epi = k == 0 ? K::EPI_WRITE : (k + 1 == R ? K::EPI_FINALIZE : K::EPI_ACCUMULATE);
        }
        launchDecode(*readers[k], dIds + k * idStride + first, 0, n, specs[k], epi, stream);
    }
    if (symWordsBuf) {
        CUDA_CHECK(cudaFreeAsync(symWordsBuf, stream));
    }
    for (size_t b = 0; b < stitched.size(); ++b) {
        const unsigned grid = static_cast<unsigned>((2 * n + 255) / 256);
        K::stitch_sectors_kernel<<<grid, 256, 0, stream>>>(side + (2 * b) * n, side + (2 * b + 1) * n, n, dOut, unionDim,
                                                           specs[stitched[b] + 1].colOffset);
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
    }
    if (side) {
        CUDA_CHECK(cudaFreeAsync(side, stream));
    }
}

// A batch that repeats its words (Zipf-distributed text; kernels.cuh "repeated queries"): the distinct id tuples
// are decoded once into a temporary and full_gather_kernel copies the rows out.  Two host round trips of 4 bytes
// (how many distinct tuples a 8192-query sample holds, then the whole batch) decide; a batch that does not repeat
// pays the sample only.  Never under stream capture (the round trips synchronise), never for small batches.
constexpr uint64_t DEDUP_MIN_QUERIES = 1u << 17;
constexpr uint32_t DEDUP_SAMPLE = 8192;
bool tryLaunchUnionDistinct(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode, const int32_t* dIds,
                            uint64_t idStride, uint64_t first, uint64_t n, float* dOut, uint64_t unionDim,
                            cudaStream_t stream)
{
    const int64_t opt = g_optDedup.load();
    if (opt == 0 || n > 0x3fffffffull || (opt == 1 && n < DEDUP_MIN_QUERIES) || n < 2) {
        return false;
    }
    cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(stream, &capturing) != cudaSuccess || capturing != cudaStreamCaptureStatusNone) {
        cudaGetLastError();
        return false;
    }
    const memb_b200_reader& r0 = *readers[0];
    const uint32_t nq = static_cast<uint32_t>(n);
    // Is there anything to win?  Per query, in HBM-byte equivalents measured on configs[4] (r02): decoding a trained
    // row costs what ~1.65 x its fp32 bytes cost to move (0.30 ms per 1 M rows of 1200 B), uniform ~1.4 x, full 2 x
    // (read + write); the copy-out moves the union row twice.  The distinct path pays d decodes + n copies (+ the
    // claim, ~10 %): worth it below this fraction of distinct tuples.  Two trained readers: average 0.29 (the copy
    // is half a decode), concatenate: never (the copy-out of 2400 B rows costs more than decoding them).
    double decodeCost = 0;
    for (uint32_t k = 0; k < R; ++k) {
        const double perByte = readers[k]->file->storage == STORAGE_TRAINED ? 1.65 : readers[k]->file->storage == STORAGE_UNIFORM ? 1.4 : 2.0;
        decodeCost += perByte * 4.0 * readers[k]->file->dim;
    }
    const double maxDistinctFraction = 0.9 - 2.0 * 4.0 * static_cast<double>(unionDim) / decodeCost;
    if (opt == 1 && maxDistinctFraction < 0.05) {
        return false;
    }
    uint32_t tableSlots = 1024;
    while (tableSlots < 2 * n) {
        tableSlots *= 2;
    }
    // one or two readers: the tuple is the slot's 64-bit key (dedup_claim64_kernel); more: slots hold query indices
    const bool packed = R <= 2 && g_optDedupKeys.load() != 0;
    // one block: count | slots (x2 when 64-bit, + compactOfSlot) | rep or slotOf | compact | distinct ids (R x n)
    const uint64_t words = 64 + uint64_t(tableSlots) * (packed ? 3 : 1) + 2 * n + uint64_t(R) * n;
    int32_t* block = nullptr;
    if (cudaMallocAsync(reinterpret_cast<void**>(&block), words * 4, stream) != cudaSuccess) {
        cudaGetLastError();  // no room for the tables: decode every query
        return false;
    }
    struct Release {
        int32_t* p;
        float* rows = nullptr;
        cudaStream_t s;
        ~Release()
        {
            if (rows) {
                cudaFreeAsync(rows, s);
            }
            cudaFreeAsync(p, s);
        }
    } release{block, nullptr, stream};
    uint32_t* dCount = reinterpret_cast<uint32_t*>(block);
    int32_t* slots = block + 64;
    int32_t* compactOfSlot = slots + 2 * uint64_t(tableSlots);  // (packed only)
    int32_t* rep = slots + uint64_t(tableSlots) * (packed ? 3 : 1);
    int32_t* compact = rep + n;
    int32_t* distinctIds = compact + n;
    const int32_t* ids = dIds + first;
    auto claim = [&](uint32_t queries, uint32_t slotsUsed) -> uint32_t {
        CUDA_CHECK(cudaMemsetAsync(dCount, 0, 4, stream));
        CUDA_CHECK(cudaMemsetAsync(slots, 0xff, uint64_t(slotsUsed) * (packed ? 8 : 4), stream));
        if (packed) {
            K::dedup_claim64_kernel<<<(queries + K::DEDUP_BLOCK - 1) / K::DEDUP_BLOCK, K::DEDUP_BLOCK, 0, stream>>>(
                ids, idStride, R, queries, readers[0]->file->rows, R > 1 ? readers[1]->file->rows : 0u,
                reinterpret_cast<unsigned long long*>(slots), slotsUsed - 1, reinterpret_cast<uint32_t*>(rep), compactOfSlot,
                compact, distinctIds, n, dCount);
        } else {
            K::dedup_claim_kernel<<<(queries + K::DEDUP_BLOCK - 1) / K::DEDUP_BLOCK, K::DEDUP_BLOCK, 0, stream>>>(ids, idStride, R, queries, slots, slotsUsed - 1,
                                                                              rep, compact, distinctIds, n, dCount);
        }
        ++g_launches;
        CUDA_CHECK(cudaGetLastError());
        uint32_t count = 0;
        CUDA_CHECK(cudaMemcpyAsync(&count, dCount, 4, cudaMemcpyDeviceToHost, stream));
        CUDA_CHECK(cudaStreamSynchronize(stream));
        return count;
    };
    if (opt == 1) {
        const uint32_t sample = std::min<uint32_t>(nq, DEDUP_SAMPLE);
        if (claim(sample, 4 * DEDUP_SAMPLE) > sample - sample / 8) {
            return false;  // fewer than one query in eight repeats within the sample: decode them all
        }
    }
    const uint32_t distinct = claim(nq, tableSlots);
    if (opt == 1 && distinct > maxDistinctFraction * nq) {
        return false;  // the copy-out of every row costs more than the repeats save
    }
    if (packed) {
        K::dedup_follow64_kernel<<<(nq + 255) / 256, 256, 0, stream>>>(reinterpret_cast<const uint32_t*>(rep), compactOfSlot,
                                                                       compact, nq);
    } else {
        K::dedup_follow_kernel<<<(nq + 255) / 256, 256, 0, stream>>>(rep, compact, nq);
    }
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
    // Average of two trained readers with small codebooks: the temporary holds the distinct tuples' SYMBOL words
    // (two symbol-word launches) and pair_gather_kernel merges them on the way out -- half the bytes per output row
    // to read back, no fp32 flush for the distinct tuples.
    if (g_optDedupSymbols.load() && mode == MEMB_B200_AVERAGE && R == 2 && distinct > 0) {
        const memb_b200_reader& ra = *readers[0];
        const memb_b200_reader& rb = *readers[1];
        K::OutputSpec probe{dOut, unionDim, 0, 1.0f};
        auto tiled = [&](const memb_b200_reader& r) {
            return r.file->storage == STORAGE_TRAINED && r.trainedWarps != 0 && !r.trainedEscape && r.trained.hasZeroSlot &&
                   r.file->dim == unionDim && vec4Ok(r.file->dim, probe);
        };
        const uint32_t entries = (ra.trained.codebookSize + 1) * (rb.trained.codebookSize + 1);
        if (tiled(ra) && tiled(rb) && entries * 128u <= 48u * 1024u && g_optSymAverage.load()) {
            const uint32_t words = static_cast<uint32_t>(unionDim / 4);
            const uint64_t symWordsPerReader = uint64_t(distinct) * words;
            if (cudaMallocAsync(reinterpret_cast<void**>(&release.rows), 2 * symWordsPerReader * 4, stream) != cudaSuccess) {
                cudaGetLastError();
                release.rows = nullptr;
                return false;
            }
            uint32_t* sym = reinterpret_cast<uint32_t*>(release.rows);
            for (uint32_t k = 0; k < 2; ++k) {
                K::OutputSpec o{dOut, unionDim, 0, 1.0f};  // (a symbol-word launch writes nothing to `out`)
                o.sym = sym + k * symWordsPerReader;
                o.symWords = words;
                launchDecode(*readers[k], distinctIds + k * n, 0, distinct, o, K::EPI_WRITE, stream);
            }
            const uint64_t groups = (n + 31) / 32;
            const unsigned blocks = static_cast<unsigned>(std::min<uint64_t>(groups, uint64_t(r0.numSMs) * 6));
            K::pair_gather_kernel<<<blocks, 256, entries * 128u, stream>>>(
                sym, sym + symWordsPerReader, compact, n, words, ra.trained.codebook, ra.trained.codebookSize,
                2u - ra.trained.symShift, rb.trained.codebook, rb.trained.codebookSize, 2u - rb.trained.symShift,
                static_cast<float>(R), reinterpret_cast<float4*>(dOut));
            ++g_launches;
            CUDA_CHECK(cudaGetLastError());
            return true;
        }
    }
    if (cudaMallocAsync(reinterpret_cast<void**>(&release.rows), uint64_t(distinct) * unionDim * 4, stream) != cudaSuccess) {
        cudaGetLastError();  // no room for the temporary: decode every query
        release.rows = nullptr;
        return false;
    }
    launchUnion(readers, R, mode, distinctIds, n, 0, distinct, release.rows, unionDim, stream);
    // the copy-out: K3 over the temporary (compact index = row of the temporary; every index is in range)
    K::FullParams rows{release.rows, distinct, static_cast<uint32_t>(unionDim)};
    K::OutputSpec o{dOut, unionDim, 0, 1.0f};
    const bool vec = vec4Ok(static_cast<uint32_t>(unionDim), o);
    const uint64_t elems = n * (vec ? unionDim / 4 : unionDim);
    const unsigned grid = static_cast<unsigned>(std::min<uint64_t>((elems + 255) / 256, uint64_t(r0.numSMs) * 32));
    if (vec && g_optDedupKeep.load()) {
        K::full_gather_kernel<true, K::EPI_WRITE, false, true><<<grid, 256, 0, stream>>>(rows, compact, 0, n, o);
    } else if (vec) {
        K::full_gather_kernel<true, K::EPI_WRITE><<<grid, 256, 0, stream>>>(rows, compact, 0, n, o);
    } else {
        K::full_gather_kernel<false, K::EPI_WRITE><<<grid, 256, 0, stream>>>(rows, compact, 0, n, o);
    }
    ++g_launches;
    CUDA_CHECK(cudaGetLastError());
    return true;
}

void launchUnionBatch(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode, const int32_t* dIds,
                      uint64_t idStride, uint64_t first, uint64_t n, float* dOut, uint64_t unionDim, cudaStream_t stream)
{
    if (!tryLaunchUnionDistinct(readers, R, mode, dIds, idStride, first, n, dOut, unionDim, stream)) {
        launchUnion(readers, R, mode, dIds, idStride, first, n, dOut, unionDim, stream);
    }
}

uint64_t unionDimOf(memb_b200_reader* const* readers, uint32_t R, memb_b200_union_mode mode)
{
    if (R < 2) {
        // python/memb/readers_union.py:59-60
        fail(MEMB_B200_ERR_ARGUMENT, "You must pass at least 2 readers to create a union");
    }
    uint64_t total = 0;
    for (uint32_t k = 0; k < R; ++k) {
        if (!readers[k]) {
            fail(MEMB_B200_ERR_ARGUMENT, "null reader in union");
        }
        if (readers[k]->device != readers[0]->device) {
            fail(MEMB_B200_ERR_ARGUMENT, "all readers of a union must live on the same device");
        }
        if (mode == MEMB_B200_AVERAGE && readers[k]->file->dim != readers[0]->file->dim) {
            // python/memb/readers_union.py:7-10
            fail(MEMB_B200_ERR_ARGUMENT, "Dimensions of all readers must be equal for average mode");
        }
        total += readers[k]->file->dim;
    }
    return mode == MEMB_B200_AVERAGE ? readers[0]->file->dim : total;
}

}  // namespace

// ----------------------------------------------------------------- C ABI ----
extern "C" {

const char* memb_b200_last_error(void)
{
    return g_lastError.c_str();
}

int memb_b200_device_count(void)
{
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int usable = 0;
    for (int d = 0; d < count; ++d) {
        usable += deviceIsBlackwell(d) ? 1 : 0;
    }
    return usable;
}

uint64_t memb_b200_launch_count(void)
{
    return g_launches.load();
}

memb_b200_status memb_b200_set_option(const char* name, int64_t value)
{
    return guarded([&] {
        const std::string key = name ? name : "";
        if (key == "trained_lut_replicas" && (value == 0 || value == 1 || value == 32)) {
            g_optLutReplicas = value;
        } else if (key == "trained_warps" && value >= 0 && value <= K::TRAINED_MAX_WARPS) {
            g_optWarps = value;
        } else if (key == "trained_generic" && (value == 0 || value == 1)) {
            g_optGeneric = value;
        } else if (key == "union_symbol_average" && (value == 0 || value == 1)) {
            g_optSymAverage = value;
        } else if (key == "union_dedup" && value >= 0 && value <= 2) {
            g_optDedup = value;
        } else if (key == "union_dedup_symbols" && (value == 0 || value == 1)) {
            g_optDedupSymbols = value;
        } else if (key == "union_dedup_keep" && (value == 0 || value == 1)) {
            g_optDedupKeep = value;
        } else if (key == "union_dedup_keys" && (value == 0 || value == 1)) {
            g_optDedupKeys = value;
        } else if (key == "union_pair_table" && value >= 0 && value <= 3) {
            g_optPairTable = value;
        } else if (key == "uniform_rows_kernel" && value >= 0 && value <= 2) {
            g_optUniformRows = value;
        } else if (key == "trained_rows_kernel" && value >= 0 && value <= 2) {
            g_optTrainedRows = value;
        } else if (key == "trained_rows_warps" && value >= 0 && value <= K::ROWS_MAX_WARPS) {
            g_optRowsWarps = value;
        } else if (key == "trained_tma" && value >= 0 && value <= 2) {
            g_optTma = value;
        } else if (key == "trained_record_align" && (value == 1 || value == 2 || value == 4)) {
            g_optRecordAlign = value;
        } else if (key == "trained_fixed_records" && (value == 0 || value == 1)) {
            g_optFixedRecords = value;
        } else if (key == "host_pass_rows" && value >= 0) {
            g_optPassRows = value;  // rows staged per pass for host outputs (0 = 4 GiB of fp32)
        } else {
            fail(MEMB_B200_ERR_ARGUMENT, "unknown option or value out of range: " + key);
        }
    });
}

memb_b200_status memb_b200_open(const char* path, int device, unsigned num_threads, memb_b200_reader** out)
{
    return guarded([&] {
        if (!path || !out) {
            fail(MEMB_B200_ERR_ARGUMENT, "open: null argument");
        }
        *out = nullptr;
        auto reader = std::make_unique<memb_b200_reader>();
        const auto t0 = std::chrono::steady_clock::now();
        const unsigned threads = num_threads ? num_threads : std::max(std::thread::hardware_concurrency(), 2u);
        try {
            reader->file = parseMembFile(path, threads);
        } catch (const fb::FormatError&) {
            throw;
        } catch (const std::runtime_error& e) {
            const std::string what = e.what();
            if (what == "File format verification failed") {
                fail(MEMB_B200_ERR_FORMAT, what);
            }
            if (what.rfind("Storage strategy", 0) == 0) {
                fail(MEMB_B200_ERR_UNSUPPORTED, what);
            }
            fail(MEMB_B200_ERR_IO, what);
        }
        reader->stats.parse_ms = msSince(t0);

        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
            cudaGetLastError();
            fail(MEMB_B200_ERR_NO_DEVICE, "no CUDA device is visible; memb_b200 has no CPU path");
        }
        if (device < 0 || device >= count) {
            fail(MEMB_B200_ERR_ARGUMENT, "device ordinal out of range");
        }
        if (!deviceIsBlackwell(device)) {
            fail(MEMB_B200_ERR_NO_DEVICE, "device is not compute capability 10.x; kernels are built for sm_100a only");
        }
        reader->device = device;
        DeviceGuard guard(device);
        int value = 0;
        CUDA_CHECK(cudaDeviceGetAttribute(&value, cudaDevAttrMultiProcessorCount, device));
        reader->numSMs = value;
        CUDA_CHECK(cudaDeviceGetAttribute(&value, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));
        reader->smemOptin = static_cast<size_t>(value);
        CUDA_CHECK(cudaStreamCreateWithFlags(&reader->stream, cudaStreamNonBlocking));
        CUDA_CHECK(cudaStreamCreateWithFlags(&reader->copyStream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            CUDA_CHECK(cudaEventCreateWithFlags(&reader->passDecoded[k], cudaEventDisableTiming));
            CUDA_CHECK(cudaEventCreateWithFlags(&reader->passCopied[k], cudaEventDisableTiming));
        }
        {
            // stream-ordered scratch of launchUnion (symbol words of an average: dim bytes per row, 1.2 GB for
            // 4 M rows; side buffers of a concatenate): let the device's default pool KEEP what it has reserved
            // instead of returning it at every synchronisation.  With the 256 MiB of round 1 a 1 M-row average
            // (300 MB) paid a real device allocation after every sync: 0.76 -> up to 1.0 ms per call (r02).
            cudaMemPool_t pool = nullptr;
            if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
                uint64_t threshold = 0;
                cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
                if (threshold < (uint64_t(8) << 30)) {
                    threshold = uint64_t(8) << 30;
                    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
                }
            }
            cudaGetLastError();
        }

        // work counters of the dynamically scheduled kernels (trained tiles, uniform row groups)
        CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&reader->dTileCounters),
                              memb_b200_reader::TILE_COUNTERS * sizeof(uint32_t)));
        CUDA_CHECK(cudaMemset(reader->dTileCounters, 0, memb_b200_reader::TILE_COUNTERS * sizeof(uint32_t)));
        const auto t1 = std::chrono::steady_clock::now();
        buildVocabulary(*reader);
        reader->stats.upload_ms += msSince(t1);
        switch (reader->file->storage) {
        case STORAGE_TRAINED:
            loadTrained(*reader);
            break;
        case STORAGE_UNIFORM:
            loadUniform(*reader);
            break;
        default:
            loadFull(*reader);
            break;
        }
        const auto t2 = std::chrono::steady_clock::now();
        CUDA_CHECK(cudaStreamSynchronize(reader->stream));
        reader->stats.prepare_ms += msSince(t2);
        try {
            noallocIdsFor(device);  // set-up of memb_b200_lookup_words_noalloc: symbol address, once per device
        } catch (const StatusError&) {
            cudaGetLastError();  // not fatal for open: the entry point retries and reports
        }
        // the mmap is only needed for the vocabulary text of trained files
        if (reader->file->storage != STORAGE_TRAINED) {
            reader->file->map.reset();
        }
        *out = reader.release();
    });
}

void memb_b200_close(memb_b200_reader* reader)
{
    if (!reader) {
        return;
    }
    int previous = -1;
    cudaGetDevice(&previous);
    cudaSetDevice(reader->device);
    delete reader;
    if (previous >= 0) {
        cudaSetDevice(previous);
    }
}

uint32_t memb_b200_dim(const memb_b200_reader* reader)
{
    return reader ? reader->file->dim : 0;
}

uint64_t memb_b200_rows(const memb_b200_reader* reader)
{
    return reader ? reader->file->rows : 0;
}

int memb_b200_storage_type(const memb_b200_reader* reader)
{
    return reader ? reader->file->storage : 0;
}

int memb_b200_device(const memb_b200_reader* reader)
{
    return reader ? reader->device : -1;
}

memb_b200_status memb_b200_get_load_stats(const memb_b200_reader* reader, memb_b200_load_stats* stats)
{
    return guarded([&] {
        if (!reader || !stats) {
            fail(MEMB_B200_ERR_ARGUMENT, "get_load_stats: null argument");
        }
        *stats = reader->stats;
    });
}

memb_b200_status memb_b200_keys(const memb_b200_reader* reader, const char** words, const uint32_t** offsets,
                                uint64_t* rows)
{
    return guarded([&] {
        if (!reader || !words || !offsets || !rows) {
            fail(MEMB_B200_ERR_ARGUMENT, "keys: null argument");
        }
        *words = reader->file->words;
        *offsets = reader->file->wordOffsets.data();
        *rows = reader->file->rows;
    });
}

}  // extern "C"

namespace {

// Uploads the query text and resolves it to device ids in scratchIds.
int32_t* resolveToDevice(memb_b200_reader& r, const char* queryBytes, const uint64_t* queryOffsets, uint64_t n,
                         cudaStream_t stream)
{
    int32_t* dIds = static_cast<int32_t*>(r.scratchIds.ensure(std::max<uint64_t>(n, 1) * 4));
    if (n == 0) {
        return dIds;
    }
    if (!queryBytes || !queryOffsets) {
        fail(MEMB_B200_ERR_ARGUMENT, "null query buffer");
    }
    const uint64_t base = queryOffsets[0];
    const uint64_t bytes = queryOffsets[n] - base;
    char* dQuery = static_cast<char*>(r.scratchQuery.ensure(std::max<uint64_t>(bytes, 1)));
    uint64_t* dOffsets = static_cast<uint64_t*>(r.scratchOffsets.ensure((n + 1) * 8));
    CUDA_CHECK(cudaMemcpyAsync(dQuery, queryBytes + base, bytes, cudaMemcpyHostToDevice, stream));
    CUDA_CHECK(cudaMemcpyAsync(dOffsets, queryOffsets, (n + 1) * 8, cudaMemcpyHostToDevice, stream));
    // offsets are relative to the blob start; the kernel sees the blob from `base`
    launchResolve(r, dQuery - base, dOffsets, n, dIds, stream);
    return dIds;
}

// Decodes n rows of a reader/union into `out` (host or device), in passes when
// the result has to be staged for a host buffer.
template <class Launch>
void decodeTo(memb_b200_reader& r, uint64_t n, uint64_t rowFloats, float* out, memb_b200_mem_kind outKind,
              cudaStream_t stream, Launch launch)
{
    if (n == 0 || rowFloats == 0) {
        return;
    }
    if (!out) {
        fail(MEMB_B200_ERR_ARGUMENT, "null output buffer");
    }
    if (outKind == MEMB_B200_DEVICE) {
        launch(uint64_t(0), n, out);
        return;
    }
    // Host output (always on r.stream, under r.mutex): passes of rowsPerPass rows through two staging
    // buffers.  Pass k is decoded on `stream`; its D2H copy runs on r.copyStream once the decode is done,
    // while pass k + 1 is already being decoded into the other buffer.  (One pass for results of up to
    // 4 GiB, see rowsPerPass.)
    const uint64_t pass = rowsPerPass(rowFloats);
    const uint64_t passElems = std::min(n, pass) * rowFloats;
    float* dOut = static_cast<float*>(r.scratchOut.ensure((n > pass ? 2 : 1) * passElems * 4));
    uint64_t k = 0;
    for (uint64_t first = 0; first < n; first += pass, ++k) {
        const uint64_t count = std::min(pass, n - first);
        const int buf = static_cast<int>(k & 1);
        float* dPass = dOut + buf * passElems;
        if (k >= 2) {
            CUDA_CHECK(cudaStreamWaitEvent(stream, r.passCopied[buf], 0));  // the buffer's previous pass has left it
        }
        launch(first, count, dPass);
        CUDA_CHECK(cudaEventRecord(r.passDecoded[buf], stream));
        CUDA_CHECK(cudaStreamWaitEvent(r.copyStream, r.passDecoded[buf], 0));
        CUDA_CHECK(cudaMemcpyAsync(out + first * rowFloats, dPass, count * rowFloats * 4, cudaMemcpyDeviceToHost,
                                   r.copyStream));
        CUDA_CHECK(cudaEventRecord(r.passCopied[buf], r.copyStream));
    }
    CUDA_CHECK(cudaStreamSynchronize(r.copyStream));
    CUDA_CHECK(cudaStreamSynchronize(stream));
}

const int32_t* idsOnDevice(memb_b200_reader& r, const int32_t* ids, memb_b200_mem_kind kind, uint64_t count,
                           cudaStream_t stream)
{
    if (count == 0) {
        return ids;
    }
    if (!ids) {
        fail(MEMB_B200_ERR_ARGUMENT, "null ids buffer");
    }
    if (kind == MEMB_B200_DEVICE) {
        return ids;
    }
    int32_t* dIds = static_cast<int32_t*>(r.scratchIds.ensure(count * 4));
    CUDA_CHECK(cudaMemcpyAsync(dIds, ids, count * 4, cudaMemcpyHostToDevice, stream));
    return dIds;
}

}  // namespace

extern "C" {

memb_b200_status memb_b200_resolve(memb_b200_reader* reader, const char* query_bytes, const uint64_t* query_offsets,
                                   uint64_t n, int32_t* ids_host)
{
    return guarded([&] {
        if (!reader || (n && !ids_host)) {
            fail(MEMB_B200_ERR_ARGUMENT, "resolve: null argument");
        }
        DeviceGuard guard(reader->device);
        std::lock_guard<std::mutex> lock(reader->mutex);
        const int32_t* dIds = resolveToDevice(*reader, query_bytes, query_offsets, n, reader->stream);
        if (n) {
            CUDA_CHECK(cudaMemcpyAsync(ids_host, dIds, n * 4, cudaMemcpyDeviceToHost, reader->stream));
        }
        CUDA_CHECK(cudaStreamSynchronize(reader->stream));
    });
}

memb_b200_status memb_b200_resolve_device(memb_b200_reader* reader, const char* query_bytes_device,
                                          const uint64_t* query_offsets_device, uint64_t n, int32_t* ids_device,
                                          void* stream)
{
    return guarded([&] {
        if (!reader || (n && (!query_bytes_device || !query_offsets_device || !ids_device))) {
            fail(MEMB_B200_ERR_ARGUMENT, "resolve_device: null argument");
        }
        DeviceGuard guard(reader->device);
        launchResolve(*reader, query_bytes_device, query_offsets_device, n, ids_device, static_cast<cudaStream_t>(stream));
    });
}

memb_b200_status memb_b200_lookup_words_noalloc(memb_b200_reader* reader, const char* query_bytes_device,
                                                const uint64_t* query_offsets_device, uint64_t n, float* out_device,
                                                void* stream)
{
    return guarded([&] {
        if (!reader) {
            fail(MEMB_B200_ERR_ARGUMENT, "lookup_words_noalloc: null reader");
        }
        if (n > MEMB_B200_NOALLOC_MAX_WORDS) {
            fail(MEMB_B200_ERR_ARGUMENT, "lookup_words_noalloc: " + std::to_string(n) + " words exceed the static scratch of " +
                                             std::to_string(MEMB_B200_NOALLOC_MAX_WORDS) +
                                             " (MEMB_B200_NOALLOC_MAX_WORDS); split the batch or use memb_b200_lookup_words_on");
        }
        if (n == 0) {
            return;
        }
        if (!query_bytes_device || !query_offsets_device || !out_device) {
            fail(MEMB_B200_ERR_ARGUMENT, "lookup_words_noalloc: null buffer");
        }
        memb_b200_reader& r = *reader;
        DeviceGuard guard(r.device);
        int32_t* dIds = noallocIdsFor(r.device);  // a device address, NOT the host-side name g_noallocIds
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        launchResolve(r, query_bytes_device, query_offsets_device, n, dIds, s);
        const uint32_t dim = r.file->dim;
        K::OutputSpec o{out_device, dim, 0, 1.0f};
        launchDecode(r, dIds, 0, n, o, K::EPI_WRITE, s);
    });
}

static memb_b200_status lookupWords(memb_b200_reader* reader, const char* query_bytes, const uint64_t* query_offsets,
                                    uint64_t n, float* out, memb_b200_mem_kind out_kind, bool callerStream,
                                    void* stream)
{
    return guarded([&] {
        if (!reader) {
            fail(MEMB_B200_ERR_ARGUMENT, "lookup_words: null reader");
        }
        DeviceGuard guard(reader->device);
        std::lock_guard<std::mutex> lock(reader->mutex);
        memb_b200_reader& r = *reader;
        cudaStream_t s = callerStream ? static_cast<cudaStream_t>(stream) : r.stream;
        const int32_t* dIds = resolveToDevice(r, query_bytes, query_offsets, n, s);
        const uint32_t dim = r.file->dim;
        decodeTo(r, n, dim, out, out_kind, s, [&](uint64_t first, uint64_t count, float* dOut) {
            K::OutputSpec o{dOut, dim, 0, 1.0f};
            launchDecode(r, dIds + first, 0, count, o, K::EPI_WRITE, s);
        });
        // the scratch ids/query buffers are reused by the next call: finish before returning
        CUDA_CHECK(cudaStreamSynchronize(s));
    });
}

memb_b200_status memb_b200_lookup_words_to(memb_b200_reader* reader, const char* query_bytes,
                                           const uint64_t* query_offsets, uint64_t n, float* out,
                                           memb_b200_mem_kind out_kind)
{
    return lookupWords(reader, query_bytes, query_offsets, n, out, out_kind, false, nullptr);
}

memb_b200_status memb_b200_lookup_words_on(memb_b200_reader* reader, const char* query_bytes,
                                           const uint64_t* query_offsets, uint64_t n, float* out_device, void* stream)
{
    return lookupWords(reader, query_bytes, query_offsets, n, out_device, MEMB_B200_DEVICE, true, stream);
}

memb_b200_status memb_b200_lookup_words(memb_b200_reader* reader, const char* query_bytes,
                                        const uint64_t* query_offsets, uint64_t n, float* out_host)
{
    return memb_b200_lookup_words_to(reader, query_bytes, query_offsets, n, out_host, MEMB_B200_HOST);
}

memb_b200_status memb_b200_decode_ids(memb_b200_reader* reader, const int32_t* ids, memb_b200_mem_kind ids_kind,
                                      uint64_t n, float* out, memb_b200_mem_kind out_kind, void* stream)
{
    return guarded([&] {
        if (!reader) {
            fail(MEMB_B200_ERR_ARGUMENT, "decode_ids: null reader");
        }
        DeviceGuard guard(reader->device);
        memb_b200_reader& r = *reader;
        const uint32_t dim = r.file->dim;
        const bool async = ids_kind == MEMB_B200_DEVICE && out_kind == MEMB_B200_DEVICE;
        cudaStream_t s = out_kind == MEMB_B200_DEVICE ? static_cast<cudaStream_t>(stream) : r.stream;
        std::unique_lock<std::mutex> lock(r.mutex, std::defer_lock);
        if (!async) {
            lock.lock();
        }
        const int32_t* dIds = idsOnDevice(r, ids, ids_kind, n, s);
        decodeTo(r, n, dim, out, out_kind, s, [&](uint64_t first, uint64_t count, float* dOut) {
            K::OutputSpec o{dOut, dim, 0, 1.0f};
            launchDecode(r, dIds + first, 0, count, o, K::EPI_WRITE, s);
        });
        if (!async) {
            CUDA_CHECK(cudaStreamSynchronize(s));
        }
    });
}

memb_b200_status memb_b200_decode_range(memb_b200_reader* reader, uint64_t row_begin, uint64_t row_end, float* out,
                                        memb_b200_mem_kind out_kind, void* stream)
{
    return guarded([&] {
        if (!reader) {
            fail(MEMB_B200_ERR_ARGUMENT, "decode_range: null reader");
        }
        if (row_begin > row_end || row_end > reader->file->rows) {
            fail(MEMB_B200_ERR_ARGUMENT, "decode_range: rows out of range");
        }
        DeviceGuard guard(reader->device);
        memb_b200_reader& r = *reader;
        const uint32_t dim = r.file->dim;
        const bool async = out_kind == MEMB_B200_DEVICE;
        cudaStream_t s = async ? static_cast<cudaStream_t>(stream) : r.stream;
        std::unique_lock<std::mutex> lock(r.mutex, std::defer_lock);
        if (!async) {
            lock.lock();
        }
        decodeTo(r, row_end - row_begin, dim, out, out_kind, s, [&](uint64_t first, uint64_t count, float* dOut) {
            K::OutputSpec o{dOut, dim, 0, 1.0f};
            launchDecode(r, nullptr, static_cast<uint32_t>(row_begin + first), count, o, K::EPI_WRITE, s);
        });
        if (!async) {
            CUDA_CHECK(cudaStreamSynchronize(s));
        }
    });
}

memb_b200_status memb_b200_decode_range_multicast(memb_b200_reader* reader, uint64_t row_begin, uint64_t row_end,
                                                  float* out_multicast, void* stream)
{
    return guarded([&] {
        if (!reader || !out_multicast) {
            fail(MEMB_B200_ERR_ARGUMENT, "decode_range_multicast: null argument");
        }
        if (row_begin > row_end || row_end > reader->file->rows) {
            fail(MEMB_B200_ERR_ARGUMENT, "decode_range_multicast: rows out of range");
        }
        memb_b200_reader& r = *reader;
        const uint32_t dim = r.file->dim;
        if (r.file->storage != STORAGE_TRAINED || r.trainedWarps == 0 || dim % 4 != 0 ||
            reinterpret_cast<uintptr_t>(out_multicast) % 16 != 0) {
            fail(MEMB_B200_ERR_UNSUPPORTED, "decode_range_multicast: trained storage with dim % 4 == 0 only");
        }
        DeviceGuard guard(r.device);
        K::OutputSpec o{out_multicast, dim, 0, 1.0f};
        o.multicast = 1;
        if (row_end > row_begin) {
            launchTrainedG<true>(r, nullptr, static_cast<uint32_t>(row_begin), row_end - row_begin, o, K::EPI_WRITE,
                                 static_cast<cudaStream_t>(stream));
            ++g_launches;
            CUDA_CHECK(cudaGetLastError());
        }
    });
}

memb_b200_status memb_b200_union_decode_ids(memb_b200_reader* const* readers, uint32_t num_readers,
                                            memb_b200_union_mode mode, const int32_t* ids,
                                            memb_b200_mem_kind ids_kind, uint64_t n, float* out,
                                            memb_b200_mem_kind out_kind, void* stream)
{
    return guarded([&] {
        if (!readers) {
            fail(MEMB_B200_ERR_ARGUMENT, "union_decode_ids: null readers");
        }
        const uint64_t unionDim = unionDimOf(readers, num_readers, mode);
        memb_b200_reader& r = *readers[0];
        DeviceGuard guard(r.device);
        const bool async = ids_kind == MEMB_B200_DEVICE && out_kind == MEMB_B200_DEVICE;
        cudaStream_t s = out_kind == MEMB_B200_DEVICE ? static_cast<cudaStream_t>(stream) : r.stream;
        std::unique_lock<std::mutex> lock(r.mutex, std::defer_lock);
        if (!async) {
            lock.lock();
        }
        const int32_t* dIds = idsOnDevice(r, ids, ids_kind, n * num_readers, s);
        decodeTo(r, n, unionDim, out, out_kind, s, [&](uint64_t first, uint64_t count, float* dOut) {
            launchUnionBatch(readers, num_readers, mode, dIds, n, first, count, dOut, unionDim, s);
        });
        if (!async) {
            CUDA_CHECK(cudaStreamSynchronize(s));
        }
    });
}

static memb_b200_status unionLookupWords(memb_b200_reader* const* readers, uint32_t num_readers,
                                         memb_b200_union_mode mode, const char* query_bytes,
                                         const uint64_t* query_offsets, uint64_t n, float* out,
                                         memb_b200_mem_kind out_kind, bool callerStream, void* stream)
{
    return guarded([&] {
        if (!readers) {
            fail(MEMB_B200_ERR_ARGUMENT, "union_lookup_words: null readers");
        }
        const uint64_t unionDim = unionDimOf(readers, num_readers, mode);
        memb_b200_reader& r = *readers[0];
        DeviceGuard guard(r.device);
        std::lock_guard<std::mutex> lock(r.mutex);
        if (n == 0) {
            return;
        }
        if (!query_bytes || !query_offsets) {
            fail(MEMB_B200_ERR_ARGUMENT, "null query buffer");
        }
        cudaStream_t s = callerStream ? static_cast<cudaStream_t>(stream) : r.stream;
        // one upload of the query text, one resolve per reader (each has its own vocabulary)
        const uint64_t base = query_offsets[0];
        const uint64_t bytes = query_offsets[n] - base;
        char* dQuery = static_cast<char*>(r.scratchQuery.ensure(std::max<uint64_t>(bytes, 1)));
        uint64_t* dOffsets = static_cast<uint64_t*>(r.scratchOffsets.ensure((n + 1) * 8));
        int32_t* dIds = static_cast<int32_t*>(r.scratchIds.ensure(n * num_readers * 4));
        CUDA_CHECK(cudaMemcpyAsync(dQuery, query_bytes + base, bytes, cudaMemcpyHostToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(dOffsets, query_offsets, (n + 1) * 8, cudaMemcpyHostToDevice, s));
        for (uint32_t k = 0; k < num_readers; ++k) {
            launchResolve(*readers[k], dQuery - base, dOffsets, n, dIds + k * n, s);
        }
        decodeTo(r, n, unionDim, out, out_kind, s, [&](uint64_t first, uint64_t count, float* dOut) {
            launchUnionBatch(readers, num_readers, mode, dIds, n, first, count, dOut, unionDim, s);
        });
        CUDA_CHECK(cudaStreamSynchronize(s));
    });
}

memb_b200_status memb_b200_union_lookup_words_to(memb_b200_reader* const* readers, uint32_t num_readers,
                                                 memb_b200_union_mode mode, const char* query_bytes,
                                                 const uint64_t* query_offsets, uint64_t n, float* out,
                                                 memb_b200_mem_kind out_kind)
{
    return unionLookupWords(readers, num_readers, mode, query_bytes, query_offsets, n, out, out_kind, false, nullptr);
}

memb_b200_status memb_b200_union_lookup_words_on(memb_b200_reader* const* readers, uint32_t num_readers,
                                                 memb_b200_union_mode mode, const char* query_bytes,
                                                 const uint64_t* query_offsets, uint64_t n, float* out_device,
                                                 void* stream)
{
    return unionLookupWords(readers, num_readers, mode, query_bytes, query_offsets, n, out_device, MEMB_B200_DEVICE,
                            true, stream);
}

memb_b200_status memb_b200_union_lookup_words(memb_b200_reader* const* readers, uint32_t num_readers,
                                              memb_b200_union_mode mode, const char* query_bytes,
                                              const uint64_t* query_offsets, uint64_t n, float* out_host)
{
    return memb_b200_union_lookup_words_to(readers, num_readers, mode, query_bytes, query_offsets, n, out_host,
                                           MEMB_B200_HOST);
}

memb_b200_status memb_b200_host_alloc(size_t bytes, void** ptr)
{
    return guarded([&] {
        if (!ptr) {
            fail(MEMB_B200_ERR_ARGUMENT, "host_alloc: null argument");
        }
        *ptr = nullptr;
        CUDA_CHECK(cudaHostAlloc(ptr, std::max<size_t>(bytes, 1), cudaHostAllocPortable));
    });
}

void memb_b200_host_free(void* ptr)
{
    if (ptr) {
        cudaFreeHost(ptr);
    }
}

// ---- Builder ----
struct memb_b200_builder {
    Builder impl;
    memb_b200_builder(size_t dim, const std::string& name, size_t bits) : impl(dim, name, bits) {}
};

memb_b200_status memb_b200_builder_create(uint64_t dim, const char* storage_name, uint64_t bits_per_weight,
                                          memb_b200_builder** out)
{
    const memb_b200_status st = guarded([&] {
        if (!storage_name || !out) {
            fail(MEMB_B200_ERR_ARGUMENT, "builder_create: null argument");
        }
        *out = nullptr;
        try {
            *out = new memb_b200_builder(dim, storage_name, bits_per_weight);
        } catch (const std::runtime_error& e) {
            fail(MEMB_B200_ERR_UNSUPPORTED, e.what());
        }
    });
    return st;
}

void memb_b200_builder_destroy(memb_b200_builder* builder)
{
    delete builder;
}

memb_b200_status memb_b200_builder_add_word(memb_b200_builder* builder, const char* word, size_t word_len,
                                            const float* values, uint64_t count)
{
    return guarded([&] {
        if (!builder || !word || (count && !values)) {
            fail(MEMB_B200_ERR_ARGUMENT, "builder_add_word: null argument");
        }
        try {
            builder->impl.addWord(std::string(word, word_len), values, count);
        } catch (const std::runtime_error& e) {
            fail(MEMB_B200_ERR_BUILDER, e.what());
        }
    });
}

memb_b200_status memb_b200_builder_add_words(memb_b200_builder* builder, const char* blob, const uint64_t* offsets,
                                             uint64_t n, const float* matrix)
{
    return guarded([&] {
        if (!builder || (n && (!blob || !offsets || !matrix))) {
            fail(MEMB_B200_ERR_ARGUMENT, "builder_add_words: null argument");
        }
        try {
            builder->impl.addWords(blob, offsets, n, matrix);
        } catch (const std::runtime_error& e) {
            fail(MEMB_B200_ERR_BUILDER, e.what());
        }
    });
}

memb_b200_status memb_b200_builder_set_device(memb_b200_builder* builder, int device)
{
    return guarded([&] {
        if (!builder) {
            fail(MEMB_B200_ERR_ARGUMENT, "builder_set_device: null builder");
        }
        if (device < -1 || device >= deviceEncoderCount()) {
            fail(device >= 0 && deviceEncoderCount() == 0 ? MEMB_B200_ERR_NO_DEVICE : MEMB_B200_ERR_ARGUMENT,
                 "builder_set_device: no CUDA device " + std::to_string(device));
        }
        builder->impl.setDevice(device);
    });
}

memb_b200_status memb_b200_builder_save(memb_b200_builder* builder, const char* path, unsigned num_threads)
{
    return guarded([&] {
        if (!builder || !path) {
            fail(MEMB_B200_ERR_ARGUMENT, "builder_save: null argument");
        }
        builder->impl.save(path, num_threads);
    });
}

memb_b200_status memb_b200_kmeans_fit_predict(const float* data, uint64_t n, uint32_t clusters, uint8_t* ids,
                                              float* centroids, uint32_t* num_centroids)
{
    return guarded([&] {
        if (!data || !ids || !centroids || !num_centroids || n == 0 || clusters < 2 || clusters > 255) {
            fail(MEMB_B200_ERR_ARGUMENT, "kmeans_fit_predict: bad argument");
        }
        KMeans1D km;
        km.fit(data, n, clusters);
        for (uint64_t i = 0; i < n; ++i) {
            ids[i] = km.predictOne(data[i]);
        }
        *num_centroids = static_cast<uint32_t>(km.centroids.size());
        std::copy(km.centroids.begin(), km.centroids.end(), centroids);
    });
}

const char* memb_b200_available_compression_strategies(void)
{
    return "full\0uniform\0trained\0";
}

}  // extern "C"
