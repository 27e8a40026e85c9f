// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  The product (memb_b200/) never links, imports or calls it.
//
// CPU restatement of the reference's batched embedding lookup, one function per
// reference function, each citing the file:line it follows (relative to
// /root/reference).  Plain C++14, built `-O3 -ffp-contract=off` (the reference is
// built -O3 for baseline x86-64, CMakeLists.txt:15, so no FMA is ever formed).
//
// Parity status: PINNED.
//   - bit order of the stream: known-answer test src/bit_stream_tests.cpp:31-59
//   - keys order, per-word values (1 %), miss -> zeros: src/tests.cpp:20-57,61-74
//   - table width does not change results (L = 1 vs 10): src/tests.cpp:76-88
//   - threaded == serial for 1025 words: src/tests.cpp:90-113
//   - bad identifier / missing file raise: src/tests.cpp:134-153
//   - and, where /root/reference exists, bit-for-bit against the reference's own
//     huffman_table_decoder.h / bit_stream_reader.h / prefix_code.cpp /
//     bit_stream.h compiled from where they lie (oracle/Makefile -> oracle/_ref/).
// With -DMEMB_ORACLE_USE_REFERENCE_CODEC (the oracle/_ref build) the Huffman
// decode below IS the reference's header, not the restatement.
//
// The FlatBuffers dependency (1.9.0, pinned in
// tools/development_image/install_deps_linux.sh:6) is absent from the reference
// tree and this image; table/vector access below restates its published binary
// layout for the six schemas in src/flatbuffers/*.fbs.  Like the reference
// (src/reader.cpp:104-111) the oracle checks the identifier and nothing else.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <future>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
#include <stddef.h>
#include <limits>
#include "huffman_table_decoder.h"  // from /root/reference/src (see oracle/Makefile)
#endif

namespace oracle {

// ------------------------------------------------------------ flatbuffers ----
struct Buf {
    const uint8_t* p;
    size_t n;
    template <class T>
    T rd(size_t pos) const
    {
        T v;
        std::memcpy(&v, p + pos, sizeof(T));
        return v;
    }
    size_t follow(size_t pos) const { return pos + rd<uint32_t>(pos); }
    // absolute position of field `id` of the table at `t`, 0 when absent
    size_t field(size_t t, unsigned id) const
    {
        const size_t vt = t - rd<int32_t>(t);
        const unsigned slot = 4 + 2 * id;
        if (slot + 2 > rd<uint16_t>(vt)) {
            return 0;
        }
        const uint16_t rel = rd<uint16_t>(vt + slot);
        return rel ? t + rel : 0;
    }
    size_t sub(size_t t, unsigned id) const
    {
        const size_t f = field(t, id);
        return f ? follow(f) : 0;
    }
    uint32_t vecLen(size_t v) const { return rd<uint32_t>(v); }
    size_t vecData(size_t v) const { return v + 4; }
};

// --------------------------------------------------------------- codec ----
// createCanonicalPrefixCodes, src/prefix_code.cpp:5-22; lengths from
// size_offsets as in src/huffman_table_decoder.h:31-37.
struct Code {
    uint16_t code;
    uint8_t len;
};

static std::vector<Code> canonicalCodes(const std::vector<uint8_t>& keys, const std::vector<uint32_t>& sizeOffsets)
{
    std::vector<Code> codes(keys.size());
    size_t current = 0;
    uint16_t code = 0;
    size_t bits = 0;
    for (size_t i = 0; i < keys.size(); ++i) {
        while (i >= sizeOffsets[current]) {
            ++current;
        }
        while (bits < current) {
            ++bits;
            code = static_cast<uint16_t>(code << 1);
        }
        codes[i] = Code{code, static_cast<uint8_t>(bits)};
        ++code;
    }
    return codes;
}

// MSB-first bit window over [p, end); bytes past `end` read as zero
// (BitStreamReader::pull, src/bit_stream_reader.h:16-31).
struct BitReader {
    const uint8_t* p;
    const uint8_t* end;
    uint64_t acc = 0;
    int have = 0;
    BitReader(const uint8_t* b, const uint8_t* e) : p(b), end(e) {}
    uint32_t peek(int n)
    {
        while (have < n) {
            acc <<= 8;
            if (p < end) {
                acc |= *p++;
            }
            have += 8;
        }
        return static_cast<uint32_t>((acc >> (have - n)) & ((1ull << n) - 1));
    }
    void skip(int n) { have -= n; }
};

// HuffmanTableDecoder, src/huffman_table_decoder.h:18-118: a direct table over
// the first L bits; longer codes are finished by canonical search (the reference
// uses per-prefix sub-tables, which decode the same prefix code).
struct TableDecoder {
    std::vector<uint8_t> keys;
    std::vector<Code> codes;
    std::vector<uint16_t> direct;  // key | len << 8 ; len == 0xff => long code
    int L = 10;
    int maxLen = 0;

    void build(const std::vector<uint8_t>& k, const std::vector<uint32_t>& so, int tableBits)
    {
        keys = k;
        codes = canonicalCodes(k, so);
        L = tableBits;
        maxLen = codes.empty() ? 0 : codes.back().len;
        direct.assign(size_t(1) << L, 0xff00);
        for (size_t i = 0; i < keys.size(); ++i) {
            const int len = codes[i].len;
            if (len <= L) {
                const size_t first = static_cast<size_t>(codes[i].code) << (L - len);
                for (size_t r = 0; r < (size_t(1) << (L - len)); ++r) {
                    direct[first + r] = static_cast<uint16_t>(keys[i] | (len << 8));
                }
            }
        }
    }

    uint8_t next(BitReader& br) const
    {
        const uint16_t e = direct[br.peek(L)];
        if ((e >> 8) != 0xff) {
            br.skip(e >> 8);
            return static_cast<uint8_t>(e);
        }
        const uint32_t w = br.peek(16);
        for (size_t i = 0; i < keys.size(); ++i) {
            const int len = codes[i].len;
            if (len > L && (w >> (16 - len)) == codes[i].code) {
                br.skip(len);
                return keys[i];
            }
        }
        return 0;  // not a code word: cannot happen for a Huffman code
    }
};

// ------------------------------------------------------------- storages ----
struct Reader {
    std::vector<uint8_t> bytes;
    Buf b{nullptr, 0};
    int storage = 0;
    uint32_t dim = 0;
    size_t storageTable = 0;
    size_t numThreads = 1;
    uint64_t rows = 0;

    // trained
    size_t wordOffsets = 0, valueOffsets = 0, packedWords = 0, packedValues = 0;
    uint32_t packedSize = 0;
    std::vector<uint8_t> hkeys;
    std::vector<uint32_t> hsizes;
    TableDecoder table;
#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
    std::unique_ptr<memb::HuffmanTableDecoder> refTable;
#endif
    std::vector<float> centroids;
    // uniform / full
    size_t nodes = 0;
    uint8_t levels = 0;

    void setTableBits(int L)
    {
        table.build(hkeys, hsizes, L);
#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
        refTable.reset(new memb::HuffmanTableDecoder(hkeys, hsizes, static_cast<size_t>(L)));
#endif
    }

    // Reader::Reader + getIndexChecked, src/reader.cpp:23-29,104-111;
    // adjustedNumThreads, src/reader.cpp:113-120
    void open(const char* path, size_t threads)
    {
        numThreads = threads > 0 ? threads : std::max(std::thread::hardware_concurrency(), 2u);
        FILE* f = std::fopen(path, "rb");
        if (!f) {
            throw std::runtime_error(std::string("cannot open ") + path);
        }
        std::fseek(f, 0, SEEK_END);
        const long size = std::ftell(f);
        std::fseek(f, 0, SEEK_SET);
        bytes.resize(static_cast<size_t>(size));
        if (size > 0 && std::fread(bytes.data(), 1, bytes.size(), f) != bytes.size()) {
            std::fclose(f);
            throw std::runtime_error("short read");
        }
        std::fclose(f);
        if (bytes.size() < 8 || std::memcmp(bytes.data() + 4, "memb", 4) != 0) {
            throw std::runtime_error("File format verification failed");
        }
        b = Buf{bytes.data(), bytes.size()};
        const size_t root = b.rd<uint32_t>(0);
        // table Index { storage_type (id 0), storage (id 1), dim (id 2) } — embeddings.fbs:13-16
        const size_t ft = b.field(root, 0);
        storage = ft ? b.rd<uint8_t>(ft) : 0;
        const size_t fd = b.field(root, 2);
        dim = fd ? b.rd<uint32_t>(fd) : 0;
        // createCompressionStrategy, src/compression_strategy.cpp:25-42
        if (storage < 1 || storage > 3) {
            throw std::runtime_error("Storage strategy " + std::to_string(storage) + " is not supported");
        }
        storageTable = b.sub(root, 1);
        if (storage == 3) {
            // TrainedCompressedStorage ctor, src/trained_compression.cpp:103-111
            wordOffsets = b.sub(storageTable, 0);
            valueOffsets = b.sub(storageTable, 1);
            packedWords = b.sub(storageTable, 2);
            packedValues = b.sub(storageTable, 3);
            packedSize = b.vecLen(packedValues);
            rows = b.vecLen(wordOffsets);
            const size_t dec = b.sub(storageTable, 4);
            const size_t km = b.sub(storageTable, 5);
            // HuffmanDecoder::load, src/huffman_decoder.cpp:25-30
            const size_t k = b.sub(dec, 0), so = b.sub(dec, 1);
            hkeys.assign(b.p + b.vecData(k), b.p + b.vecData(k) + b.vecLen(k));
            hsizes.resize(b.vecLen(so));
            std::memcpy(hsizes.data(), b.p + b.vecData(so), 4 * hsizes.size());
            // DEFAULT_DECODE_TABLE_BIT_LENGTH = 10, src/trained_compression.h:11
            setTableBits(10);
            // KMeansClusterizer::load / centroids, src/kmeans.cpp:121-128,81-84
            const size_t c = b.sub(km, 0);
            centroids.resize(b.vecLen(c));
            std::memcpy(centroids.data(), b.p + b.vecData(c), 4 * centroids.size());
        } else {
            nodes = b.sub(storageTable, 0);
            rows = b.vecLen(nodes);
            if (storage == 2) {
                const size_t fl = b.field(storageTable, 1);
                levels = fl ? b.rd<uint8_t>(fl) : 0;
            }
        }
    }

    const char* trainedWord(uint64_t r) const
    {
        return reinterpret_cast<const char*>(b.p + b.vecData(packedWords) +
                                             b.rd<uint32_t>(b.vecData(wordOffsets) + 4 * r));
    }
    size_t node(uint64_t r) const { return b.follow(b.vecData(nodes) + 4 * r); }
    const char* nodeWord(uint64_t r) const
    {
        return reinterpret_cast<const char*>(b.p + b.vecData(b.sub(node(r), 0)));
    }
    const char* key(uint64_t r) const { return storage == 3 ? trainedWord(r) : nodeWord(r); }

    // Row id of `word`: std::lower_bound with strcmp then an equality check
    // (src/trained_compression.cpp:115-125) / flatbuffers LookupByKey over the
    // sorted nodes (src/uniform_compression.cpp:56, src/full_compression.cpp:39).
    // The reference dereferences end() when the word sorts after every key
    // (UB); the evident intent — "not found" — is what is returned.
    int64_t resolve(const char* word) const
    {
        uint64_t lo = 0, hi = rows;
        while (lo < hi) {
            const uint64_t mid = lo + (hi - lo) / 2;
            if (std::strcmp(key(mid), word) < 0) {
                lo = mid + 1;
            } else {
                hi = mid;
            }
        }
        if (lo < rows && std::strcmp(key(lo), word) == 0) {
            return static_cast<int64_t>(lo);
        }
        return -1;
    }

    void decodeSymbols(uint64_t row, uint8_t* out) const
    {
        const uint32_t offset = b.rd<uint32_t>(b.vecData(valueOffsets) + 4 * row);
        const uint8_t* data = b.p + b.vecData(packedValues);
#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
        auto state = refTable->decode(data + offset, packedSize - offset);
        for (uint32_t i = 0; i < dim; ++i) {
            out[i] = refTable->next(state);
        }
#else
        BitReader br(data + offset, data + packedSize);
        for (uint32_t i = 0; i < dim; ++i) {
            out[i] = table.next(br);
        }
#endif
    }

    // The decode half of CompressedStorage::extract for a known row.
    void extractRow(uint64_t row, float* dst) const
    {
        if (storage == 3) {
            // src/trained_compression.cpp:126-136
            const uint32_t offset = b.rd<uint32_t>(b.vecData(valueOffsets) + 4 * row);
            const uint8_t* data = b.p + b.vecData(packedValues);
#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
            auto state = refTable->decode(data + offset, packedSize - offset);
            for (uint32_t i = 0; i < dim; ++i) {
                dst[i] = centroids[refTable->next(state)];
            }
#else
            BitReader br(data + offset, data + packedSize);
            for (uint32_t i = 0; i < dim; ++i) {
                dst[i] = centroids[table.next(br)];
            }
#endif
        } else if (storage == 2) {
            // src/uniform_compression.cpp:58-73; min/max default to 0 when the
            // builder omitted them (uniform_compression.fbs:3-7)
            const size_t vec = b.sub(node(row), 1);
            const size_t fmin = b.field(vec, 0), fmax = b.field(vec, 1);
            const float minValue = fmin ? b.rd<float>(fmin) : 0.0f;
            const float maxValue = fmax ? b.rd<float>(fmax) : 0.0f;
            const size_t values = b.sub(vec, 2);
            const uint32_t count = b.vecLen(values);
            const uint8_t* v = b.p + b.vecData(values);
            const uint8_t quantizationLevels = levels;
            for (uint32_t i = 0; i < count; ++i) {
                const float floatValue = static_cast<float>(v[i]);
                dst[i] = minValue + (maxValue - minValue) * floatValue / quantizationLevels;
            }
        } else {
            // src/full_compression.cpp:41-42
            const size_t values = b.sub(node(row), 1);
            std::memcpy(dst, b.p + b.vecData(values), 4 * size_t(b.vecLen(values)));
        }
    }

    // Reader::wordEmbeddingToBuffer, src/reader.cpp:41-47
    void wordEmbedding(const char* word, float* dst) const
    {
        const int64_t row = resolve(word);
        if (row >= 0) {
            extractRow(static_cast<uint64_t>(row), dst);
        } else {
            std::fill(dst, dst + dim, 0.0f);
        }
    }

    void rowOrZeros(int64_t row, float* dst) const
    {
        if (row >= 0 && static_cast<uint64_t>(row) < rows) {
            extractRow(static_cast<uint64_t>(row), dst);
        } else {
            std::fill(dst, dst + dim, 0.0f);
        }
    }

    // Reader::batchEmbeddingToBuffer, src/reader.cpp:59-86: serial below 1024
    // items or with one thread, else ceil(n/T) items per std::async task.
    template <class Fn>
    void chunked(uint64_t n, Fn fn) const
    {
        const uint64_t THREADED_DECODER_THRESHOLD = 1024;  // src/reader.cpp:9
        if (n < THREADED_DECODER_THRESHOLD || numThreads == 1) {
            fn(0, n);
            return;
        }
        const uint64_t jobSize = (n + numThreads - 1) / numThreads;
        std::vector<std::future<void>> results;
        for (uint64_t start = 0; start < n; start += jobSize) {
            const uint64_t end = std::min(start + jobSize, n);
            results.push_back(std::async([=] { fn(start, end); }));
        }
        for (auto& r : results) {
            r.get();
        }
    }
};

}  // namespace oracle

// ------------------------------------------------------------------ C API ----
static thread_local std::string g_error;

extern "C" {

const char* oracle_last_error() { return g_error.c_str(); }

#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
int oracle_uses_reference_codec() { return 1; }
#else
int oracle_uses_reference_codec() { return 0; }
#endif

void* oracle_open(const char* path, uint64_t num_threads)
{
    try {
        auto* r = new oracle::Reader();
        try {
            r->open(path, num_threads);
        } catch (...) {
            delete r;
            throw;
        }
        return r;
    } catch (const std::exception& e) {
        g_error = e.what();
        return nullptr;
    }
}

void oracle_close(void* h) { delete static_cast<oracle::Reader*>(h); }
uint32_t oracle_dim(void* h) { return static_cast<oracle::Reader*>(h)->dim; }
uint64_t oracle_rows(void* h) { return static_cast<oracle::Reader*>(h)->rows; }
int oracle_storage(void* h) { return static_cast<oracle::Reader*>(h)->storage; }
uint64_t oracle_threads(void* h) { return static_cast<oracle::Reader*>(h)->numThreads; }
const char* oracle_key(void* h, uint64_t row) { return static_cast<oracle::Reader*>(h)->key(row); }
void oracle_set_table_bits(void* h, int bits) { static_cast<oracle::Reader*>(h)->setTableBits(bits); }

void oracle_word_embedding(void* h, const char* word, float* out)
{
    static_cast<oracle::Reader*>(h)->wordEmbedding(word, out);
}

// words: n byte ranges [offsets[i], offsets[i+1]) of blob; like std::string::c_str()
// + strcmp each is cut at its first NUL.
void oracle_batch_embedding(void* h, const char* blob, const uint64_t* offsets, uint64_t n, float* out)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    std::vector<std::string> words(n);
    for (uint64_t i = 0; i < n; ++i) {
        words[i].assign(blob + offsets[i], blob + offsets[i + 1]);
    }
    const uint32_t dim = r->dim;
    r->chunked(n, [&, dim](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            r->wordEmbedding(words[i].c_str(), out + i * dim);
        }
    });
}

void oracle_resolve(void* h, const char* blob, const uint64_t* offsets, uint64_t n, int32_t* ids)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    for (uint64_t i = 0; i < n; ++i) {
        const std::string w(blob + offsets[i], blob + offsets[i + 1]);
        ids[i] = static_cast<int32_t>(r->resolve(w.c_str()));
    }
}

// Same chunking as batch_embedding but starting from row ids (the decode half
// only); negative / out-of-range ids give zero rows.
void oracle_batch_ids(void* h, const int32_t* ids, uint64_t n, float* out)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    const uint32_t dim = r->dim;
    r->chunked(n, [&, dim](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            r->rowOrZeros(ids[i], out + i * dim);
        }
    });
}

void oracle_decode_range(void* h, uint64_t begin, uint64_t end, float* out)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    const uint32_t dim = r->dim;
    r->chunked(end - begin, [&, dim](uint64_t b, uint64_t e) {
        for (uint64_t i = b; i < e; ++i) {
            r->extractRow(begin + i, out + i * dim);
        }
    });
}

// trained only: the `dim` decoded symbols of one row
int oracle_decode_symbols(void* h, uint64_t row, uint8_t* out)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    if (r->storage != 3 || row >= r->rows) {
        return 1;
    }
    r->decodeSymbols(row, out);
    return 0;
}

uint32_t oracle_codebook(void* h, float* out, uint32_t cap)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    const uint32_t n = static_cast<uint32_t>(r->centroids.size());
    std::memcpy(out, r->centroids.data(), 4 * std::min(n, cap));
    return n;
}

// Serialised prefix code of a trained file: keys, size_offsets and the derived
// canonical (code, length) per key.
uint32_t oracle_prefix_code(void* h, uint8_t* keys, uint16_t* codes, uint8_t* lens, uint32_t cap)
{
    const auto* r = static_cast<oracle::Reader*>(h);
    const uint32_t n = static_cast<uint32_t>(r->hkeys.size());
    for (uint32_t i = 0; i < std::min(n, cap); ++i) {
        keys[i] = r->hkeys[i];
        codes[i] = r->table.codes[i].code;
        lens[i] = r->table.codes[i].len;
    }
    return n;
}

// ---- codec-level entry points for the known-answer tests ----
// BitStream::push, src/bit_stream.h:18-34: codes MSB first, zero padded to a byte.
uint64_t oracle_bitstream_pack(const uint16_t* codes, const uint8_t* lens, uint64_t n, uint8_t* out, uint64_t cap)
{
    uint64_t bits = 0;
    for (uint64_t i = 0; i < n; ++i) {
        for (int k = lens[i] - 1; k >= 0; --k) {
            const uint64_t byte = bits >> 3;
            if (byte < cap) {
                if ((bits & 7) == 0) {
                    out[byte] = 0;
                }
                out[byte] = static_cast<uint8_t>(out[byte] | (((codes[i] >> k) & 1) << (7 - (bits & 7))));
            }
            ++bits;
        }
    }
    return (bits + 7) >> 3;
}

// Decode `count` symbols of a stream with an explicit (keys, size_offsets, L).
void oracle_decode_stream(const uint8_t* keys, uint32_t nkeys, const uint32_t* size_offsets, uint32_t nsizes,
                          int table_bits, const uint8_t* data, uint64_t size, uint64_t count, uint8_t* out)
{
    const std::vector<uint8_t> k(keys, keys + nkeys);
    const std::vector<uint32_t> so(size_offsets, size_offsets + nsizes);
#ifdef MEMB_ORACLE_USE_REFERENCE_CODEC
    memb::HuffmanTableDecoder decoder(k, so, static_cast<size_t>(table_bits));
    auto state = decoder.decode(data, size);
    for (uint64_t i = 0; i < count; ++i) {
        out[i] = decoder.next(state);
    }
#else
    oracle::TableDecoder decoder;
    decoder.build(k, so, table_bits);
    oracle::BitReader br(data, data + size);
    for (uint64_t i = 0; i < count; ++i) {
        out[i] = decoder.next(br);
    }
#endif
}

void oracle_canonical_codes(const uint8_t* keys, uint32_t nkeys, const uint32_t* size_offsets, uint32_t nsizes,
                            uint16_t* codes, uint8_t* lens)
{
    const std::vector<uint8_t> k(keys, keys + nkeys);
    const std::vector<uint32_t> so(size_offsets, size_offsets + nsizes);
    const auto c = oracle::canonicalCodes(k, so);
    for (uint32_t i = 0; i < nkeys; ++i) {
        codes[i] = c[i].code;
        lens[i] = c[i].len;
    }
}

}  // extern "C"
