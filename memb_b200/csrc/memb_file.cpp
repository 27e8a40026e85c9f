#include "memb_file.h"
#include "fb_format.h"

#include <algorithm>
#include <cerrno>
#include <cstring>
#include <thread>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace memb_b200 {

using fb::FormatError;
using fb::View;

MappedFile::MappedFile(const std::string& path)
{
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) {
        throw std::runtime_error("Failed to open " + path + ": " + std::strerror(errno));
    }
    struct stat st;
    if (::fstat(fd, &st) != 0) {
        const int err = errno;
        ::close(fd);
        throw std::runtime_error("Failed to stat " + path + ": " + std::strerror(err));
    }
    size_ = static_cast<size_t>(st.st_size);
    if (size_ > 0) {
        void* p = ::mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd, 0);
        if (p == MAP_FAILED) {
            const int err = errno;
            ::close(fd);
            throw std::runtime_error("Failed to mmap " + path + ": " + std::strerror(err));
        }
        data_ = static_cast<const uint8_t*>(p);
    }
    ::close(fd);
}

MappedFile::~MappedFile()
{
    if (data_) {
        ::munmap(const_cast<uint8_t*>(data_), size_);
    }
}

// Code length of keys[i] = number of entries s with sizeOffsets[s] <= i
// (src/huffman_table_decoder.h:31-37); canonical assignment walks the keys in
// stored order, left-shifting the running code whenever the length grows and
// incrementing after every symbol (src/prefix_code.cpp:5-22).
void buildPrefixCodeTable(PrefixCodeTable& t)
{
    const size_t n = t.keys.size();
    if (n == 0) {
        throw FormatError("memb file: Huffman decoder has no symbols");
    }
    if (n > 256) {
        throw FormatError("memb file: Huffman decoder has more than 256 symbols");
    }
    if (t.sizeOffsets.empty() || t.sizeOffsets.size() > 17) {
        throw FormatError("memb file: Huffman size_offsets must hold 1..17 entries (codes are <= 16 bits)");
    }
    for (size_t s = 0; s + 1 < t.sizeOffsets.size(); ++s) {
        if (t.sizeOffsets[s] > t.sizeOffsets[s + 1]) {
            throw FormatError("memb file: Huffman size_offsets are not monotone");
        }
    }
    if (t.sizeOffsets.back() != n) {
        throw FormatError("memb file: Huffman size_offsets do not end at the symbol count");
    }
    t.lengths.resize(n);
    t.codes.resize(n);
    size_t current = 0;
    for (size_t i = 0; i < n; ++i) {
        while (i >= t.sizeOffsets[current]) {
            ++current;
        }
        t.lengths[i] = static_cast<uint8_t>(current);
    }
    std::vector<bool> seen(256, false);
    for (uint8_t k : t.keys) {
        if (seen[k]) {
            throw FormatError("memb file: duplicate symbol in Huffman decoder");
        }
        seen[k] = true;
    }
    uint32_t code = 0;
    unsigned bits = 0;
    uint64_t kraft = 0;  // in units of 2^-16
    for (size_t i = 0; i < n; ++i) {
        while (bits < t.lengths[i]) {
            ++bits;
            code <<= 1;
        }
        if (bits > 0 && code >= (1u << bits)) {
            throw FormatError("memb file: Huffman code is over-subscribed");
        }
        t.codes[i] = static_cast<uint16_t>(code);
        ++code;
        kraft += 1ull << (16 - t.lengths[i]);
    }
    t.minLen = t.lengths.front();
    t.maxLen = t.lengths.back();
    if (n == 1) {
        if (t.maxLen != 0) {
            throw FormatError("memb file: single-symbol Huffman code must have length 0");
        }
    } else if (kraft != (1ull << 16)) {
        // The reference's table decoder indexes out of bounds on a window that
        // matches no code (src/huffman_table_decoder.h:106-116); a Huffman tree
        // with >= 2 leaves is always complete, so anything else is corruption.
        throw FormatError("memb file: Huffman code is not complete");
    }
}

namespace {

template <class Fn>
void parallelFor(uint64_t n, unsigned threads, Fn fn)
{
    threads = std::max(1u, threads);
    if (n < 4096 || threads == 1) {
        fn(0, n);
        return;
    }
    const uint64_t chunk = (n + threads - 1) / threads;
    std::vector<std::thread> pool;
    std::vector<std::exception_ptr> errors(threads);
    for (unsigned t = 0; t < threads; ++t) {
        const uint64_t b = std::min<uint64_t>(n, t * chunk), e = std::min<uint64_t>(n, b + chunk);
        if (b >= e) {
            break;
        }
        pool.emplace_back([&, t, b, e] {
            try {
                fn(b, e);
            } catch (...) {
                errors[t] = std::current_exception();
            }
        });
    }
    for (auto& th : pool) {
        th.join();
    }
    for (auto& err : errors) {
        if (err) {
            std::rethrow_exception(err);
        }
    }
}

void checkSorted(const MembFile& f, unsigned threads)
{
    // Row id := index in keys(); lookups in the reference are binary searches
    // with strcmp (src/trained_compression.cpp:115-125, flatbuffers
    // LookupByKey), which only agree with an exact-match hash when the keys
    // are strictly ascending.
    parallelFor(f.rows ? f.rows - 1 : 0, threads, [&](uint64_t b, uint64_t e) {
        for (uint64_t r = b; r < e; ++r) {
            if (std::strcmp(f.words + f.wordOffsets[r], f.words + f.wordOffsets[r + 1]) >= 0) {
                throw FormatError("memb file: words are not strictly ascending (row " +
                                  std::to_string(r) + ")");
            }
        }
    });
}

void parseTrained(const View& v, size_t t, MembFile& f, unsigned threads)
{
    namespace I = fb::ids;
    const auto wordOffsets = v.vector(t, I::TRAINED_WORD_OFFSETS, 4);
    const auto valueOffsets = v.vector(t, I::TRAINED_VALUE_OFFSETS, 4);
    const auto packedWords = v.vector(t, I::TRAINED_PACKED_WORDS, 1);
    const auto packedValues = v.vector(t, I::TRAINED_PACKED_VALUES, 1);
    const size_t decoder = v.table(t, I::TRAINED_DECODER);
    const size_t clusterizer = v.table(t, I::TRAINED_CLUSTERIZER);
    if (!wordOffsets.present || !valueOffsets.present || !packedWords.present ||
        !packedValues.present || !decoder || !clusterizer) {
        throw FormatError("memb file: trained storage is missing a required field");
    }
    if (wordOffsets.len != valueOffsets.len) {
        throw FormatError("memb file: word_offsets and value_offsets differ in length");
    }
    f.rows = wordOffsets.len;
    f.words = reinterpret_cast<const char*>(v.data() + packedWords.pos);
    // flatbuffers strings carry a trailing NUL that is not counted in len; it
    // terminates the last word even if packed_words itself lacks one.
    f.wordsBytes = static_cast<uint64_t>(packedWords.len) + 1;
    if (packedWords.pos + f.wordsBytes > v.size() || f.words[packedWords.len] != '\0') {
        throw FormatError("memb file: packed_words is not NUL-terminated");
    }
    f.wordOffsets.resize(f.rows + 1);
    if (f.rows) {
        std::memcpy(f.wordOffsets.data(), v.data() + wordOffsets.pos, 4 * f.rows);
    }
    f.wordOffsets[f.rows] = static_cast<uint32_t>(f.wordsBytes);
    for (uint64_t r = 0; r < f.rows; ++r) {
        if (f.wordOffsets[r] > packedWords.len) {
            throw FormatError("memb file: word offset outside packed_words");
        }
    }
    f.packedValues = v.data() + packedValues.pos;
    f.packedSize = packedValues.len;
    f.valueOffsets.resize(f.rows);
    if (f.rows) {
        std::memcpy(f.valueOffsets.data(), v.data() + valueOffsets.pos, 4 * f.rows);
    }
    for (uint64_t r = 0; r < f.rows; ++r) {
        if (f.valueOffsets[r] > f.packedSize) {
            throw FormatError("memb file: value offset outside packed_values");
        }
    }

    const auto keys = v.vector(decoder, I::HUFFMAN_KEYS, 1);
    const auto sizeOffsets = v.vector(decoder, I::HUFFMAN_SIZE_OFFSETS, 4);
    if (!keys.present || !sizeOffsets.present) {
        throw FormatError("memb file: Huffman decoder is missing a required field");
    }
    f.code.keys.assign(v.data() + keys.pos, v.data() + keys.pos + keys.len);
    f.code.sizeOffsets.resize(sizeOffsets.len);
    if (sizeOffsets.len) {
        std::memcpy(f.code.sizeOffsets.data(), v.data() + sizeOffsets.pos, 4 * sizeOffsets.len);
    }
    buildPrefixCodeTable(f.code);
    // The load decodes every row once (dim symbols each) to measure it: bound that work by what
    // the file can hold.  A row of dim codes of at least minLen bits each that starts inside
    // packed_values cannot be longer than the vector plus the zero bits the reference reads
    // past its end (src/bit_stream_reader.h:22-25) -- allow 64 KiB of those; a one-symbol code
    // (zero-bit rows) is bounded by a plain cap.
    const uint64_t rowBits = static_cast<uint64_t>(f.dim) * std::max(1u, f.code.minLen);
    if (f.dim > (1u << 24) || rowBits > 8 * (static_cast<uint64_t>(f.packedSize) + 65536)) {
        throw FormatError("memb file: dim is larger than packed_values can hold rows for");
    }

    const auto centroids = v.vector(clusterizer, I::KMEANS_CENTROIDS, 4);
    if (!centroids.present || centroids.len == 0) {
        throw FormatError("memb file: clusterizer has no centroids");
    }
    f.centroids.resize(centroids.len);
    std::memcpy(f.centroids.data(), v.data() + centroids.pos, 4 * centroids.len);
    for (uint8_t k : f.code.keys) {
        // Symbols index centroids_ directly (src/trained_compression.cpp:133-135).
        if (k >= f.centroids.size()) {
            throw FormatError("memb file: Huffman symbol exceeds the codebook size");
        }
    }
    checkSorted(f, threads);
}

// Uniform and full storages keep one flatbuffers table per word
// (uniform_compression.fbs:3-17, full_compression.fbs:3-10); they are
// flattened here into SoA arrays so the device sees dense rows.
template <bool UNIFORM>
void parseNodeStorage(const View& v, size_t t, MembFile& f, unsigned threads)
{
    namespace I = fb::ids;
    const auto nodes = v.vector(t, UNIFORM ? I::UNIFORM_NODES : I::FULL_NODES, 4);
    if (!nodes.present) {
        throw FormatError("memb file: storage has no nodes vector");
    }
    f.rows = nodes.len;
    const uint64_t dim = f.dim;
    // every value of every row lives in the mapping: a header that promises more than the file
    // holds is corrupt (and would otherwise drive a terabyte-sized allocation below)
    if (dim != 0 && f.rows > v.size() / (dim * (UNIFORM ? 1 : 4))) {
        throw FormatError("memb file: rows x dim exceeds the file size");
    }
    std::vector<View::Span> wordSpans(f.rows), valueSpans(f.rows);
    if (UNIFORM) {
        f.quantizationLevels = v.scalar<uint8_t>(t, I::UNIFORM_LEVELS, 0);
        f.uniformMin.resize(f.rows);
        f.uniformMax.resize(f.rows);
        f.uniformValues.resize(f.rows * dim);
    } else {
        f.fullValues.resize(f.rows * dim);
    }
    parallelFor(f.rows, threads, [&](uint64_t b, uint64_t e) {
        for (uint64_t r = b; r < e; ++r) {
            const size_t node = v.tableAt(nodes, static_cast<uint32_t>(r));
            wordSpans[r] = v.vector(node, UNIFORM ? I::UNODE_WORD : I::FNODE_WORD, 1);
            if (!wordSpans[r].present) {
                throw FormatError("memb file: node without a word");
            }
            if (UNIFORM) {
                const size_t vec = v.table(node, I::UNODE_VALUES);
                if (!vec) {
                    throw FormatError("memb file: uniform node without compressed_values");
                }
                f.uniformMin[r] = v.scalar<float>(vec, I::UVEC_MIN, 0.0f);
                f.uniformMax[r] = v.scalar<float>(vec, I::UVEC_MAX, 0.0f);
                valueSpans[r] = v.vector(vec, I::UVEC_VALUES, 1);
            } else {
                valueSpans[r] = v.vector(node, I::FNODE_VALUES, 4);
            }
            // The reference writes values->size() floats into a dim-sized row
            // (src/uniform_compression.cpp:64-73, src/full_compression.cpp:41-42);
            // any other length is a buffer overrun or an uninitialised tail there.
            if (!valueSpans[r].present || valueSpans[r].len != dim) {
                throw FormatError("memb file: row " + std::to_string(r) +
                                  " does not hold exactly dim values");
            }
            if (UNIFORM) {
                std::memcpy(f.uniformValues.data() + r * dim, v.data() + valueSpans[r].pos, dim);
            } else {
                std::memcpy(f.fullValues.data() + r * dim, v.data() + valueSpans[r].pos, 4 * dim);
            }
        }
    });
    f.wordOffsets.resize(f.rows + 1);
    uint64_t total = 0;
    for (uint64_t r = 0; r < f.rows; ++r) {
        f.wordOffsets[r] = static_cast<uint32_t>(total);
        // keys are compared with strcmp => cut at the first NUL like c_str()
        const char* w = reinterpret_cast<const char*>(v.data() + wordSpans[r].pos);
        total += ::strnlen(w, wordSpans[r].len) + 1;
        if (total > 0xffffffffULL) {
            throw FormatError("memb file: vocabulary text exceeds 4 GiB");
        }
    }
    f.wordOffsets[f.rows] = static_cast<uint32_t>(total);
    f.wordsOwned.assign(total, '\0');
    parallelFor(f.rows, threads, [&](uint64_t b, uint64_t e) {
        for (uint64_t r = b; r < e; ++r) {
            const size_t len = f.wordOffsets[r + 1] - f.wordOffsets[r] - 1;
            std::memcpy(f.wordsOwned.data() + f.wordOffsets[r], v.data() + wordSpans[r].pos, len);
        }
    });
    f.words = f.wordsOwned.data();
    f.wordsBytes = total;
    checkSorted(f, threads);
}

}  // namespace

std::unique_ptr<MembFile> parseMembFile(const std::string& path, unsigned numThreads)
{
    auto f = std::make_unique<MembFile>();
    f->map = std::make_unique<MappedFile>(path);
    const View v(f->map->data(), f->map->size());
    // src/reader.cpp:104-111: size >= 8 and identifier "memb" at bytes [4, 8).
    if (v.size() < 8 || std::memcmp(v.data() + 4, "memb", 4) != 0) {
        throw std::runtime_error("File format verification failed");
    }
    namespace I = fb::ids;
    const size_t root = v.root();
    const uint8_t storageType = v.scalar<uint8_t>(root, I::INDEX_STORAGE_TYPE, I::STORAGE_NONE);
    f->dim = v.scalar<uint32_t>(root, I::INDEX_DIM, 0);
    if (storageType != I::STORAGE_FULL && storageType != I::STORAGE_UNIFORM &&
        storageType != I::STORAGE_TRAINED) {
        // src/compression_strategy.cpp:36-38
        throw std::runtime_error("Storage strategy " + std::to_string(storageType) +
                                 " is not supported");
    }
    const size_t storage = v.table(root, I::INDEX_STORAGE);
    if (!storage) {
        throw FormatError("memb file: index has no storage table");
    }
    if (f->dim == 0) {
        throw FormatError("memb file: dim is 0");
    }
    f->storage = storageType;
    switch (storageType) {
    case I::STORAGE_TRAINED:
        parseTrained(v, storage, *f, numThreads);
        break;
    case I::STORAGE_UNIFORM:
        parseNodeStorage<true>(v, storage, *f, numThreads);
        break;
    default:
        parseNodeStorage<false>(v, storage, *f, numThreads);
        break;
    }
    if (f->rows >= 0x7fffffffULL) {
        throw FormatError("memb file: too many rows for 31-bit row ids");
    }
    return f;
}

}  // namespace memb_b200
