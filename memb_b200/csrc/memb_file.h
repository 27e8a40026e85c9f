// Host-side view of one memb file after parsing: everything the device loader
// needs, in flat arrays, with every offset validated.
//
// Replaces the read side of the reference's storage strategies:
//   Reader::Reader / getIndexChecked            src/reader.cpp:13-29,104-111
//   createCompressionStrategy(wire::Storage)    src/compression_strategy.cpp:25-42
//   TrainedCompressedStorage ctor               src/trained_compression.cpp:103-111
//   HuffmanDecoder::load / KMeansClusterizer::load  src/huffman_decoder.cpp:25-30, src/kmeans.cpp:121-128
//   Uniform/Full storage ctors                  src/uniform_compression.cpp:50-52, src/full_compression.cpp:33-35
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace memb_b200 {

enum StorageKind : int { STORAGE_FULL = 1, STORAGE_UNIFORM = 2, STORAGE_TRAINED = 3 };

// Read-only mmap of the file (the reference uses boost mapped_file_source,
// src/reader.h:37).
class MappedFile {
public:
    explicit MappedFile(const std::string& path);
    ~MappedFile();
    MappedFile(const MappedFile&) = delete;
    MappedFile& operator=(const MappedFile&) = delete;
    const uint8_t* data() const { return data_; }
    size_t size() const { return size_; }

private:
    const uint8_t* data_ = nullptr;
    size_t size_ = 0;
};

// Canonical prefix code in the serialised form of the file
// (src/flatbuffers/huffman_decoder.fbs:3-6) plus the derived per-symbol codes
// (src/prefix_code.cpp:5-22, src/huffman_table_decoder.h:31-39).
struct PrefixCodeTable {
    std::vector<uint8_t> keys;           // symbols, ascending code length
    std::vector<uint32_t> sizeOffsets;   // as stored
    std::vector<uint8_t> lengths;        // lengths[i] = code length of keys[i]
    std::vector<uint16_t> codes;         // codes[i]   = canonical code of keys[i] (low `length` bits)
    unsigned minLen = 0, maxLen = 0;
};

struct MembFile {
    std::unique_ptr<MappedFile> map;
    int storage = 0;
    uint32_t dim = 0;
    uint64_t rows = 0;

    // Vocabulary in row-id order (= sorted order): NUL-terminated
    // words back to back, wordOffsets[r] = start of word r, wordOffsets[rows] =
    // total bytes.  For trained files this is the file's own packed_words.
    const char* words = nullptr;
    uint64_t wordsBytes = 0;
    std::vector<char> wordsOwned;           // uniform/full: rebuilt here
    std::vector<uint32_t> wordOffsets;      // rows + 1 entries

    // trained
    const uint8_t* packedValues = nullptr;  // into the mmap
    uint64_t packedSize = 0;
    std::vector<uint32_t> valueOffsets;     // rows entries, byte offsets into packedValues
    PrefixCodeTable code;
    std::vector<float> centroids;

    // uniform (flattened AoS -> SoA)
    std::vector<uint8_t> uniformValues;     // rows * dim
    std::vector<float> uniformMin, uniformMax;
    uint8_t quantizationLevels = 0;

    // full
    std::vector<float> fullValues;          // rows * dim

    std::string word(uint64_t row) const { return std::string(words + wordOffsets[row]); }
};

// Derives lengths/codes from (keys, sizeOffsets); throws FormatError when the
// table cannot have been produced by a Huffman construction (length > 16,
// non-monotone offsets, over-subscribed or incomplete code).
void buildPrefixCodeTable(PrefixCodeTable& table);

// Throws std::runtime_error("File format verification failed") for a bad
// identifier exactly like src/reader.cpp:104-111, std::runtime_error for an
// unsupported storage type like src/compression_strategy.cpp:36-38, and
// fb::FormatError (a runtime_error) for any structural problem.
std::unique_ptr<MembFile> parseMembFile(const std::string& path, unsigned numThreads);

}  // namespace memb_b200
