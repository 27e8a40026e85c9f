// Write side of the memb format: the fixture generator for tests/bench and the
// implementation behind `_memb.Builder`.
//
// Follows the decisions of the reference's compressors so that files look like
// the ones memb_converter produces (same k-means, same Huffman lengths, same
// byte-aligned MSB-first rows), but is written for bulk input and uses all host
// threads for the per-row work:
//   Builder facade         src/builder.cpp:20-68
//   trained compressor     src/trained_compression.cpp:25-101
//   1-D k-means            src/kmeans.cpp:26-112
//   Huffman lengths/codes  src/huffman_encoder.cpp:43-117, src/prefix_code.cpp:5-22
//   bit packing            src/bit_stream.h:18-34
//   uniform compressor     src/uniform_compression.cpp:5-48
//   full compressor        src/full_compression.cpp:5-31
#pragma once

#include <cstdint>
#include <string>
#include <unordered_set>
#include <vector>

namespace memb_b200 {

class Builder {
public:
    // storageName in {"full", "uniform", "trained"}; anything else throws
    // runtime_error("Storage strategy %s is not supported")
    // (src/compression_strategy.cpp:44-61).
    Builder(size_t dim, const std::string& storageName, size_t bitsPerWeight);

    // Throws runtime_error on dimension mismatch / duplicate word with the
    // reference's messages (src/builder.cpp:12-16,34-48).
    void addWord(const std::string& word, const float* values, size_t count);

    // Bulk variant: n words given as NUL-free byte ranges [offsets[i], offsets[i+1])
    // of `blob`, vectors as a row-major (n, dim) matrix.
    void addWords(const char* blob, const uint64_t* offsets, size_t n, const float* matrix);

    std::vector<uint8_t> serialize(unsigned numThreads = 0);
    void save(const std::string& filename, unsigned numThreads = 0);

    // Additive: run the bulk passes of the trained compressor (predict, row sizes,
    // bit packing) on CUDA device `device` instead of the host threads; -1 = host.  The file is
    // byte-identical either way.  Other storages ignore it.
    void setDevice(int device) { device_ = device; }

    size_t dim() const { return dim_; }
    size_t size() const { return words_.size(); }

private:
    size_t dim_;
    int storage_;
    size_t bits_;
    std::vector<std::string> words_;
    std::vector<float> values_;  // row-major, insertion order
    std::unordered_set<std::string> seen_;
    int device_ = -1;
};

const std::vector<std::string>& availableCompressionStrategies();

// Exposed for unit tests -------------------------------------------------------
struct KMeans1D {
    std::vector<float> centroids;
    std::vector<float> splits;
    void setCentroids(const std::vector<float>& c);
    void fit(const float* data, size_t n, size_t clusters);
    void update(const float* data, size_t n);
    uint8_t predictOne(float v) const;
};

struct HuffmanCode {
    std::vector<uint8_t> keys;          // ascending code length
    std::vector<uint8_t> lengths;       // per key (same order)
    std::vector<uint32_t> sizeOffsets;  // serialised form
    uint16_t codeOf[256];
    uint8_t lengthOf[256];
};
// counts in first-occurrence order of the symbols (this is what fixes the
// tie-breaks of the reference's unordered_map-driven construction).
HuffmanCode buildHuffmanCode(const std::vector<std::pair<uint8_t, size_t>>& countsInFirstSeenOrder);

// Everything of a trained file that depends on ALL values: the prefix code, where every row (in
// insertion order) starts in packed_values, and packed_values itself
// (src/trained_compression.cpp:54-71).
struct TrainedEncoding {
    HuffmanCode code;
    std::vector<uint64_t> rowStart;  // n + 1 byte offsets
    std::vector<uint8_t> packed;
};
TrainedEncoding encodeTrainedOnHost(const float* values, size_t n, size_t dim, const KMeans1D& km, unsigned threads);
// device_builder.cu (nvcc): the same on a GPU, byte-identical; throws runtime_error on CUDA errors
TrainedEncoding encodeTrainedOnDevice(int device, const float* values, size_t n, size_t dim, const KMeans1D& km);
int deviceEncoderCount();

}  // namespace memb_b200
