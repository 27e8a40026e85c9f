#include "builder.h"
#include "fb_format.h"
#include "memb_file.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <queue>
#include <stdexcept>
#include <thread>
#include <unordered_map>

namespace memb_b200 {

namespace {

const size_t CLUSTER_SAMPLE_ROWS = 10000;  // src/trained_compression.cpp:21
const size_t KMEANS_EPOCHS = 30;           // src/kmeans.cpp:11
const double SMALL_CLUSTER_FACTOR = 128;   // src/kmeans.cpp:12

unsigned hostThreads(unsigned requested)
{
    if (requested) {
        return requested;
    }
    return std::max(std::thread::hardware_concurrency(), 2u);
}

template <class Fn>
void parallelRows(size_t n, unsigned threads, Fn fn)
{
    if (n < 2048 || threads <= 1) {
        fn(0, size_t(0), n);
        return;
    }
    const size_t chunk = (n + threads - 1) / threads;
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < threads; ++t) {
        const size_t b = std::min(n, t * chunk), e = std::min(n, b + chunk);
        if (b >= e) {
            break;
        }
        pool.emplace_back([=, &fn] { fn(t, b, e); });
    }
    for (auto& th : pool) {
        th.join();
    }
}

int storageFromName(const std::string& name)
{
    if (name == "full") return STORAGE_FULL;
    if (name == "uniform") return STORAGE_UNIFORM;
    if (name == "trained") return STORAGE_TRAINED;
    throw std::runtime_error("Storage strategy " + name + " is not supported");
}

}  // namespace

const std::vector<std::string>& availableCompressionStrategies()
{
    // order of src/compression_strategy.cpp:14-18
    static const std::vector<std::string> names = {"full", "uniform", "trained"};
    return names;
}

// ------------------------------------------------------------- k-means ----
void KMeans1D::setCentroids(const std::vector<float>& c)
{
    centroids = c;
    splits.clear();
    for (size_t i = 0; i + 1 < centroids.size(); ++i) {
        // midpoint evaluated in double, stored as float (src/kmeans.cpp:108-110)
        splits.push_back(static_cast<float>(0.5 * (centroids[i] + centroids[i + 1])));
    }
}

uint8_t KMeans1D::predictOne(float v) const
{
    // number of split points strictly below v (lower_bound, src/kmeans.cpp:73-76)
    return static_cast<uint8_t>(std::lower_bound(splits.begin(), splits.end(), v) - splits.begin());
}

// One Lloyd step with the reference's *running-mean* update; the running mean is
// order dependent in fp32, so this stays a sequential pass (src/kmeans.cpp:86-102).
void KMeans1D::update(const float* data, size_t n)
{
    const size_t k = centroids.size();
    std::vector<size_t> counts(k, 0);
    std::vector<float> means(k, 0.0f);
    for (size_t i = 0; i < n; ++i) {
        const uint8_t a = predictOne(data[i]);
        const float c = static_cast<float>(counts[a]);
        means[a] = c / (c + 1) * means[a] + 1 / (c + 1) * data[i];
        counts[a] += 1;
    }
    std::sort(means.begin(), means.end());
    setCentroids(means);
}

void KMeans1D::fit(const float* data, size_t n, size_t clusters)
{
    if (n == 0) {
        throw std::runtime_error("Attempt to use KMeansClusterizer before fitting");
    }
    const auto mm = std::minmax_element(data, data + n);
    const float lo = *mm.first, hi = *mm.second;
    std::vector<float> init;
    for (size_t i = 0; i < clusters; ++i) {
        init.push_back(lo + i / static_cast<float>(clusters - 1) * (hi - lo));
    }
    setCentroids(init);
    for (size_t epoch = 0; epoch < KMEANS_EPOCHS; ++epoch) {
        update(data, n);
    }
    // prune clusters holding <= max/128 of the sample (src/kmeans.cpp:44-60)
    std::vector<size_t> counts(centroids.size(), 0);
    for (size_t i = 0; i < n; ++i) {
        counts[predictOne(data[i])] += 1;
    }
    const size_t maxCount = *std::max_element(counts.begin(), counts.end());
    const double limit = maxCount / SMALL_CLUSTER_FACTOR;
    std::vector<float> kept;
    for (size_t i = 0; i < centroids.size(); ++i) {
        if (static_cast<double>(counts[i]) > limit) {
            kept.push_back(centroids[i]);
        }
    }
    setCentroids(kept);
    update(data, n);
}

// ------------------------------------------------------------- Huffman ----
namespace {

struct HeapNode {
    size_t count;
    int index;  // into the node pool
};
struct HeapOrder {
    bool operator()(const HeapNode& a, const HeapNode& b) const { return a.count > b.count; }
};
struct PoolNode {
    uint8_t key;
    int left, right;
};

void collectLengths(const std::vector<PoolNode>& pool, int node, uint8_t depth,
                    std::vector<std::pair<uint8_t, uint8_t>>& out)
{
    if (pool[node].left >= 0 && pool[node].right >= 0) {
        collectLengths(pool, pool[node].left, depth + 1, out);
        collectLengths(pool, pool[node].right, depth + 1, out);
    } else {
        out.push_back({pool[node].key, depth});
    }
}

}  // namespace

HuffmanCode buildHuffmanCode(const std::vector<std::pair<uint8_t, size_t>>& firstSeen)
{
    if (firstSeen.empty()) {
        throw std::runtime_error("Cannot build a Huffman code without symbols");
    }
    // The reference walks an unordered_map<uint8_t,size_t> filled in data order
    // (src/huffman_encoder.cpp:52-55,120-124); using the same container with the
    // same insertion order reproduces its iteration order and therefore its
    // tie-breaks between equal counts.
    std::unordered_map<uint8_t, size_t> counts;
    for (const auto& kv : firstSeen) {
        counts[kv.first] += kv.second;
    }
    std::vector<PoolNode> pool;
    std::priority_queue<HeapNode, std::vector<HeapNode>, HeapOrder> heap;
    for (const auto& kv : counts) {
        pool.push_back({kv.first, -1, -1});
        heap.push({kv.second, static_cast<int>(pool.size()) - 1});
    }
    while (heap.size() > 1) {
        const HeapNode a = heap.top();
        heap.pop();
        const HeapNode b = heap.top();
        heap.pop();
        pool.push_back({0, a.index, b.index});
        heap.push({a.count + b.count, static_cast<int>(pool.size()) - 1});
    }
    std::vector<std::pair<uint8_t, uint8_t>> lengths;  // (key, length) in DFS order
    collectLengths(pool, heap.top().index, 0, lengths);
    // unstable sort by length only, like src/huffman_encoder.cpp:84-90
    std::sort(lengths.begin(), lengths.end(),
              [](const std::pair<uint8_t, uint8_t>& a, const std::pair<uint8_t, uint8_t>& b) {
                  return a.second < b.second;
              });
    if (lengths.back().second > 16) {
        throw std::runtime_error("Huffman code length exceeds 16 bits");
    }
    HuffmanCode hc;
    std::memset(hc.codeOf, 0, sizeof(hc.codeOf));
    std::memset(hc.lengthOf, 0, sizeof(hc.lengthOf));
    uint16_t code = 0;
    unsigned bits = 0;
    unsigned currentSize = 0;
    for (size_t i = 0; i < lengths.size(); ++i) {
        while (bits < lengths[i].second) {
            ++bits;
            code = static_cast<uint16_t>(code << 1);
        }
        while (currentSize < lengths[i].second) {
            ++currentSize;
            hc.sizeOffsets.push_back(static_cast<uint32_t>(i));
        }
        hc.keys.push_back(lengths[i].first);
        hc.lengths.push_back(lengths[i].second);
        hc.codeOf[lengths[i].first] = code;
        hc.lengthOf[lengths[i].first] = lengths[i].second;
        ++code;
    }
    hc.sizeOffsets.push_back(static_cast<uint32_t>(lengths.size()));
    return hc;
}

// ------------------------------------------------------------- Builder ----
Builder::Builder(size_t dim, const std::string& storageName, size_t bitsPerWeight)
    : dim_(dim), storage_(storageFromName(storageName)), bits_(bitsPerWeight)
{}

void Builder::addWord(const std::string& word, const float* values, size_t count)
{
    if (count != dim_) {
        throw std::runtime_error("Vector dimension (" + std::to_string(count) + ") for word " + word +
                                 " doesn't match builder dimension (" + std::to_string(dim_) + ")");
    }
    if (!seen_.insert(word).second) {
        throw std::runtime_error("Attempt to add duplicate word " + word + " to index");
    }
    words_.push_back(word);
    values_.insert(values_.end(), values, values + count);
}

void Builder::addWords(const char* blob, const uint64_t* offsets, size_t n, const float* matrix)
{
    words_.reserve(words_.size() + n);
    values_.reserve(values_.size() + n * dim_);
    for (size_t i = 0; i < n; ++i) {
        addWord(std::string(blob + offsets[i], blob + offsets[i + 1]), matrix + i * dim_, dim_);
    }
}

namespace {

// Rows sorted the way the reference sorts them: std::string operator< for
// trained (src/trained_compression.cpp:73-79), strcmp-keyed sorted tables for
// uniform/full (CreateVectorOfSortedTables, src/uniform_compression.cpp:45).
// Both are unsigned byte-wise order for NUL-free words.
std::vector<uint32_t> sortedOrder(const std::vector<std::string>& words)
{
    std::vector<uint32_t> order(words.size());
    for (size_t i = 0; i < order.size(); ++i) {
        order[i] = static_cast<uint32_t>(i);
    }
    std::sort(order.begin(), order.end(),
              [&](uint32_t a, uint32_t b) { return words[a] < words[b]; });
    return order;
}

struct IndexHeader {
    fb::Writer::TableLayout layout;
    size_t table;
};

// [root uoffset]["memb"][Index vtable][Index table]; the storage table follows.
IndexHeader writeIndex(fb::Writer& w, uint8_t storageType, uint32_t dim)
{
    namespace I = fb::ids;
    const size_t rootPos = w.put<uint32_t>(0);
    w.put<uint8_t>('m');
    w.put<uint8_t>('e');
    w.put<uint8_t>('m');
    w.put<uint8_t>('b');
    IndexHeader h;
    h.layout = w.vtable({{I::INDEX_STORAGE, 4}, {I::INDEX_DIM, 4}, {I::INDEX_STORAGE_TYPE, 1}});
    h.table = w.table(h.layout);
    w.setScalar<uint32_t>(h.layout, h.table, I::INDEX_DIM, dim);
    w.setScalar<uint8_t>(h.layout, h.table, I::INDEX_STORAGE_TYPE, storageType);
    w.patch<uint32_t>(rootPos, static_cast<uint32_t>(h.table - rootPos));
    return h;
}

}  // namespace

// The bulk passes of the trained compressor on the host threads (src/trained_compression.cpp:54-71);
// device_builder.cu holds the GPU version, which must stay byte-identical to this one.
TrainedEncoding encodeTrainedOnHost(const float* values, size_t n, size_t dim, const KMeans1D& km, unsigned threads)
{
    // quantise every row; remember the order in which symbols first appear
    std::vector<uint8_t> symbols(n * dim);
    std::vector<std::vector<size_t>> partCounts(threads, std::vector<size_t>(256, 0));
    std::vector<std::vector<uint64_t>> partFirst(threads, std::vector<uint64_t>(256, ~0ull));
    parallelRows(n, threads, [&](unsigned t, size_t b, size_t e) {
        auto& counts = partCounts[t];
        auto& first = partFirst[t];
        for (size_t i = b * dim; i < e * dim; ++i) {
            const uint8_t s = km.predictOne(values[i]);
            symbols[i] = s;
            if (counts[s]++ == 0) {
                first[s] = i;
            }
        }
    });
    std::vector<size_t> counts(256, 0);
    std::vector<uint64_t> first(256, ~0ull);
    for (unsigned t = 0; t < threads; ++t) {
        for (int s = 0; s < 256; ++s) {
            counts[s] += partCounts[t][s];
            first[s] = std::min(first[s], partFirst[t][s]);
        }
    }
    std::vector<std::pair<uint64_t, uint8_t>> seenOrder;
    for (int s = 0; s < 256; ++s) {
        if (counts[s]) {
            seenOrder.push_back({first[s], static_cast<uint8_t>(s)});
        }
    }
    std::sort(seenOrder.begin(), seenOrder.end());
    std::vector<std::pair<uint8_t, size_t>> firstSeen;
    for (const auto& p : seenOrder) {
        firstSeen.push_back({p.second, counts[p.second]});
    }
    if (firstSeen.empty()) {
        throw std::runtime_error("Cannot save an empty trained index");
    }
    TrainedEncoding enc;
    enc.code = buildHuffmanCode(firstSeen);
    const HuffmanCode& hc = enc.code;

    // every row is its own byte-aligned, zero-padded, MSB-first bit stream,
    // appended in insertion order (src/trained_compression.cpp:65-71)
    std::vector<uint64_t>& rowStart = enc.rowStart;
    rowStart.assign(n + 1, 0);
    parallelRows(n, threads, [&](unsigned, size_t b, size_t e) {
        for (size_t r = b; r < e; ++r) {
            uint64_t bitsTotal = 0;
            for (size_t i = 0; i < dim; ++i) {
                bitsTotal += hc.lengthOf[symbols[r * dim + i]];
            }
            rowStart[r + 1] = (bitsTotal + 7) / 8;
        }
    });
    for (size_t r = 0; r < n; ++r) {
        rowStart[r + 1] += rowStart[r];
    }
    if (rowStart[n] > 0xffffffffULL) {
        throw std::runtime_error("packed_values exceeds the 32-bit offsets of the format");
    }
    std::vector<uint8_t>& packed = enc.packed;
    packed.assign(rowStart[n], 0);
    parallelRows(n, threads, [&](unsigned, size_t b, size_t e) {
        for (size_t r = b; r < e; ++r) {
            uint8_t* dst = packed.data() + rowStart[r];
            uint64_t acc = 0;
            unsigned have = 0;
            for (size_t i = 0; i < dim; ++i) {
                const uint8_t s = symbols[r * dim + i];
                acc = (acc << hc.lengthOf[s]) | hc.codeOf[s];
                have += hc.lengthOf[s];
                while (have >= 8) {
                    *dst++ = static_cast<uint8_t>(acc >> (have - 8));
                    have -= 8;
                }
            }
            if (have) {
                *dst++ = static_cast<uint8_t>(acc << (8 - have));
            }
        }
    });
    return enc;
}

std::vector<uint8_t> Builder::serialize(unsigned numThreads)
{
    namespace I = fb::ids;
    const unsigned threads = hostThreads(numThreads);
    const size_t n = words_.size();
    const size_t dim = dim_;
    fb::Writer w;
    const std::vector<uint32_t> order = sortedOrder(words_);

    if (storage_ == STORAGE_TRAINED) {
        // levels = min(1 << bits, 255) as uint8 (src/trained_compression.cpp:25-30)
        const size_t levels = static_cast<uint8_t>(std::min<long long>(1LL << std::min<size_t>(bits_, 30), 255));
        KMeans1D km;
        const size_t sampleRows = std::min(CLUSTER_SAMPLE_ROWS, n);
        km.fit(values_.data(), sampleRows * dim, levels);

        const TrainedEncoding enc = device_ >= 0 ? encodeTrainedOnDevice(device_, values_.data(), n, dim, km)
                                                  : encodeTrainedOnHost(values_.data(), n, dim, km, threads);
        const HuffmanCode& hc = enc.code;
        const std::vector<uint64_t>& rowStart = enc.rowStart;
        const std::vector<uint8_t>& packed = enc.packed;

        std::string packedWords;
        std::vector<uint32_t> wordOffsets(n), valueOffsets(n);
        for (size_t i = 0; i < n; ++i) {
            const uint32_t src = order[i];
            wordOffsets[i] = static_cast<uint32_t>(packedWords.size());
            valueOffsets[i] = static_cast<uint32_t>(rowStart[src]);
            packedWords.append(words_[src].c_str(), std::strlen(words_[src].c_str()) + 1);
        }

        const IndexHeader idx = writeIndex(w, I::STORAGE_TRAINED, static_cast<uint32_t>(dim));
        const auto lay = w.vtable({{I::TRAINED_WORD_OFFSETS, 4}, {I::TRAINED_VALUE_OFFSETS, 4},
                                   {I::TRAINED_PACKED_WORDS, 4}, {I::TRAINED_PACKED_VALUES, 4},
                                   {I::TRAINED_DECODER, 4}, {I::TRAINED_CLUSTERIZER, 4}});
        const size_t t = w.table(lay);
        w.link(w.fieldPos(idx.layout, idx.table, I::INDEX_STORAGE), t);
        w.link(w.fieldPos(lay, t, I::TRAINED_WORD_OFFSETS), w.vector(wordOffsets.data(), n, 4));
        w.link(w.fieldPos(lay, t, I::TRAINED_VALUE_OFFSETS), w.vector(valueOffsets.data(), n, 4));
        w.link(w.fieldPos(lay, t, I::TRAINED_PACKED_WORDS),
               w.string(packedWords.data(), packedWords.size()));
        w.link(w.fieldPos(lay, t, I::TRAINED_PACKED_VALUES), w.vector(packed.data(), packed.size(), 1));
        const auto decLay = w.vtable({{I::HUFFMAN_KEYS, 4}, {I::HUFFMAN_SIZE_OFFSETS, 4}});
        const size_t dec = w.table(decLay);
        w.link(w.fieldPos(lay, t, I::TRAINED_DECODER), dec);
        w.link(w.fieldPos(decLay, dec, I::HUFFMAN_KEYS), w.vector(hc.keys.data(), hc.keys.size(), 1));
        w.link(w.fieldPos(decLay, dec, I::HUFFMAN_SIZE_OFFSETS),
               w.vector(hc.sizeOffsets.data(), hc.sizeOffsets.size(), 4));
        const auto kmLay = w.vtable({{I::KMEANS_CENTROIDS, 4}});
        const size_t kmTable = w.table(kmLay);
        w.link(w.fieldPos(lay, t, I::TRAINED_CLUSTERIZER), kmTable);
        w.link(w.fieldPos(kmLay, kmTable, I::KMEANS_CENTROIDS),
               w.vector(km.centroids.data(), km.centroids.size(), 4));
        return std::move(w.buffer());
    }

    if (storage_ == STORAGE_UNIFORM) {
        // levels = min(1 << bits, 255) as uint8 (src/uniform_compression.cpp:5-8)
        const uint8_t levels = static_cast<uint8_t>(std::min<long long>(1LL << std::min<size_t>(bits_, 30), 255));
        std::vector<uint8_t> q(n * dim);
        std::vector<float> mins(n), maxs(n);
        parallelRows(n, threads, [&](unsigned, size_t b, size_t e) {
            for (size_t r = b; r < e; ++r) {
                const float* src = values_.data() + r * dim;
                const auto mm = std::minmax_element(src, src + dim);
                const float lo = *mm.first, hi = *mm.second;
                mins[r] = lo;
                maxs[r] = hi;
                for (size_t i = 0; i < dim; ++i) {
                    // levels * (v - min) / (max - min), truncated (src/uniform_compression.cpp:24-28).
                    // max == min divides 0 by 0 there (UB on the cast); such rows are stored as 0 here.
                    const float f = levels * (src[i] - lo) / (hi - lo);
                    q[r * dim + i] = (f == f) ? static_cast<uint8_t>(f) : 0;
                }
            }
        });
        const IndexHeader idx = writeIndex(w, I::STORAGE_UNIFORM, static_cast<uint32_t>(dim));
        const auto lay = w.vtable({{I::UNIFORM_NODES, 4}, {I::UNIFORM_LEVELS, 1}});
        const size_t t = w.table(lay);
        w.link(w.fieldPos(idx.layout, idx.table, I::INDEX_STORAGE), t);
        w.setScalar<uint8_t>(lay, t, I::UNIFORM_LEVELS, levels);
        std::vector<uint32_t> zeros(n, 0);
        const size_t nodesVec = w.vector(zeros.data(), n, 4);
        w.link(w.fieldPos(lay, t, I::UNIFORM_NODES), nodesVec);
        const auto nodeLay = w.vtable({{I::UNODE_WORD, 4}, {I::UNODE_VALUES, 4}});
        // scalars equal to their default are omitted by the flatbuffers builder:
        // one vtable per (min present, max present) combination
        fb::Writer::TableLayout vecLay[4];
        for (int m = 0; m < 4; ++m) {
            std::vector<fb::Writer::FieldSpec> fields = {{I::UVEC_VALUES, 4}};
            if (m & 1) fields.push_back({I::UVEC_MIN, 4});
            if (m & 2) fields.push_back({I::UVEC_MAX, 4});
            vecLay[m] = w.vtable(fields);
        }
        for (size_t i = 0; i < n; ++i) {
            const uint32_t src = order[i];
            const size_t node = w.table(nodeLay);
            w.link(nodesVec + 4 + 4 * i, node);
            w.link(w.fieldPos(nodeLay, node, I::UNODE_WORD),
                   w.string(words_[src].c_str(), std::strlen(words_[src].c_str())));
            const int m = (mins[src] != 0.0f ? 1 : 0) | (maxs[src] != 0.0f ? 2 : 0);
            const size_t vec = w.table(vecLay[m]);
            w.link(w.fieldPos(nodeLay, node, I::UNODE_VALUES), vec);
            if (m & 1) w.setScalar<float>(vecLay[m], vec, I::UVEC_MIN, mins[src]);
            if (m & 2) w.setScalar<float>(vecLay[m], vec, I::UVEC_MAX, maxs[src]);
            w.link(w.fieldPos(vecLay[m], vec, I::UVEC_VALUES), w.vector(q.data() + src * dim, dim, 1));
        }
        return std::move(w.buffer());
    }

    // full: one (word, fp32 vector) node per row (src/full_compression.cpp:9-31)
    const IndexHeader idx = writeIndex(w, I::STORAGE_FULL, static_cast<uint32_t>(dim));
    const auto lay = w.vtable({{I::FULL_NODES, 4}});
    const size_t t = w.table(lay);
    w.link(w.fieldPos(idx.layout, idx.table, I::INDEX_STORAGE), t);
    std::vector<uint32_t> zeros(n, 0);
    const size_t nodesVec = w.vector(zeros.data(), n, 4);
    w.link(w.fieldPos(lay, t, I::FULL_NODES), nodesVec);
    const auto nodeLay = w.vtable({{I::FNODE_WORD, 4}, {I::FNODE_VALUES, 4}});
    for (size_t i = 0; i < n; ++i) {
        const uint32_t src = order[i];
        const size_t node = w.table(nodeLay);
        w.link(nodesVec + 4 + 4 * i, node);
        w.link(w.fieldPos(nodeLay, node, I::FNODE_WORD),
               w.string(words_[src].c_str(), std::strlen(words_[src].c_str())));
        w.link(w.fieldPos(nodeLay, node, I::FNODE_VALUES), w.vector(values_.data() + src * dim, dim, 4));
    }
    return std::move(w.buffer());
}

void Builder::save(const std::string& filename, unsigned numThreads)
{
    const std::vector<uint8_t> bytes = serialize(numThreads);
    std::ofstream f(filename, std::ios::binary);
    if (!f) {
        throw std::runtime_error("Failed to open " + filename + " for writing");
    }
    f.write(reinterpret_cast<const char*>(bytes.data()), static_cast<std::streamsize>(bytes.size()));
    if (!f) {
        throw std::runtime_error("Failed to write " + filename);
    }
}

}  // namespace memb_b200
