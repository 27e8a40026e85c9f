// Extension module `_memb`: same classes, method names and argument meaning as
// the reference binding (python/memb_bindings.cpp:11-72), implemented purely on
// the C ABI of include/memb_b200.h.  Additive members (not in the reference) are
// grouped at the end of each class.
#include "../../include/memb_b200.h"

#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <memory>
#include <mutex>
#include <thread>
#include <stdexcept>
#include <string>
#include <vector>

namespace py = pybind11;

namespace {

void check(memb_b200_status status)
{
    if (status == MEMB_B200_OK) {
        return;
    }
    const std::string message = memb_b200_last_error();
    if (status == MEMB_B200_ERR_ARGUMENT) {
        throw std::invalid_argument(message);
    }
    // the reference surfaces every failure as std::runtime_error / std::exception
    // (src/tests.cpp:134-153), which pybind11 maps to RuntimeError
    throw std::runtime_error(message);
}

// Query words as one byte blob + n+1 offsets, the form the C ABI takes.
// list[str]: Python str -> UTF-8 like pybind11's std::string caster used by the
// reference (python/memb_bindings.cpp:54-56).  Additive fast paths:
// a (blob, offsets) pair is passed through without a copy, NumPy 'S' / 'U' arrays are
// packed without creating Python objects.
struct Marshalled {
    std::string blob;
    std::vector<uint64_t> ownedOffsets;
    std::unique_ptr<char[]> rawBlob;  // list path: filled by several threads, never zeroed
    std::unique_ptr<uint64_t[]> rawOffsets;
    py::object keepBlob, keepOffsets;  // zero-copy sources
    const char* data = "";
    const uint64_t* offsets = nullptr;
    size_t n = 0;

    void adoptOwned()
    {
        data = blob.data();
        offsets = ownedOffsets.data();
        n = ownedOffsets.size() - 1;
    }
};

// One word of a list: where its UTF-8 bytes are.  ASCII str objects and bytes expose
// their buffer without any call into the interpreter; anything else (non-ASCII str) goes
// through PyUnicode_AsUTF8AndSize, which may allocate and therefore needs the GIL.
struct WordView {
    const char* data;
    size_t size;
};

inline bool viewWithoutInterpreter(PyObject* item, WordView& v)
{
    if (PyUnicode_Check(item)) {
        if (PyUnicode_IS_COMPACT_ASCII(item)) {
            v.data = reinterpret_cast<const char*>(PyUnicode_1BYTE_DATA(item));
            v.size = static_cast<size_t>(PyUnicode_GET_LENGTH(item));
            return true;
        }
        return false;
    }
    if (PyBytes_Check(item)) {
        v.data = PyBytes_AS_STRING(item);
        v.size = static_cast<size_t>(PyBytes_GET_SIZE(item));
        return true;
    }
    return false;
}

inline void viewWithInterpreter(PyObject* item, WordView& v)
{
    if (!PyUnicode_Check(item)) {
        throw py::type_error("batch_embedding(): words must be str");
    }
    Py_ssize_t size = 0;
    const char* utf8 = PyUnicode_AsUTF8AndSize(item, &size);
    if (!utf8) {
        throw py::error_already_set();
    }
    v.data = utf8;
    v.size = static_cast<size_t>(size);
}

constexpr size_t PARALLEL_MARSHAL_MIN_WORDS = 1u << 16;

unsigned marshalThreads(size_t n)
{
    if (n < PARALLEL_MARSHAL_MIN_WORDS) {
        return 1;
    }
    if (const char* forced = std::getenv("MEMB_B200_MARSHAL_THREADS")) {
        return static_cast<unsigned>(std::max(1, std::atoi(forced)));
    }
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    return static_cast<unsigned>(std::min<size_t>({hw, 16, n / (PARALLEL_MARSHAL_MIN_WORDS / 4)}));
}

template <class Body>
void forRanges(size_t n, unsigned threads, Body body)
{
    if (threads <= 1) {
        body(0u, size_t(0), n);
        return;
    }
    std::vector<std::thread> pool;
    const size_t per = (n + threads - 1) / threads;
    for (unsigned t = 1; t < threads; ++t) {
        pool.emplace_back([=] { body(t, std::min(n, t * per), std::min(n, (t + 1) * per)); });
    }
    body(0u, size_t(0), std::min(n, per));
    for (std::thread& th : pool) {
        th.join();
    }
}

// list[str] -> blob + offsets.  The caller holds the GIL for the whole call, so no other
// Python thread can touch the list or its items; the worker threads only READ object
// headers and character data (no reference counts, no allocation), which is what makes
// the two passes below safe to run on several host threads.  1 M short words: 38 ms on
// one thread, the cost that dominated words -> device tensor.
void marshalList(const py::handle& words, Marshalled& m)
{
    PyObject* seq = words.ptr();
    const size_t n = static_cast<size_t>(PySequence_Fast_GET_SIZE(seq));
    PyObject** items = PySequence_Fast_ITEMS(seq);
    const unsigned threads = marshalThreads(n);
    std::unique_ptr<WordView[]> views(new WordView[n]);
    std::vector<std::vector<size_t>> slow(threads);
    std::vector<uint64_t> rangeBytes(threads + 1, 0);
    forRanges(n, threads, [&](unsigned t, size_t begin, size_t end) {
        uint64_t sum = 0;
        for (size_t i = begin; i < end; ++i) {
            if (viewWithoutInterpreter(items[i], views[i])) {
                sum += views[i].size;
            } else {
                slow[t].push_back(i);
            }
        }
        rangeBytes[t + 1] = sum;
    });
    for (unsigned t = 0; t < threads; ++t) {
        for (size_t i : slow[t]) {
            viewWithInterpreter(items[i], views[i]);
            rangeBytes[t + 1] += views[i].size;
        }
    }
    for (unsigned t = 0; t < threads; ++t) {
        rangeBytes[t + 1] += rangeBytes[t];
    }
    m.rawBlob.reset(new char[rangeBytes[threads] + 1]);
    m.rawOffsets.reset(new uint64_t[n + 1]);
    char* blob = m.rawBlob.get();
    uint64_t* offsets = m.rawOffsets.get();
    forRanges(n, threads, [&](unsigned t, size_t begin, size_t end) {
        uint64_t at = rangeBytes[t];
        for (size_t i = begin; i < end; ++i) {
            offsets[i] = at;
            std::memcpy(blob + at, views[i].data, views[i].size);
            at += views[i].size;
        }
    });
    offsets[n] = rangeBytes[threads];
    m.data = blob;
    m.offsets = offsets;
    m.n = n;
}

void appendUtf8(std::string& out, uint32_t c)
{
    if (c < 0x80) {
        out.push_back(static_cast<char>(c));
    } else if (c < 0x800) {
        out.push_back(static_cast<char>(0xC0 | (c >> 6)));
        out.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    } else if (c < 0x10000) {
        out.push_back(static_cast<char>(0xE0 | (c >> 12)));
        out.push_back(static_cast<char>(0x80 | ((c >> 6) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    } else {
        out.push_back(static_cast<char>(0xF0 | (c >> 18)));
        out.push_back(static_cast<char>(0x80 | ((c >> 12) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | ((c >> 6) & 0x3F)));
        out.push_back(static_cast<char>(0x80 | (c & 0x3F)));
    }
}

// 1-D NumPy array of dtype 'S<w>' (bytes, NUL padded) or 'U<w>' (UCS-4, NUL padded).
void marshalArray(const py::array& array, Marshalled& m)
{
    if (array.ndim() != 1) {
        throw py::type_error("batch_embedding(): word array must be 1-dimensional");
    }
    const char kind = array.dtype().kind();
    const size_t n = static_cast<size_t>(array.shape(0));
    const size_t item = static_cast<size_t>(array.itemsize());
    const py::ssize_t stride = array.strides(0);
    const char* base = static_cast<const char*>(array.data());
    m.ownedOffsets.reserve(n + 1);
    m.ownedOffsets.push_back(0);
    if (kind == 'S') {
        m.blob.reserve(n * item);
        for (size_t i = 0; i < n; ++i) {
            const char* p = base + static_cast<py::ssize_t>(i) * stride;
            m.blob.append(p, strnlen(p, item));
            m.ownedOffsets.push_back(m.blob.size());
        }
    } else if (kind == 'U') {
        if (array.dtype().byteorder() == '>') {
            throw py::type_error("batch_embedding(): big-endian unicode arrays are not supported");
        }
        const size_t chars = item / 4;
        m.blob.reserve(n * chars);
        for (size_t i = 0; i < n; ++i) {
            const char* p = base + static_cast<py::ssize_t>(i) * stride;
            size_t len = chars;
            uint32_t c;
            while (len > 0 && (std::memcpy(&c, p + (len - 1) * 4, 4), c == 0)) {
                --len;  // NumPy strips trailing NULs only
            }
            for (size_t k = 0; k < len; ++k) {
                std::memcpy(&c, p + k * 4, 4);
                appendUtf8(m.blob, c);
            }
            m.ownedOffsets.push_back(m.blob.size());
        }
    } else {
        throw py::type_error("batch_embedding(): word array must have dtype 'S' or 'U'");
    }
    m.adoptOwned();
}

// (blob, offsets): blob = any contiguous byte buffer, offsets = uint64[n+1], ascending.
void marshalPacked(const py::tuple& packed, Marshalled& m)
{
    if (packed.size() != 2) {
        throw py::type_error("batch_embedding(): packed words must be a (blob, offsets) pair");
    }
    py::buffer blob = packed[0].cast<py::buffer>();
    const py::buffer_info blobInfo = blob.request();
    auto offsets = py::array_t<uint64_t, py::array::c_style | py::array::forcecast>::ensure(packed[1]);
    if (!offsets || offsets.ndim() != 1 || offsets.shape(0) < 1) {
        throw py::type_error("batch_embedding(): offsets must be a 1-D uint64 array with n+1 entries");
    }
    const size_t blobBytes = static_cast<size_t>(blobInfo.size * blobInfo.itemsize);
    const uint64_t* o = offsets.data();
    const size_t n = static_cast<size_t>(offsets.shape(0)) - 1;
    for (size_t i = 0; i < n; ++i) {
        if (o[i] > o[i + 1]) {
            throw py::value_error("batch_embedding(): offsets must be ascending");
        }
    }
    if (o[n] > blobBytes) {
        throw py::value_error("batch_embedding(): offsets run past the end of the blob");
    }
    m.keepBlob = blob;
    m.keepOffsets = offsets;
    m.data = static_cast<const char*>(blobInfo.ptr);
    m.offsets = o;
    m.n = n;
}

std::unique_ptr<Marshalled> marshal(const py::handle& words)
{
    auto m = std::make_unique<Marshalled>();
    if (PyList_Check(words.ptr())) {
        marshalList(words, *m);
    } else if (PyTuple_Check(words.ptr())) {
        marshalPacked(py::reinterpret_borrow<py::tuple>(words), *m);
    } else if (py::isinstance<py::array>(words)) {
        const py::array array = py::reinterpret_borrow<py::array>(words);
        if (array.dtype().kind() == 'O') {
            const py::list list = array.attr("tolist")();
            marshalList(list, *m);
        } else {
            marshalArray(array, *m);
        }
    } else {
        throw py::type_error("batch_embedding(): words must be a list of str");
    }
    return m;
}

// Results that go back to Python as NumPy arrays are DMA targets: a fresh pageable
// array costs a page fault per 4 KB plus the driver's bounce buffer (1 M rows: 250 ms,
// 4.8 GB/s), a pinned one runs at PCIe rate (21.5 ms).  Pinning is itself slow, so
// buffers are recycled: the array's base object hands its buffer back when NumPy frees
// the array.  At most `limit` bytes stay cached; 0 switches pinned results off.
class PinnedResults {
public:
    static PinnedResults& instance()
    {
        static PinnedResults* pool = new PinnedResults;  // never destroyed: outlives CUDA teardown order
        return *pool;
    }

    void setLimit(uint64_t bytes)
    {
        std::vector<Block> drop;
        {
            std::lock_guard<std::mutex> lock(mutex_);
            limit_ = bytes;
            while (cached_ > limit_ && !free_.empty()) {
                cached_ -= free_.back().capacity;
                drop.push_back(free_.back());
                free_.pop_back();
            }
        }
        for (const Block& b : drop) {
            memb_b200_host_free(b.ptr);
        }
    }
    uint64_t limit() const { return limit_; }
    uint64_t cachedBytes() const { return cached_; }

    py::array_t<float> make(size_t rows, size_t cols)
    {
        const size_t bytes = rows * cols * sizeof(float);
        const std::vector<py::ssize_t> shape{static_cast<py::ssize_t>(rows), static_cast<py::ssize_t>(cols)};
        if (bytes < MIN_BYTES || limit_ == 0) {
            return py::array_t<float>(shape);
        }
        Block block = take(bytes);
        if (!block.ptr) {
            return py::array_t<float>(shape);  // no pinned memory to be had: pageable result
        }
        auto* owned = new Block(block);
        py::capsule base(owned, [](void* p) {
            Block* b = static_cast<Block*>(p);
            PinnedResults::instance().give(*b);
            delete b;
        });
        return py::array_t<float>(shape, static_cast<float*>(block.ptr), base);
    }

private:
    struct Block {
        void* ptr = nullptr;
        size_t capacity = 0;
    };
    static constexpr size_t MIN_BYTES = size_t(1) << 20;
    static constexpr size_t GRANULE = size_t(2) << 20;

    Block take(size_t bytes)
    {
        {
            std::lock_guard<std::mutex> lock(mutex_);
            size_t best = free_.size();
            for (size_t i = 0; i < free_.size(); ++i) {
                // a small result must not keep a huge block alive
                if (free_[i].capacity >= bytes && free_[i].capacity / 4 <= bytes &&
                    (best == free_.size() || free_[i].capacity < free_[best].capacity)) {
                    best = i;
                }
            }
            if (best != free_.size()) {
                const Block b = free_[best];
                free_.erase(free_.begin() + static_cast<std::ptrdiff_t>(best));
                cached_ -= b.capacity;
                return b;
            }
        }
        Block b;
        b.capacity = (bytes + GRANULE - 1) / GRANULE * GRANULE;
        if (memb_b200_host_alloc(b.capacity, &b.ptr) != MEMB_B200_OK) {
            b.ptr = nullptr;
        }
        return b;
    }

    void give(const Block& b)
    {
        {
            std::lock_guard<std::mutex> lock(mutex_);
            if (cached_ + b.capacity <= limit_) {
                free_.push_back(b);
                cached_ += b.capacity;
                return;
            }
        }
        memb_b200_host_free(b.ptr);
    }

    std::mutex mutex_;
    std::vector<Block> free_;
    uint64_t cached_ = 0;
    uint64_t limit_ = uint64_t(8) << 30;
};

py::array_t<float> makeResult(size_t rows, size_t cols)
{
    return PinnedResults::instance().make(rows, cols);
}

class Reader {
public:
    Reader(const std::string& filename, size_t numThreads, int device)
    {
        check(memb_b200_open(filename.c_str(), device, static_cast<unsigned>(numThreads), &impl_));
    }
    ~Reader() { memb_b200_close(impl_); }
    Reader(const Reader&) = delete;
    Reader& operator=(const Reader&) = delete;

    memb_b200_reader* handle() const { return impl_; }
    size_t dim() const { return memb_b200_dim(impl_); }
    size_t rows() const { return memb_b200_rows(impl_); }
    int device() const { return memb_b200_device(impl_); }

    py::array_t<float> wordEmbedding(const std::string& word) const
    {
        py::array_t<float> result(static_cast<py::ssize_t>(dim()));
        const uint64_t offsets[2] = {0, word.size()};
        check(memb_b200_lookup_words(impl_, word.data(), offsets, 1, result.mutable_data()));
        return result;
    }

    // words: list[str] as in the reference; additive: (blob, offsets) or a NumPy 'S'/'U' array
    py::array_t<float> batchEmbedding(const py::object& words) const
    {
        const auto m = marshal(words);
        py::array_t<float> result = makeResult(m->n, dim());
        float* out = result.mutable_data();
        {
            py::gil_scoped_release release;
            check(memb_b200_lookup_words(impl_, m->data, m->offsets, m->n, out));
        }
        return result;
    }

    // additive: rows written to caller memory (torch data_ptr); returns the row count
    // (device output: queued on the caller's CUDA stream, memb_b200_lookup_words_on)
    uint64_t batchEmbeddingInto(const py::object& words, uintptr_t out, uint64_t capacityRows, bool outOnDevice,
                                uintptr_t stream) const
    {
        const auto m = marshal(words);
        if (m->n > capacityRows) {
            throw std::invalid_argument("batch_embedding_into: output holds fewer rows than words given");
        }
        py::gil_scoped_release release;
        if (outOnDevice) {
            check(memb_b200_lookup_words_on(impl_, m->data, m->offsets, m->n, reinterpret_cast<float*>(out),
                                            reinterpret_cast<void*>(stream)));
        } else {
            check(memb_b200_lookup_words_to(impl_, m->data, m->offsets, m->n, reinterpret_cast<float*>(out),
                                            MEMB_B200_HOST));
        }
        return m->n;
    }

    static uint64_t countWords(const py::object& words) { return marshal(words)->n; }

    std::vector<std::string> keys() const
    {
        const char* words = nullptr;
        const uint32_t* offsets = nullptr;
        uint64_t rows = 0;
        check(memb_b200_keys(impl_, &words, &offsets, &rows));
        std::vector<std::string> result;
        result.reserve(rows);
        for (uint64_t r = 0; r < rows; ++r) {
            result.emplace_back(words + offsets[r]);
        }
        return result;
    }

    // ---- additive ----
    py::array_t<int32_t> resolve(const py::object& words) const
    {
        const auto m = marshal(words);
        py::array_t<int32_t> ids(static_cast<py::ssize_t>(m->n));
        int32_t* out = ids.mutable_data();
        {
            py::gil_scoped_release release;
            check(memb_b200_resolve(impl_, m->data, m->offsets, m->n, out));
        }
        return ids;
    }

    // raw-pointer form: query blob, offsets and ids already on the device; asynchronous on `stream`
    void resolveInto(uintptr_t queryBytes, uintptr_t queryOffsets, uint64_t n, uintptr_t ids, uintptr_t stream) const
    {
        check(memb_b200_resolve_device(impl_, reinterpret_cast<const char*>(queryBytes),
                                       reinterpret_cast<const uint64_t*>(queryOffsets), n,
                                       reinterpret_cast<int32_t*>(ids), reinterpret_cast<void*>(stream)));
    }

    py::array_t<float> batchEmbeddingIds(py::array_t<int32_t, py::array::c_style | py::array::forcecast> ids) const
    {
        const size_t n = static_cast<size_t>(ids.size());
        py::array_t<float> result = makeResult(n, dim());
        float* out = result.mutable_data();
        const int32_t* in = ids.data();
        {
            py::gil_scoped_release release;
            check(memb_b200_decode_ids(impl_, in, MEMB_B200_HOST, n, out, MEMB_B200_HOST, nullptr));
        }
        return result;
    }

    py::array_t<float> decodeRange(uint64_t begin, uint64_t end) const
    {
        if (begin > end) {
            throw std::invalid_argument("decode_range: begin > end");
        }
        py::array_t<float> result = makeResult(end - begin, dim());
        float* out = result.mutable_data();
        {
            py::gil_scoped_release release;
            check(memb_b200_decode_range(impl_, begin, end, out, MEMB_B200_HOST, nullptr));
        }
        return result;
    }

    // raw-pointer forms for device-resident tensors (torch data_ptr / stream)
    void decodeIdsInto(uintptr_t ids, bool idsOnDevice, uint64_t n, uintptr_t out, bool outOnDevice,
                       uintptr_t stream) const
    {
        py::gil_scoped_release release;
        check(memb_b200_decode_ids(impl_, reinterpret_cast<const int32_t*>(ids),
                                   idsOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST, n, reinterpret_cast<float*>(out),
                                   outOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST, reinterpret_cast<void*>(stream)));
    }

    void decodeRangeInto(uint64_t begin, uint64_t end, uintptr_t out, bool outOnDevice, uintptr_t stream) const
    {
        py::gil_scoped_release release;
        check(memb_b200_decode_range(impl_, begin, end, reinterpret_cast<float*>(out),
                                     outOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST,
                                     reinterpret_cast<void*>(stream)));
    }

    // rows [begin, end) stored through an NVLS multicast address (memb_b200_decode_range_multicast)
    void decodeRangeMulticast(uint64_t begin, uint64_t end, uintptr_t outMulticast, uintptr_t stream) const
    {
        py::gil_scoped_release release;
        check(memb_b200_decode_range_multicast(impl_, begin, end, reinterpret_cast<float*>(outMulticast),
                                               reinterpret_cast<void*>(stream)));
    }

    py::dict loadStats() const
    {
        memb_b200_load_stats s;
        check(memb_b200_get_load_stats(impl_, &s));
        py::dict d;
        d["parse_ms"] = s.parse_ms;
        d["upload_ms"] = s.upload_ms;
        d["prepare_ms"] = s.prepare_ms;
        d["device_bytes"] = s.device_bytes;
        d["payload_bytes"] = s.payload_bytes;
        d["max_code_len"] = s.max_code_len;
        d["lut_bits"] = s.lut_bits;
        d["lut_replicas"] = s.lut_replicas;
        d["codebook_size"] = s.codebook_size;
        d["storage_type"] = memb_b200_storage_type(impl_);
        d["device"] = memb_b200_device(impl_);
        return d;
    }

private:
    memb_b200_reader* impl_ = nullptr;
};

memb_b200_union_mode parseMode(const std::string& mode)
{
    if (mode == "average") return MEMB_B200_AVERAGE;
    if (mode == "concatenate") return MEMB_B200_CONCATENATE;
    throw std::invalid_argument("unknown union mode " + mode);
}

struct UnionHandles {
    memb_b200_union_mode mode;
    std::vector<memb_b200_reader*> handles;
    size_t dim = 0;
};

UnionHandles unionHandles(const std::vector<Reader*>& readers, const std::string& mode)
{
    UnionHandles u;
    u.mode = parseMode(mode);
    for (const Reader* r : readers) {
        u.handles.push_back(r->handle());
        u.dim = (u.mode == MEMB_B200_AVERAGE) ? r->dim() : u.dim + r->dim();
    }
    return u;
}

py::array_t<float> unionBatchEmbedding(const std::vector<Reader*>& readers, const std::string& mode,
                                       const py::object& words)
{
    UnionHandles u = unionHandles(readers, mode);
    const auto q = marshal(words);
    py::array_t<float> result = makeResult(q->n, u.dim);
    float* out = result.mutable_data();
    {
        py::gil_scoped_release release;
        check(memb_b200_union_lookup_words(u.handles.data(), static_cast<uint32_t>(u.handles.size()), u.mode, q->data,
                                           q->offsets, q->n, out));
    }
    return result;
}

uint64_t unionBatchEmbeddingInto(const std::vector<Reader*>& readers, const std::string& mode, const py::object& words,
                                 uintptr_t out, uint64_t capacityRows, bool outOnDevice, uintptr_t stream)
{
    UnionHandles u = unionHandles(readers, mode);
    const auto q = marshal(words);
    if (q->n > capacityRows) {
        throw std::invalid_argument("union_batch_embedding_into: output holds fewer rows than words given");
    }
    py::gil_scoped_release release;
    if (outOnDevice) {
        check(memb_b200_union_lookup_words_on(u.handles.data(), static_cast<uint32_t>(u.handles.size()), u.mode, q->data,
                                              q->offsets, q->n, reinterpret_cast<float*>(out),
                                              reinterpret_cast<void*>(stream)));
    } else {
        check(memb_b200_union_lookup_words_to(u.handles.data(), static_cast<uint32_t>(u.handles.size()), u.mode, q->data,
                                              q->offsets, q->n, reinterpret_cast<float*>(out), MEMB_B200_HOST));
    }
    return q->n;
}

void unionDecodeIdsInto(const std::vector<Reader*>& readers, const std::string& mode, uintptr_t ids, bool idsOnDevice,
                        uint64_t n, uintptr_t out, bool outOnDevice, uintptr_t stream)
{
    const memb_b200_union_mode m = parseMode(mode);
    std::vector<memb_b200_reader*> handles;
    for (const Reader* r : readers) {
        handles.push_back(r->handle());
    }
    py::gil_scoped_release release;
    check(memb_b200_union_decode_ids(handles.data(), static_cast<uint32_t>(handles.size()), m,
                                     reinterpret_cast<const int32_t*>(ids),
                                     idsOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST, n, reinterpret_cast<float*>(out),
                                     outOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST, reinterpret_cast<void*>(stream)));
}

class Builder {
public:
    Builder(size_t dim, const std::string& storageName, size_t bitsPerWeight)
    {
        check(memb_b200_builder_create(dim, storageName.c_str(), bitsPerWeight, &impl_));
    }
    ~Builder() { memb_b200_builder_destroy(impl_); }
    Builder(const Builder&) = delete;
    Builder& operator=(const Builder&) = delete;

    void addWord(const std::string& word, py::array_t<float, py::array::c_style> values)
    {
        const py::buffer_info info = values.request();
        if (info.ndim != 1) {
            throw std::runtime_error("Word vector must be 1-dimensional");  // python/memb_bindings.cpp:19-21
        }
        check(memb_b200_builder_add_word(impl_, word.data(), word.size(), static_cast<const float*>(info.ptr),
                                         static_cast<uint64_t>(info.shape[0])));
    }

    void addWords(const py::list& words, py::array_t<float, py::array::c_style | py::array::forcecast> matrix)
    {
        const py::buffer_info info = matrix.request();
        if (info.ndim != 2 || static_cast<size_t>(info.shape[0]) != words.size()) {
            throw std::runtime_error("add_words expects a (len(words), dim) float32 matrix");
        }
        const auto m = marshal(words);
        // a width mismatch is reported per word by the builder itself
        if (words.size() && static_cast<uint64_t>(info.shape[1]) != dim_) {
            check(memb_b200_builder_add_word(impl_, m->data, m->offsets[1], static_cast<const float*>(info.ptr),
                                             static_cast<uint64_t>(info.shape[1])));
            return;
        }
        check(memb_b200_builder_add_words(impl_, m->data, m->offsets, words.size(),
                                          static_cast<const float*>(info.ptr)));
    }

    void save(const std::string& filename, unsigned numThreads)
    {
        py::gil_scoped_release release;
        check(memb_b200_builder_save(impl_, filename.c_str(), numThreads));
    }

    void setDim(size_t dim) { dim_ = dim; }
    void setDevice(int device) { check(memb_b200_builder_set_device(impl_, device)); }

private:
    memb_b200_builder* impl_ = nullptr;
    uint64_t dim_ = 0;
};

std::vector<std::string> availableCompressionStrategies()
{
    std::vector<std::string> names;
    for (const char* p = memb_b200_available_compression_strategies(); *p; p += std::strlen(p) + 1) {
        names.emplace_back(p);
    }
    return names;
}

}  // namespace

PYBIND11_MODULE(_memb, m)
{
    m.doc() = "memb_b200: B200-native drop-in for memb's _memb extension module";

    py::class_<Builder>(m, "Builder")
        .def(py::init([](size_t dim, const std::string& storage, size_t bits, int device) {
                 auto b = std::make_unique<Builder>(dim, storage, bits);
                 b->setDim(dim);
                 if (device >= 0) {
                     b->setDevice(device);
                 }
                 return b;
             }),
             py::arg("dim"), py::arg("storage_name"), py::arg("bits_per_weight"), py::arg("device") = -1)
        .def("set_device", &Builder::setDevice, py::arg("device"))
        .def("add_word", &Builder::addWord)
        .def("save", [](Builder& b, const std::string& filename) { b.save(filename, 0); })
        // additive
        .def("add_words", &Builder::addWords, py::arg("words"), py::arg("matrix"))
        .def("save_threads", &Builder::save, py::arg("filename"), py::arg("num_threads"));

    py::class_<Reader>(m, "Reader")
        .def(py::init<std::string, size_t, int>(), py::arg("filename"), py::arg("num_threads"), py::arg("device") = 0)
        .def("dim", &Reader::dim)
        .def("word_embedding", &Reader::wordEmbedding)
        .def("batch_embedding", &Reader::batchEmbedding)
        .def("keys", &Reader::keys)
        // additive
        .def("rows", &Reader::rows)
        .def("device", &Reader::device)
        .def("resolve", &Reader::resolve)
        .def("resolve_into", &Reader::resolveInto, py::arg("query_bytes"), py::arg("query_offsets"), py::arg("n"),
             py::arg("ids"), py::arg("stream") = 0)
        .def("batch_embedding_ids", &Reader::batchEmbeddingIds)
        .def("batch_embedding_into", &Reader::batchEmbeddingInto, py::arg("words"), py::arg("out"),
             py::arg("capacity_rows"), py::arg("out_on_device"), py::arg("stream") = 0)
        .def_static("count_words", &Reader::countWords)
        .def("decode_range", &Reader::decodeRange)
        .def("decode_ids_into", &Reader::decodeIdsInto, py::arg("ids"), py::arg("ids_on_device"), py::arg("n"),
             py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0)
        .def("decode_range_into", &Reader::decodeRangeInto, py::arg("begin"), py::arg("end"), py::arg("out"),
             py::arg("out_on_device"), py::arg("stream") = 0)
        .def("decode_range_multicast", &Reader::decodeRangeMulticast, py::arg("begin"), py::arg("end"),
             py::arg("out_multicast"), py::arg("stream") = 0)
        .def("load_stats", &Reader::loadStats);

    m.def("available_compression_strategies", &availableCompressionStrategies);
    // additive
    m.def("union_batch_embedding", &unionBatchEmbedding);
    m.def("union_batch_embedding_into", &unionBatchEmbeddingInto, py::arg("readers"), py::arg("mode"), py::arg("words"),
          py::arg("out"), py::arg("capacity_rows"), py::arg("out_on_device"), py::arg("stream") = 0);
    m.def("union_decode_ids_into", &unionDecodeIdsInto, py::arg("readers"), py::arg("mode"), py::arg("ids"),
          py::arg("ids_on_device"), py::arg("n"), py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0);
    // the marshalled form of a word collection: (UTF-8 bytes, uint64 offsets[n+1])
    m.def("pack_words", [](const py::object& words) {
        const auto q = marshal(words);
        py::array_t<uint64_t> offsets(static_cast<py::ssize_t>(q->n + 1));
        uint64_t* rebased = offsets.mutable_data();
        for (size_t i = 0; i <= q->n; ++i) {
            rebased[i] = q->offsets[i] - q->offsets[0];
        }
        return py::make_tuple(py::bytes(q->data + q->offsets[0], q->offsets[q->n] - q->offsets[0]), offsets);
    });
    m.def("set_pinned_result_limit", [](uint64_t bytes) { PinnedResults::instance().setLimit(bytes); },
          "bytes of pinned result buffers kept for reuse (default 8 GiB); 0 = results in pageable memory");
    m.def("pinned_result_state", [] {
        return py::make_tuple(PinnedResults::instance().limit(), PinnedResults::instance().cachedBytes());
    });
    m.def("device_count", &memb_b200_device_count);
    m.def("launch_count", &memb_b200_launch_count);
    m.def("set_option", [](const std::string& name, int64_t value) { check(memb_b200_set_option(name.c_str(), value)); });
}
