// Extension module `_memb`: same classes, method names and argument meaning as
// the reference binding (python/memb_bindings.cpp:11-72), implemented purely on
// the C ABI of include/memb_b200.h.  Additive members (not in the reference) are
// grouped at the end of each class.
#include "../../include/memb_b200.h"

#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstring>
#include <memory>
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

void marshalList(const py::handle& words, Marshalled& m)
{
    PyObject* seq = words.ptr();
    const Py_ssize_t n = PySequence_Fast_GET_SIZE(seq);
    PyObject** items = PySequence_Fast_ITEMS(seq);
    m.ownedOffsets.reserve(static_cast<size_t>(n) + 1);
    m.ownedOffsets.push_back(0);
    m.blob.reserve(static_cast<size_t>(n) * 8);
    for (Py_ssize_t i = 0; i < n; ++i) {
        PyObject* item = items[i];
        if (PyUnicode_Check(item)) {
            Py_ssize_t size = 0;
            const char* utf8 = PyUnicode_AsUTF8AndSize(item, &size);
            if (!utf8) {
                throw py::error_already_set();
            }
            m.blob.append(utf8, static_cast<size_t>(size));
        } else if (PyBytes_Check(item)) {
            m.blob.append(PyBytes_AS_STRING(item), static_cast<size_t>(PyBytes_GET_SIZE(item)));
        } else {
            throw py::type_error("batch_embedding(): words must be str");
        }
        m.ownedOffsets.push_back(m.blob.size());
    }
    m.adoptOwned();
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
        py::array_t<float> result({static_cast<py::ssize_t>(m->n), static_cast<py::ssize_t>(dim())});
        float* out = result.mutable_data();
        {
            py::gil_scoped_release release;
            check(memb_b200_lookup_words(impl_, m->data, m->offsets, m->n, out));
        }
        return result;
    }

    // additive: rows written to caller memory (torch data_ptr); returns the row count
    uint64_t batchEmbeddingInto(const py::object& words, uintptr_t out, uint64_t capacityRows, bool outOnDevice) const
    {
        const auto m = marshal(words);
        if (m->n > capacityRows) {
            throw std::invalid_argument("batch_embedding_into: output holds fewer rows than words given");
        }
        py::gil_scoped_release release;
        check(memb_b200_lookup_words_to(impl_, m->data, m->offsets, m->n, reinterpret_cast<float*>(out),
                                        outOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST));
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

    py::array_t<float> batchEmbeddingIds(py::array_t<int32_t, py::array::c_style | py::array::forcecast> ids) const
    {
        const size_t n = static_cast<size_t>(ids.size());
        py::array_t<float> result({static_cast<py::ssize_t>(n), static_cast<py::ssize_t>(dim())});
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
        py::array_t<float> result({static_cast<py::ssize_t>(end - begin), static_cast<py::ssize_t>(dim())});
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
    py::array_t<float> result({static_cast<py::ssize_t>(q->n), static_cast<py::ssize_t>(u.dim)});
    float* out = result.mutable_data();
    {
        py::gil_scoped_release release;
        check(memb_b200_union_lookup_words(u.handles.data(), static_cast<uint32_t>(u.handles.size()), u.mode, q->data,
                                           q->offsets, q->n, out));
    }
    return result;
}

uint64_t unionBatchEmbeddingInto(const std::vector<Reader*>& readers, const std::string& mode, const py::object& words,
                                 uintptr_t out, uint64_t capacityRows, bool outOnDevice)
{
    UnionHandles u = unionHandles(readers, mode);
    const auto q = marshal(words);
    if (q->n > capacityRows) {
        throw std::invalid_argument("union_batch_embedding_into: output holds fewer rows than words given");
    }
    py::gil_scoped_release release;
    check(memb_b200_union_lookup_words_to(u.handles.data(), static_cast<uint32_t>(u.handles.size()), u.mode, q->data,
                                          q->offsets, q->n, reinterpret_cast<float*>(out),
                                          outOnDevice ? MEMB_B200_DEVICE : MEMB_B200_HOST));
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
        .def(py::init([](size_t dim, const std::string& storage, size_t bits) {
            auto b = std::make_unique<Builder>(dim, storage, bits);
            b->setDim(dim);
            return b;
        }))
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
        .def("batch_embedding_ids", &Reader::batchEmbeddingIds)
        .def("batch_embedding_into", &Reader::batchEmbeddingInto, py::arg("words"), py::arg("out"),
             py::arg("capacity_rows"), py::arg("out_on_device"))
        .def_static("count_words", &Reader::countWords)
        .def("decode_range", &Reader::decodeRange)
        .def("decode_ids_into", &Reader::decodeIdsInto, py::arg("ids"), py::arg("ids_on_device"), py::arg("n"),
             py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0)
        .def("decode_range_into", &Reader::decodeRangeInto, py::arg("begin"), py::arg("end"), py::arg("out"),
             py::arg("out_on_device"), py::arg("stream") = 0)
        .def("load_stats", &Reader::loadStats);

    m.def("available_compression_strategies", &availableCompressionStrategies);
    // additive
    m.def("union_batch_embedding", &unionBatchEmbedding);
    m.def("union_batch_embedding_into", &unionBatchEmbeddingInto, py::arg("readers"), py::arg("mode"), py::arg("words"),
          py::arg("out"), py::arg("capacity_rows"), py::arg("out_on_device"));
    m.def("union_decode_ids_into", &unionDecodeIdsInto, py::arg("readers"), py::arg("mode"), py::arg("ids"),
          py::arg("ids_on_device"), py::arg("n"), py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0);
    // the marshalled form of a word collection: (UTF-8 bytes, uint64 offsets[n+1])
    m.def("pack_words", [](const py::object& words) {
        const auto q = marshal(words);
        py::array_t<uint64_t> offsets(static_cast<py::ssize_t>(q->n + 1));
        std::memcpy(offsets.mutable_data(), q->offsets, (q->n + 1) * sizeof(uint64_t));
        return py::make_tuple(py::bytes(q->data + q->offsets[0], q->offsets[q->n] - q->offsets[0]), offsets);
    });
    m.def("device_count", &memb_b200_device_count);
    m.def("launch_count", &memb_b200_launch_count);
    m.def("set_option", [](const std::string& name, int64_t value) { check(memb_b200_set_option(name.c_str(), value)); });
}
