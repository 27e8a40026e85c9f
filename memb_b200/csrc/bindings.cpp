// Extension module `_memb`: same classes, method names and argument meaning as
// the reference binding (python/memb_bindings.cpp:11-72), implemented purely on
// the C ABI of include/memb_b200.h.  Additive members (not in the reference) are
// grouped at the end of each class.
#include "../../include/memb_b200.h"

#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstring>
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

// list[str] -> (blob, offsets).  Python str -> UTF-8 like pybind11's
// std::string caster used by the reference (python/memb_bindings.cpp:54-56).
struct Marshalled {
    std::string blob;
    std::vector<uint64_t> offsets;
};

Marshalled marshal(const py::list& words)
{
    Marshalled m;
    const size_t n = words.size();
    m.offsets.reserve(n + 1);
    m.offsets.push_back(0);
    m.blob.reserve(n * 8);
    for (const py::handle& item : words) {
        if (PyUnicode_Check(item.ptr())) {
            Py_ssize_t size = 0;
            const char* utf8 = PyUnicode_AsUTF8AndSize(item.ptr(), &size);
            if (!utf8) {
                throw py::error_already_set();
            }
            m.blob.append(utf8, static_cast<size_t>(size));
        } else if (PyBytes_Check(item.ptr())) {
            m.blob.append(PyBytes_AS_STRING(item.ptr()), static_cast<size_t>(PyBytes_GET_SIZE(item.ptr())));
        } else {
            throw py::type_error("batch_embedding(): words must be str");
        }
        m.offsets.push_back(m.blob.size());
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

    py::array_t<float> batchEmbedding(const py::list& words) const
    {
        const Marshalled m = marshal(words);
        const size_t n = m.offsets.size() - 1;
        py::array_t<float> result({static_cast<py::ssize_t>(n), static_cast<py::ssize_t>(dim())});
        float* out = result.mutable_data();
        {
            py::gil_scoped_release release;
            check(memb_b200_lookup_words(impl_, m.blob.data(), m.offsets.data(), n, out));
        }
        return result;
    }

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
    py::array_t<int32_t> resolve(const py::list& words) const
    {
        const Marshalled m = marshal(words);
        const size_t n = m.offsets.size() - 1;
        py::array_t<int32_t> ids(static_cast<py::ssize_t>(n));
        check(memb_b200_resolve(impl_, m.blob.data(), m.offsets.data(), n, ids.mutable_data()));
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

py::array_t<float> unionBatchEmbedding(const std::vector<Reader*>& readers, const std::string& mode,
                                       const py::list& words)
{
    const memb_b200_union_mode m = parseMode(mode);
    std::vector<memb_b200_reader*> handles;
    size_t dim = 0;
    for (const Reader* r : readers) {
        handles.push_back(r->handle());
        dim = (m == MEMB_B200_AVERAGE) ? r->dim() : dim + r->dim();
    }
    const Marshalled q = marshal(words);
    const size_t n = q.offsets.size() - 1;
    py::array_t<float> result({static_cast<py::ssize_t>(n), static_cast<py::ssize_t>(dim)});
    float* out = result.mutable_data();
    {
        py::gil_scoped_release release;
        check(memb_b200_union_lookup_words(handles.data(), static_cast<uint32_t>(handles.size()), m, q.blob.data(),
                                           q.offsets.data(), n, out));
    }
    return result;
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
        const Marshalled m = marshal(words);
        // a width mismatch is reported per word by the builder itself
        if (words.size() && static_cast<uint64_t>(info.shape[1]) != dim_) {
            check(memb_b200_builder_add_word(impl_, m.blob.data(), m.offsets[1], static_cast<const float*>(info.ptr),
                                             static_cast<uint64_t>(info.shape[1])));
            return;
        }
        check(memb_b200_builder_add_words(impl_, m.blob.data(), m.offsets.data(), words.size(),
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
        .def("decode_range", &Reader::decodeRange)
        .def("decode_ids_into", &Reader::decodeIdsInto, py::arg("ids"), py::arg("ids_on_device"), py::arg("n"),
             py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0)
        .def("decode_range_into", &Reader::decodeRangeInto, py::arg("begin"), py::arg("end"), py::arg("out"),
             py::arg("out_on_device"), py::arg("stream") = 0)
        .def("load_stats", &Reader::loadStats);

    m.def("available_compression_strategies", &availableCompressionStrategies);
    // additive
    m.def("union_batch_embedding", &unionBatchEmbedding);
    m.def("union_decode_ids_into", &unionDecodeIdsInto, py::arg("readers"), py::arg("mode"), py::arg("ids"),
          py::arg("ids_on_device"), py::arg("n"), py::arg("out"), py::arg("out_on_device"), py::arg("stream") = 0);
    m.def("device_count", &memb_b200_device_count);
    m.def("launch_count", &memb_b200_launch_count);
    m.def("set_option", [](const std::string& name, int64_t value) { check(memb_b200_set_option(name.c_str(), value)); });
}
