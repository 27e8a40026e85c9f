// FlatBuffers *binary* layer for the memb file format: a bounds-checked reader
// ("View") and a forward-patching writer ("Writer").
//
// There is no flatc / flatbuffers.h in this build; everything here follows the
// public FlatBuffers wire layout (little-endian; table = soffset to a vtable of
// u16 slots; vectors/strings = u32 length + payload; uoffsets point forward)
// and the six schemas of the reference:
//   /root/reference/src/flatbuffers/embeddings.fbs:7-19
//   /root/reference/src/flatbuffers/trained_compression.fbs:6-13
//   /root/reference/src/flatbuffers/uniform_compression.fbs:3-17
//   /root/reference/src/flatbuffers/full_compression.fbs:3-10
//   /root/reference/src/flatbuffers/huffman_decoder.fbs:3-6
//   /root/reference/src/flatbuffers/kmeans.fbs:3-5
// The reference never runs a flatbuffers Verifier (src/reader.cpp:104-111); a
// corrupt file is UB there.  Here every access is range-checked because the
// offsets end up driving device reads.
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace memb_b200 {
namespace fb {

struct FormatError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// Field ids (vtable slot = 4 + 2*id), straight from the schemas above.
namespace ids {
// table Index { storage: Storage (union => _type field first); dim: uint; }
constexpr unsigned INDEX_STORAGE_TYPE = 0, INDEX_STORAGE = 1, INDEX_DIM = 2;
// union Storage { Full, Uniform, Trained }  (NONE = 0)
constexpr uint8_t STORAGE_NONE = 0, STORAGE_FULL = 1, STORAGE_UNIFORM = 2, STORAGE_TRAINED = 3;
constexpr unsigned TRAINED_WORD_OFFSETS = 0, TRAINED_VALUE_OFFSETS = 1, TRAINED_PACKED_WORDS = 2,
                   TRAINED_PACKED_VALUES = 3, TRAINED_DECODER = 4, TRAINED_CLUSTERIZER = 5;
constexpr unsigned HUFFMAN_KEYS = 0, HUFFMAN_SIZE_OFFSETS = 1;
constexpr unsigned KMEANS_CENTROIDS = 0;
constexpr unsigned UNIFORM_NODES = 0, UNIFORM_LEVELS = 1;
constexpr unsigned UNODE_WORD = 0, UNODE_VALUES = 1;
constexpr unsigned UVEC_MIN = 0, UVEC_MAX = 1, UVEC_VALUES = 2;
constexpr unsigned FULL_NODES = 0;
constexpr unsigned FNODE_WORD = 0, FNODE_VALUES = 1;
}  // namespace ids

// ---------------------------------------------------------------- reader ----
class View {
public:
    struct Span {
        size_t pos = 0;     // absolute position of element 0
        uint32_t len = 0;   // element count
        bool present = false;
    };

    View(const uint8_t* data, size_t size) : data_(data), size_(size) {}

    const uint8_t* data() const { return data_; }
    size_t size() const { return size_; }

    template <class T>
    T read(size_t pos) const
    {
        if (pos > size_ || size_ - pos < sizeof(T)) {
            throw FormatError("memb file: read past end of buffer");
        }
        T v;
        std::memcpy(&v, data_ + pos, sizeof(T));
        return v;
    }

    // Follow the uoffset stored at `pos`.
    size_t follow(size_t pos) const
    {
        const uint32_t off = read<uint32_t>(pos);
        const size_t target = pos + static_cast<size_t>(off);
        if (off == 0 || target >= size_) {
            throw FormatError("memb file: offset points outside the buffer");
        }
        return target;
    }

    size_t root() const
    {
        const size_t t = static_cast<size_t>(read<uint32_t>(0));
        checkTable(t);
        return t;
    }

    // Absolute position of field `id` inside table `table`, or 0 when the
    // vtable slot is absent/zero (=> schema default; the flatbuffers builder
    // omits scalars equal to their default).
    size_t field(size_t table, unsigned id) const
    {
        const int32_t soff = read<int32_t>(table);
        const int64_t vt = static_cast<int64_t>(table) - static_cast<int64_t>(soff);
        if (vt < 0 || static_cast<uint64_t>(vt) + 4 > size_) {
            throw FormatError("memb file: vtable outside the buffer");
        }
        const uint16_t vtSize = read<uint16_t>(static_cast<size_t>(vt));
        const size_t slot = 4 + 2 * static_cast<size_t>(id);
        if (slot + 2 > vtSize) {
            return 0;
        }
        const uint16_t rel = read<uint16_t>(static_cast<size_t>(vt) + slot);
        if (rel == 0) {
            return 0;
        }
        const size_t pos = table + rel;
        if (pos >= size_) {
            throw FormatError("memb file: table field outside the buffer");
        }
        return pos;
    }

    template <class T>
    T scalar(size_t table, unsigned id, T dflt) const
    {
        const size_t pos = field(table, id);
        return pos ? read<T>(pos) : dflt;
    }

    // Sub-table referenced by an offset field; 0 when absent.
    size_t table(size_t parent, unsigned id) const
    {
        const size_t pos = field(parent, id);
        if (!pos) {
            return 0;
        }
        const size_t t = follow(pos);
        checkTable(t);
        return t;
    }

    // Vector (or string: elemSize 1) referenced by an offset field.
    Span vector(size_t parent, unsigned id, size_t elemSize) const
    {
        Span s;
        const size_t pos = field(parent, id);
        if (!pos) {
            return s;
        }
        const size_t v = follow(pos);
        s.len = read<uint32_t>(v);
        s.pos = v + 4;
        s.present = true;
        const uint64_t bytes = static_cast<uint64_t>(s.len) * elemSize;
        if (s.pos > size_ || bytes > size_ - s.pos) {
            throw FormatError("memb file: vector extends past end of buffer");
        }
        return s;
    }

    // Element `i` of a vector of tables.
    size_t tableAt(const Span& v, uint32_t i) const
    {
        const size_t t = follow(v.pos + 4 * static_cast<size_t>(i));
        checkTable(t);
        return t;
    }

private:
    void checkTable(size_t t) const
    {
        if (t + 4 > size_) {
            throw FormatError("memb file: table outside the buffer");
        }
    }

    const uint8_t* data_;
    size_t size_;
};

// ---------------------------------------------------------------- writer ----
// Emits parents before children and back-patches the (forward, unsigned)
// uoffsets.  vtables are written immediately before the first table that uses
// them and may be shared by later tables (soffset is signed, so a table may
// point backwards to an earlier vtable, exactly as real FlatBuffers dedups).
class Writer {
public:
    struct FieldSpec {
        unsigned id;
        unsigned size;  // 1, 2, 4 (offsets are size 4)
    };
    struct TableLayout {
        size_t vtablePos = 0;
        uint16_t tableSize = 0;
        std::vector<uint16_t> rel;  // rel[id] = offset inside the table, 0 = absent
    };

    std::vector<uint8_t>& buffer() { return buf_; }
    size_t size() const { return buf_.size(); }

    void align(size_t a)
    {
        while (buf_.size() % a) {
            buf_.push_back(0);
        }
    }

    template <class T>
    size_t put(T v)
    {
        const size_t pos = buf_.size();
        buf_.resize(pos + sizeof(T));
        std::memcpy(buf_.data() + pos, &v, sizeof(T));
        return pos;
    }

    template <class T>
    void patch(size_t pos, T v)
    {
        std::memcpy(buf_.data() + pos, &v, sizeof(T));
    }

    // Store at `fieldPos` the uoffset that reaches `target` (target > fieldPos).
    void link(size_t fieldPos, size_t target)
    {
        if (target <= fieldPos || target - fieldPos > 0x7fffffffULL) {
            throw FormatError("memb writer: offset does not fit a forward 31-bit uoffset");
        }
        patch<uint32_t>(fieldPos, static_cast<uint32_t>(target - fieldPos));
    }

    // Writes a vtable for the listed fields (laid out in the given order after
    // the soffset, each naturally aligned).
    TableLayout vtable(const std::vector<FieldSpec>& fields)
    {
        TableLayout lay;
        unsigned maxId = 0;
        for (const auto& f : fields) {
            maxId = f.id > maxId ? f.id : maxId;
        }
        lay.rel.assign(fields.empty() ? 0 : maxId + 1, 0);
        uint16_t cursor = 4;
        for (const auto& f : fields) {
            while (cursor % f.size) {
                ++cursor;
            }
            lay.rel[f.id] = cursor;
            cursor = static_cast<uint16_t>(cursor + f.size);
        }
        while (cursor % 4) {
            ++cursor;
        }
        lay.tableSize = cursor;
        align(2);
        lay.vtablePos = buf_.size();
        put<uint16_t>(static_cast<uint16_t>(4 + 2 * lay.rel.size()));
        put<uint16_t>(lay.tableSize);
        for (uint16_t r : lay.rel) {
            put<uint16_t>(r);
        }
        return lay;
    }

    // Starts a zero-filled table instance using `lay`; returns its position.
    size_t table(const TableLayout& lay)
    {
        align(4);
        const size_t pos = buf_.size();
        buf_.resize(pos + lay.tableSize, 0);
        patch<int32_t>(pos, static_cast<int32_t>(static_cast<int64_t>(pos) -
                                                   static_cast<int64_t>(lay.vtablePos)));
        return pos;
    }

    static size_t fieldPos(const TableLayout& lay, size_t table, unsigned id)
    {
        return table + lay.rel[id];
    }

    template <class T>
    void setScalar(const TableLayout& lay, size_t table, unsigned id, T v)
    {
        patch<T>(table + lay.rel[id], v);
    }

    // u32 length + payload (+ NUL for strings).  Returns the position of the
    // length word, which is what uoffsets point at.
    size_t vector(const void* data, size_t count, size_t elemSize)
    {
        align(4);
        const size_t pos = put<uint32_t>(static_cast<uint32_t>(count));
        const size_t bytes = count * elemSize;
        const size_t at = buf_.size();
        buf_.resize(at + bytes);
        if (bytes) {
            std::memcpy(buf_.data() + at, data, bytes);
        }
        return pos;
    }

    size_t string(const char* data, size_t len)
    {
        const size_t pos = vector(data, len, 1);
        buf_.push_back(0);
        return pos;
    }

private:
    std::vector<uint8_t> buf_;
};

}  // namespace fb
}  // namespace memb_b200
