"""Test helpers: ctypes views of the oracle (the checker) and of the product's C ABI,
plus synthetic-file builders.  Nothing here reads /root/reference at run time."""
import ctypes
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

c_void_p, c_char_p, c_u64, c_u32, c_int = (ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_uint32,
                                           ctypes.c_int)


def marshal(words):
    """list[str|bytes] -> (blob bytes, uint64 offsets[n+1]) like the binding does."""
    encoded = [w.encode("utf-8") if isinstance(w, str) else bytes(w) for w in words]
    offsets = np.zeros(len(encoded) + 1, np.uint64)
    if encoded:
        offsets[1:] = np.cumsum([len(e) for e in encoded])
    return b"".join(encoded), offsets


def bits_equal(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


# ------------------------------------------------------------------ oracle ----
class OracleLib:
    def __init__(self, path):
        lib = ctypes.CDLL(path)
        self.lib = lib
        lib.oracle_open.restype = c_void_p
        lib.oracle_open.argtypes = [c_char_p, c_u64]
        lib.oracle_last_error.restype = c_char_p
        lib.oracle_close.argtypes = [c_void_p]
        lib.oracle_dim.argtypes = [c_void_p]
        lib.oracle_dim.restype = c_u32
        lib.oracle_rows.argtypes = [c_void_p]
        lib.oracle_rows.restype = c_u64
        lib.oracle_storage.argtypes = [c_void_p]
        lib.oracle_threads.argtypes = [c_void_p]
        lib.oracle_threads.restype = c_u64
        lib.oracle_key.argtypes = [c_void_p, c_u64]
        lib.oracle_key.restype = c_char_p
        lib.oracle_set_table_bits.argtypes = [c_void_p, c_int]
        lib.oracle_word_embedding.argtypes = [c_void_p, c_char_p, c_void_p]
        lib.oracle_batch_embedding.argtypes = [c_void_p, c_char_p, c_void_p, c_u64, c_void_p]
        lib.oracle_resolve.argtypes = [c_void_p, c_char_p, c_void_p, c_u64, c_void_p]
        lib.oracle_batch_ids.argtypes = [c_void_p, c_void_p, c_u64, c_void_p]
        lib.oracle_decode_range.argtypes = [c_void_p, c_u64, c_u64, c_void_p]
        lib.oracle_decode_symbols.argtypes = [c_void_p, c_u64, c_void_p]
        lib.oracle_codebook.argtypes = [c_void_p, c_void_p, c_u32]
        lib.oracle_codebook.restype = c_u32
        lib.oracle_prefix_code.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_u32]
        lib.oracle_prefix_code.restype = c_u32
        lib.oracle_bitstream_pack.argtypes = [c_void_p, c_void_p, c_u64, c_void_p, c_u64]
        lib.oracle_bitstream_pack.restype = c_u64
        lib.oracle_decode_stream.argtypes = [c_void_p, c_u32, c_void_p, c_u32, c_int, c_void_p, c_u64, c_u64,
                                             c_void_p]
        lib.oracle_canonical_codes.argtypes = [c_void_p, c_u32, c_void_p, c_u32, c_void_p, c_void_p]
        lib.oracle_uses_reference_codec.restype = c_int

    @property
    def uses_reference_codec(self):
        return bool(self.lib.oracle_uses_reference_codec())

    def open(self, path, num_threads=1):
        return OracleReader(self, path, num_threads)

    # codec-level
    def bitstream_pack(self, codes, lens):
        codes = np.ascontiguousarray(codes, np.uint16)
        lens = np.ascontiguousarray(lens, np.uint8)
        cap = int(lens.astype(np.int64).sum() // 8 + 2)
        out = np.zeros(cap, np.uint8)
        n = self.lib.oracle_bitstream_pack(codes.ctypes.data, lens.ctypes.data, len(codes), out.ctypes.data, cap)
        return out[:n].copy()

    def canonical_codes(self, keys, size_offsets):
        keys = np.ascontiguousarray(keys, np.uint8)
        so = np.ascontiguousarray(size_offsets, np.uint32)
        codes = np.zeros(len(keys), np.uint16)
        lens = np.zeros(len(keys), np.uint8)
        self.lib.oracle_canonical_codes(keys.ctypes.data, len(keys), so.ctypes.data, len(so), codes.ctypes.data,
                                        lens.ctypes.data)
        return codes, lens

    def decode_stream(self, keys, size_offsets, table_bits, data, count):
        keys = np.ascontiguousarray(keys, np.uint8)
        so = np.ascontiguousarray(size_offsets, np.uint32)
        data = np.ascontiguousarray(data, np.uint8)
        out = np.zeros(count, np.uint8)
        self.lib.oracle_decode_stream(keys.ctypes.data, len(keys), so.ctypes.data, len(so), table_bits,
                                      data.ctypes.data, len(data), count, out.ctypes.data)
        return out


class OracleReader:
    def __init__(self, oracle, path, num_threads=1):
        self.o = oracle
        self.lib = oracle.lib
        self.h = self.lib.oracle_open(os.fsencode(str(path)), num_threads)
        if not self.h:
            raise RuntimeError(self.lib.oracle_last_error().decode())
        self.dim = self.lib.oracle_dim(self.h)
        self.rows = self.lib.oracle_rows(self.h)
        self.storage = self.lib.oracle_storage(self.h)

    def close(self):
        if self.h:
            self.lib.oracle_close(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def keys(self):
        return [self.lib.oracle_key(self.h, r) for r in range(self.rows)]

    def set_table_bits(self, bits):
        self.lib.oracle_set_table_bits(self.h, bits)

    def word_embedding(self, word):
        out = np.empty(self.dim, np.float32)
        w = word.encode("utf-8") if isinstance(word, str) else word
        self.lib.oracle_word_embedding(self.h, w, out.ctypes.data)
        return out

    def batch_embedding(self, words):
        blob, offsets = marshal(words)
        out = np.empty((len(words), self.dim), np.float32)
        self.lib.oracle_batch_embedding(self.h, blob, offsets.ctypes.data, len(words), out.ctypes.data)
        return out

    def resolve(self, words):
        blob, offsets = marshal(words)
        ids = np.empty(len(words), np.int32)
        self.lib.oracle_resolve(self.h, blob, offsets.ctypes.data, len(words), ids.ctypes.data)
        return ids

    def batch_ids(self, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        out = np.empty((len(ids), self.dim), np.float32)
        self.lib.oracle_batch_ids(self.h, ids.ctypes.data, len(ids), out.ctypes.data)
        return out

    def decode_range(self, begin, end):
        out = np.empty((end - begin, self.dim), np.float32)
        self.lib.oracle_decode_range(self.h, begin, end, out.ctypes.data)
        return out

    def decode_symbols(self, row):
        out = np.empty(self.dim, np.uint8)
        assert self.lib.oracle_decode_symbols(self.h, row, out.ctypes.data) == 0
        return out

    def codebook(self):
        out = np.empty(256, np.float32)
        n = self.lib.oracle_codebook(self.h, out.ctypes.data, 256)
        return out[:n].copy()

    def prefix_code(self):
        keys = np.zeros(256, np.uint8)
        codes = np.zeros(256, np.uint16)
        lens = np.zeros(256, np.uint8)
        n = self.lib.oracle_prefix_code(self.h, keys.ctypes.data, codes.ctypes.data, lens.ctypes.data, 256)
        return keys[:n].copy(), codes[:n].copy(), lens[:n].copy()


def load_oracle():
    return OracleLib(os.path.join(ROOT, "oracle", "liboracle.so"))


def load_ref_oracle():
    """The oracle whose Huffman decode IS the reference's header (oracle/_ref, built only
    where /root/reference exists; the prebuilt .so travels to the GPU box)."""
    path = os.path.join(ROOT, "oracle", "_ref", "liboracle_ref.so")
    return OracleLib(path) if os.path.exists(path) else None


class RefCodec:
    """C entry points over the reference's own codec sources (oracle/ref_glue.cpp)."""

    def __init__(self, path):
        lib = ctypes.CDLL(path)
        self.lib = lib
        lib.ref_bitstream_pack.argtypes = [c_void_p, c_void_p, c_u64, c_void_p, c_u64]
        lib.ref_bitstream_pack.restype = c_u64
        lib.ref_canonical_codes.argtypes = [c_void_p, c_u32, c_void_p, c_u32, c_void_p, c_void_p]
        lib.ref_decode_stream.argtypes = [c_void_p, c_u32, c_void_p, c_u32, c_int, c_void_p, c_u64, c_u64, c_void_p]

    def bitstream_pack(self, codes, lens):
        codes = np.ascontiguousarray(codes, np.uint16)
        lens = np.ascontiguousarray(lens, np.uint8)
        cap = int(lens.astype(np.int64).sum() // 8 + 2)
        out = np.zeros(cap, np.uint8)
        n = self.lib.ref_bitstream_pack(codes.ctypes.data, lens.ctypes.data, len(codes), out.ctypes.data, cap)
        return out[:n].copy()

    def canonical_codes(self, keys, size_offsets):
        keys = np.ascontiguousarray(keys, np.uint8)
        so = np.ascontiguousarray(size_offsets, np.uint32)
        codes = np.zeros(len(keys), np.uint16)
        lens = np.zeros(len(keys), np.uint8)
        self.lib.ref_canonical_codes(keys.ctypes.data, len(keys), so.ctypes.data, len(so), codes.ctypes.data,
                                     lens.ctypes.data)
        return codes, lens

    def decode_stream(self, keys, size_offsets, table_bits, data, count):
        keys = np.ascontiguousarray(keys, np.uint8)
        so = np.ascontiguousarray(size_offsets, np.uint32)
        data = np.ascontiguousarray(data, np.uint8)
        out = np.zeros(count, np.uint8)
        self.lib.ref_decode_stream(keys.ctypes.data, len(keys), so.ctypes.data, len(so), table_bits,
                                   data.ctypes.data, len(data), count, out.ctypes.data)
        return out


def load_ref_codec():
    path = os.path.join(ROOT, "oracle", "_ref", "libmemb_ref_codec.so")
    return RefCodec(path) if os.path.exists(path) else None


# ------------------------------------------------------------------- C ABI ----
HEADER = os.path.join(ROOT, "include", "memb_b200.h")
LIBRARY = os.path.join(ROOT, "memb_b200", "libmemb_b200.so")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(memb_b200_[a-z0-9_]+)\s*\(", text)))


class LoadStats(ctypes.Structure):
    _fields_ = [("parse_ms", ctypes.c_double), ("upload_ms", ctypes.c_double), ("prepare_ms", ctypes.c_double),
                ("device_bytes", c_u64), ("payload_bytes", c_u64), ("max_code_len", c_u32), ("lut_bits", c_u32),
                ("lut_replicas", c_u32), ("codebook_size", c_u32)]


class CAbi:
    """Plain ctypes binding of include/memb_b200.h — what any FFI would do."""
    OK, ERR_IO, ERR_FORMAT, ERR_UNSUPPORTED, ERR_CUDA, ERR_ARGUMENT, ERR_NO_DEVICE, ERR_BUILDER = range(8)
    HOST, DEVICE = 0, 1
    AVERAGE, CONCATENATE = 0, 1

    def __init__(self):
        lib = ctypes.CDLL(LIBRARY)
        self.lib = lib
        P = ctypes.POINTER
        lib.memb_b200_last_error.restype = c_char_p
        lib.memb_b200_open.argtypes = [c_char_p, c_int, ctypes.c_uint, P(c_void_p)]
        lib.memb_b200_close.argtypes = [c_void_p]
        lib.memb_b200_dim.argtypes = [c_void_p]
        lib.memb_b200_dim.restype = c_u32
        lib.memb_b200_rows.argtypes = [c_void_p]
        lib.memb_b200_rows.restype = c_u64
        lib.memb_b200_storage_type.argtypes = [c_void_p]
        lib.memb_b200_device.argtypes = [c_void_p]
        lib.memb_b200_get_load_stats.argtypes = [c_void_p, P(LoadStats)]
        lib.memb_b200_keys.argtypes = [c_void_p, P(c_void_p), P(c_void_p), P(c_u64)]
        lib.memb_b200_resolve.argtypes = [c_void_p, c_char_p, c_void_p, c_u64, c_void_p]
        lib.memb_b200_lookup_words.argtypes = [c_void_p, c_char_p, c_void_p, c_u64, c_void_p]
        lib.memb_b200_decode_ids.argtypes = [c_void_p, c_void_p, c_int, c_u64, c_void_p, c_int, c_void_p]
        lib.memb_b200_decode_range.argtypes = [c_void_p, c_u64, c_u64, c_void_p, c_int, c_void_p]
        lib.memb_b200_union_lookup_words.argtypes = [P(c_void_p), c_u32, c_int, c_char_p, c_void_p, c_u64, c_void_p]
        lib.memb_b200_union_decode_ids.argtypes = [P(c_void_p), c_u32, c_int, c_void_p, c_int, c_u64, c_void_p,
                                                   c_int, c_void_p]
        lib.memb_b200_host_alloc.argtypes = [ctypes.c_size_t, P(c_void_p)]
        lib.memb_b200_host_free.argtypes = [c_void_p]
        lib.memb_b200_launch_count.restype = c_u64
        lib.memb_b200_set_option.argtypes = [c_char_p, ctypes.c_int64]
        lib.memb_b200_builder_create.argtypes = [c_u64, c_char_p, c_u64, P(c_void_p)]
        lib.memb_b200_builder_destroy.argtypes = [c_void_p]
        lib.memb_b200_builder_add_word.argtypes = [c_void_p, c_char_p, ctypes.c_size_t, c_void_p, c_u64]
        lib.memb_b200_builder_add_words.argtypes = [c_void_p, c_char_p, c_void_p, c_u64, c_void_p]
        lib.memb_b200_builder_save.argtypes = [c_void_p, c_char_p, ctypes.c_uint]
        lib.memb_b200_available_compression_strategies.restype = ctypes.POINTER(ctypes.c_char)

    def error(self):
        return self.lib.memb_b200_last_error().decode()

    def check(self, status):
        if status != 0:
            raise RuntimeError("status %d: %s" % (status, self.error()))

    def open(self, path, device=0, num_threads=0):
        handle = c_void_p()
        status = self.lib.memb_b200_open(os.fsencode(str(path)), device, num_threads, ctypes.byref(handle))
        return status, handle

    def open_checked(self, path, device=0):
        status, handle = self.open(path, device)
        self.check(status)
        return CAbiReader(self, handle)


class CAbiReader:
    def __init__(self, abi, handle):
        self.abi, self.lib, self.h = abi, abi.lib, handle
        self.dim = self.lib.memb_b200_dim(handle)
        self.rows = self.lib.memb_b200_rows(handle)

    def close(self):
        if self.h:
            self.lib.memb_b200_close(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def keys(self):
        words, offsets, rows = c_void_p(), c_void_p(), c_u64()
        self.abi.check(self.lib.memb_b200_keys(self.h, ctypes.byref(words), ctypes.byref(offsets), ctypes.byref(rows)))
        n = rows.value
        off = np.ctypeslib.as_array(ctypes.cast(offsets, ctypes.POINTER(c_u32)), shape=(n + 1,)) if n else []
        return [ctypes.string_at(words.value + int(off[r])) for r in range(n)]

    def resolve(self, words):
        blob, offsets = marshal(words)
        ids = np.empty(len(words), np.int32)
        self.abi.check(self.lib.memb_b200_resolve(self.h, blob, offsets.ctypes.data, len(words), ids.ctypes.data))
        return ids

    def lookup_words(self, words):
        blob, offsets = marshal(words)
        out = np.empty((len(words), self.dim), np.float32)
        self.abi.check(self.lib.memb_b200_lookup_words(self.h, blob, offsets.ctypes.data, len(words), out.ctypes.data))
        return out

    def decode_ids(self, ids):
        ids = np.ascontiguousarray(ids, np.int32)
        out = np.empty((len(ids), self.dim), np.float32)
        self.abi.check(self.lib.memb_b200_decode_ids(self.h, ids.ctypes.data, 0, len(ids), out.ctypes.data, 0, None))
        return out

    def decode_range(self, begin, end):
        out = np.empty((end - begin, self.dim), np.float32)
        self.abi.check(self.lib.memb_b200_decode_range(self.h, begin, end, out.ctypes.data, 0, None))
        return out

    def load_stats(self):
        stats = LoadStats()
        self.abi.check(self.lib.memb_b200_get_load_stats(self.h, ctypes.byref(stats)))
        return stats


def union_lookup_words(abi, readers, mode, words, union_dim):
    blob, offsets = marshal(words)
    handles = (c_void_p * len(readers))(*[r.h for r in readers])
    out = np.empty((len(words), union_dim), np.float32)
    abi.check(abi.lib.memb_b200_union_lookup_words(handles, len(readers), mode, blob, offsets.ctypes.data, len(words),
                                                   out.ctypes.data))
    return out


def union_decode_ids(abi, readers, mode, ids, union_dim):
    ids = np.ascontiguousarray(ids, np.int32)  # (R, n) reader-major
    n = ids.shape[1]
    handles = (c_void_p * len(readers))(*[r.h for r in readers])
    out = np.empty((n, union_dim), np.float32)
    abi.check(abi.lib.memb_b200_union_decode_ids(handles, len(readers), mode, ids.ctypes.data, 0, n, out.ctypes.data,
                                                 0, None))
    return out


# ------------------------------------------------------------ synthetic files ----
def synthetic_words(count, seed=0, utf8_fraction=0.01):
    """Unique words, 2-12 bytes, ~1 % multi-byte UTF-8, in shuffled (non-sorted)
    insertion order."""
    rng = np.random.default_rng(seed)
    words = []
    for i in range(count):
        base = np.base_repr(i, 36).lower()
        if rng.random() < utf8_fraction:
            words.append("é" + base + "ж")
        else:
            words.append("w" + base)
    order = rng.permutation(count)
    return [words[i] for i in order]


def build_file(path, storage, bits, words, matrix):
    import memb_b200

    builder = memb_b200.Builder(matrix.shape[1], storage, bits)
    builder.add_words(list(words), np.ascontiguousarray(matrix, np.float32))
    builder.save(str(path))
    return str(path)
