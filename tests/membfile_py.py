"""Pure-Python writer of memb files (test infrastructure).

An independent second implementation of the wire format — written from the six
schemas (/root/reference/src/flatbuffers/*.fbs) and the public FlatBuffers binary
layout — so that the C++ parser is not only ever tested against the C++ writer,
and so that tests can hand-craft files the Builder can never emit (a 16-bit
prefix code, a one-symbol code, omitted default fields, overlapping rows, ...).
"""
import struct

import numpy as np


class _Writer:
    """Front-to-back FlatBuffers layout: parents first, children after, every
    uoffset patched once its target is placed (targets are always forward)."""

    def __init__(self):
        self.buf = bytearray()

    def align(self, n):
        while len(self.buf) % n:
            self.buf.append(0)

    def patch_offset(self, at, target):
        struct.pack_into("<I", self.buf, at, target - at)

    def vector(self, array, elem_align=4):
        array = np.ascontiguousarray(array)
        self.align(max(4, elem_align))
        pos = len(self.buf)
        self.buf += struct.pack("<I", len(array))
        self.buf += array.tobytes()
        return pos

    def string(self, data):
        self.align(4)
        pos = len(self.buf)
        self.buf += struct.pack("<I", len(data)) + bytes(data) + b"\0"
        return pos

    def table(self, fields):
        """fields: list of (id, fmt, value) with fmt in {'B','I','f'} for scalars or
        'O' for a child; a child value is a callable emitting the child and
        returning its position.  Fields given as None are omitted (= default)."""
        fields = [f for f in fields if f is not None]
        nslots = max([f[0] for f in fields], default=-1) + 1
        sizes = {"B": 1, "I": 4, "f": 4, "O": 4}
        # inline layout: soffset, then 4-byte fields, then 1-byte fields
        layout, cursor = {}, 4
        for fid, fmt, _ in sorted(fields, key=lambda f: -sizes[f[1]]):
            layout[fid] = cursor
            cursor += sizes[fmt]
        table_size = cursor
        vt_size = 4 + 2 * nslots
        # place the vtable so that the table after it is 4-aligned
        while (len(self.buf) + vt_size) % 4:
            self.buf.append(0)
        vt_pos = len(self.buf)
        self.buf += struct.pack("<HH", vt_size, table_size)
        for slot in range(nslots):
            self.buf += struct.pack("<H", layout.get(slot, 0))
        pos = len(self.buf)
        self.buf += struct.pack("<i", pos - vt_pos)
        self.buf += bytes(table_size - 4)
        children = []
        for fid, fmt, value in fields:
            at = pos + layout[fid]
            if fmt == "O":
                children.append((at, value))
            else:
                struct.pack_into("<" + fmt, self.buf, at, value)
        for at, emit in children:
            self.patch_offset(at, emit())
        return pos

    def table_vector(self, emitters):
        self.align(4)
        pos = len(self.buf)
        self.buf += struct.pack("<I", len(emitters)) + bytes(4 * len(emitters))
        for i, emit in enumerate(emitters):
            self.patch_offset(pos + 4 + 4 * i, emit())
        return pos


def _finish(storage_type, dim, emit_storage):
    w = _Writer()
    w.buf += bytes(4) + b"memb"
    root = w.table([(0, "B", storage_type), (1, "O", emit_storage(w)), (2, "I", dim) if dim else None])
    struct.pack_into("<I", w.buf, 0, root)
    return bytes(w.buf)


def canonical_codes(lengths):
    """(code, length) per symbol for lengths sorted ascending (src/prefix_code.cpp:5-22)."""
    codes, code, bits = [], 0, 0
    for length in lengths:
        while bits < length:
            bits += 1
            code <<= 1
        codes.append(code & 0xffff)
        code += 1
    return codes


def size_offsets_from_lengths(lengths):
    """Serialised length table (src/huffman_encoder.cpp:100-117): one entry pushed each
    time the code length grows by one, plus the final key count."""
    out, current = [], 0
    for i, length in enumerate(lengths):
        while current < length:
            out.append(i)
            current += 1
    out.append(len(lengths))
    return out


def pack_row(symbols, code_of, len_of):
    """MSB-first concatenation of the codes, zero padded to a byte (src/bit_stream.h:18-34)."""
    bits = []
    for s in symbols:
        code, length = code_of[s], len_of[s]
        bits.extend((code >> k) & 1 for k in range(length - 1, -1, -1))
    while len(bits) % 8:
        bits.append(0)
    if not bits:
        return b""
    return np.packbits(np.array(bits, np.uint8)).tobytes()


def pack_rows_numpy(symbols, keys, lengths, chunk_rows=10000, threads=None):
    """Vectorised pack_row for a (rows, dim) uint8 symbol matrix: returns (packed bytes of all rows
    back to back, each zero padded to a byte; uint64 byte offset of every row).  Chunks of rows
    are packed on a thread pool (NumPy releases the GIL in the heavy calls)."""
    import os
    from concurrent.futures import ThreadPoolExecutor

    symbols = np.ascontiguousarray(symbols, np.uint8)
    codes = canonical_codes([int(v) for v in lengths])
    aligned = np.zeros(256, ">u2")  # code left-aligned in 16 bits, big-endian: unpackbits yields it MSB-first
    len_of = np.zeros(256, np.uint8)
    for k, c, n in zip(keys, codes, lengths):
        aligned[int(k)], len_of[int(k)] = (c << (16 - int(n))) & 0xffff, n
    rows, dim = symbols.shape
    bit_index = np.arange(16, dtype=np.uint8)[None, None, :]
    starts = list(range(0, rows, chunk_rows))

    def pack(begin):
        sym = symbols[begin:begin + chunk_rows]
        lens = len_of[sym]
        bits = np.unpackbits(aligned[sym].view(np.uint8).reshape(len(sym), dim, 2), axis=-1)  # (r, dim, 16)
        valid = bit_index < lens[:, :, None]
        total = lens.sum(axis=1, dtype=np.int64)
        pad = (-total) % 8
        pad_valid = np.arange(8)[None, :] < pad[:, None]     # zero bits up to the byte boundary
        flat_valid = np.concatenate([valid.reshape(len(sym), -1), pad_valid], axis=1)
        flat_bits = np.concatenate([bits.reshape(len(sym), -1), np.zeros((len(sym), 8), np.uint8)], axis=1)
        return np.packbits(flat_bits[flat_valid]), (total + pad) // 8

    with ThreadPoolExecutor(threads or min(32, os.cpu_count() or 1)) as pool:
        done = list(pool.map(pack, starts))
    row_bytes = np.concatenate([d[1] for d in done]).astype(np.uint64) if done else np.zeros(0, np.uint64)
    offsets = np.zeros(rows, np.uint64)
    np.cumsum(row_bytes[:-1], out=offsets[1:])
    return np.concatenate([d[0] for d in done]) if done else np.zeros(0, np.uint8), offsets


def write_trained_packed(path, sorted_words, packed, value_offsets, keys, lengths, centroids, dim):
    """Trained file from rows that are already packed: sorted_words in row-id order (bytes, sorted),
    value_offsets[i] = byte offset of row i in `packed`."""
    lengths = [int(v) for v in lengths]
    blob = b"\0".join(sorted_words) + (b"\0" if sorted_words else b"")
    sizes = np.fromiter((len(w) + 1 for w in sorted_words), np.uint64, len(sorted_words))
    wo = np.zeros(len(sorted_words), np.uint32)
    if len(sorted_words):
        wo[1:] = np.cumsum(sizes[:-1])
    vo = np.asarray(value_offsets).astype(np.uint32)
    so = np.array(size_offsets_from_lengths(lengths), np.uint32)
    packed = np.ascontiguousarray(packed, np.uint8)

    def storage(w):
        def decoder():
            return w.table([(0, "O", lambda: w.vector(np.array([int(k) for k in keys], np.uint8), 1)),
                            (1, "O", lambda: w.vector(so))])

        def clusterizer():
            return w.table([(0, "O", lambda: w.vector(np.asarray(centroids, np.float32)))])

        return lambda: w.table([
            (0, "O", lambda: w.vector(wo)), (1, "O", lambda: w.vector(vo)),
            (2, "O", lambda: w.string(blob)), (3, "O", lambda: w.vector(packed, 1)),
            (4, "O", decoder), (5, "O", clusterizer)])

    data = _finish(3, dim, storage)
    with open(path, "wb") as f:
        f.write(data)


def write_trained(path, words, symbol_rows, keys, lengths, centroids, dim=None, row_order=None, tail=b""):
    """words: list of bytes in ANY order (they are sorted here like the compressor does,
    src/trained_compression.cpp:73-79); symbol_rows[i] = the dim symbols (key values)
    of words[i]; keys/lengths: the prefix code, lengths ascending.  Rows are packed in
    `row_order` (default: insertion order), offsets sorted by word.  `tail` is appended
    to packed_values (junk after the last row must not matter)."""
    keys = [int(k) for k in keys]
    lengths = [int(v) for v in lengths]
    codes = canonical_codes(lengths)
    code_of = {k: c for k, c in zip(keys, codes)}
    len_of = {k: v for k, v in zip(keys, lengths)}
    dim = len(symbol_rows[0]) if dim is None else dim
    order = list(range(len(words))) if row_order is None else list(row_order)
    packed, value_offset = bytearray(), {}
    for i in order:
        value_offset[i] = len(packed)
        packed += pack_row(symbol_rows[i], code_of, len_of)
    packed += tail
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    packed_words, word_offsets = bytearray(), []
    for i in by_word:  # text in sorted order, every word with its NUL (trained_compression.cpp:85-90)
        word_offsets.append(len(packed_words))
        packed_words += words[i] + b"\0"
    wo = np.array(word_offsets, np.uint32)
    vo = np.array([value_offset[i] for i in by_word], np.uint32)
    so = np.array(size_offsets_from_lengths(lengths), np.uint32)

    def storage(w):
        def decoder():
            return w.table([(0, "O", lambda: w.vector(np.array(keys, np.uint8), 1)), (1, "O", lambda: w.vector(so))])

        def clusterizer():
            return w.table([(0, "O", lambda: w.vector(np.asarray(centroids, np.float32)))])

        return lambda: w.table([
            (0, "O", lambda: w.vector(wo)), (1, "O", lambda: w.vector(vo)),
            (2, "O", lambda: w.string(bytes(packed_words))),
            (3, "O", lambda: w.vector(np.frombuffer(bytes(packed), np.uint8), 1)),
            (4, "O", decoder), (5, "O", clusterizer)])

    data = _finish(3, dim, storage)
    with open(path, "wb") as f:
        f.write(data)
    return [words[i] for i in by_word]


def write_uniform(path, words, mins, maxs, values, levels, omit_defaults=True):
    """values: (n, dim) uint8; nodes sorted by word like CreateVectorOfSortedTables.
    A scalar equal to its default (0.0) is omitted like the flatbuffers builder does."""
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    values = np.asarray(values, np.uint8)

    def storage(w):
        def node(i):
            def vec():
                mn, mx = float(mins[i]), float(maxs[i])
                return w.table([
                    None if (omit_defaults and mn == 0.0) else (0, "f", mn),
                    None if (omit_defaults and mx == 0.0) else (1, "f", mx),
                    (2, "O", lambda: w.vector(values[i], 1))])

            return lambda: w.table([(0, "O", lambda: w.string(words[i])), (1, "O", vec)])

        return lambda: w.table([(0, "O", lambda: w.table_vector([node(i) for i in by_word])),
                                None if (omit_defaults and levels == 0) else (1, "B", levels)])

    data = _finish(2, values.shape[1], storage)
    with open(path, "wb") as f:
        f.write(data)
    return [words[i] for i in by_word]


def write_full(path, words, matrix):
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    matrix = np.asarray(matrix, np.float32)

    def storage(w):
        def node(i):
            return lambda: w.table([(0, "O", lambda: w.string(words[i])), (1, "O", lambda: w.vector(matrix[i]))])

        return lambda: w.table([(0, "O", lambda: w.table_vector([node(i) for i in by_word]))])

    data = _finish(1, matrix.shape[1], storage)
    with open(path, "wb") as f:
        f.write(data)
    return [words[i] for i in by_word]


# ---- the layout a real FlatBufferBuilder emits (flatc 1.9.0, what memb_converter produces) ----
class _BackWriter:
    """Back-to-front FlatBuffers construction like flatbuffers::FlatBufferBuilder: the buffer
    grows towards lower addresses, children are finished before their parents (so they sit at
    HIGHER addresses), fields are added in the order given (the generated Create* helpers add
    them by descending size), scalars equal to their default are dropped, and identical
    vtables are shared -- every table after the first with the same shape reaches its vtable
    through a NEGATIVE soffset.  Offsets handed around are distances from the END of the
    buffer, as in the C++ builder."""

    def __init__(self):
        self.buf = bytearray()
        self.vtables = []  # (bytes, offset from end)
        self.minalign = 1

    def size(self):
        return len(self.buf)

    def _push(self, data):
        self.buf[0:0] = data

    def prealign(self, length, alignment):
        self.minalign = max(self.minalign, alignment)
        while (self.size() + length) % alignment:
            self._push(b"\0")

    def _refer_to(self, off):
        self.prealign(0, 4)
        return self.size() - off + 4

    def vector(self, array, elem_size):
        data = np.ascontiguousarray(array).tobytes()
        self.prealign(len(data), 4)
        self.prealign(len(data), elem_size)
        self._push(data)
        self._push(struct.pack("<I", len(data) // elem_size))
        return self.size()

    def string(self, data):
        self.prealign(len(data) + 1, 4)
        self._push(bytes(data) + b"\0")
        self._push(struct.pack("<I", len(data)))
        return self.size()

    def offset_vector(self, offsets):
        self.prealign(4 * len(offsets), 4)
        for off in reversed(offsets):
            self._push(struct.pack("<I", self._refer_to(off)))
        self._push(struct.pack("<I", len(offsets)))
        return self.size()

    def table(self, fields):
        """fields: (id, fmt, value) in the order a generated Create* adds them; fmt 'O': value is an
        offset from the end (0 = absent); scalars: None = equal to the default, omitted."""
        start = self.size()
        placed = {}
        for fid, fmt, value in fields:
            if value is None or (fmt == "O" and not value):
                continue
            size = {"B": 1, "I": 4, "f": 4, "O": 4}[fmt]
            self.prealign(0, size)
            if fmt == "O":
                self._push(struct.pack("<I", self._refer_to(value)))
            else:
                self._push(struct.pack("<" + fmt, value))
            placed[fid] = self.size()
        self.prealign(0, 4)
        self._push(bytes(4))  # soffset, patched below
        table_off = self.size()
        nslots = max(placed, default=-1) + 1
        vt = struct.pack("<HH", 4 + 2 * nslots, table_off - start)
        for slot in range(nslots):
            vt += struct.pack("<H", table_off - placed[slot] if slot in placed else 0)
        for known, off in self.vtables:
            if known == vt:
                vt_off = off
                break
        else:
            self._push(vt)
            vt_off = self.size()
            self.vtables.append((vt, vt_off))
        struct.pack_into("<i", self.buf, self.size() - table_off, vt_off - table_off)
        return table_off

    def finish(self, root_off, identifier=b"memb"):
        self.prealign(4 + len(identifier), self.minalign)
        self._push(identifier)
        self._push(struct.pack("<I", self._refer_to(root_off)))
        return bytes(self.buf)


def _finish_fbb(w, storage_type, storage_off, dim):
    # Index: dim (u32), storage (offset), storage_type (u8) -- descending size like CreateIndex
    root = w.table([(2, "I", dim or None), (1, "O", storage_off), (0, "B", storage_type)])
    return w.finish(root)


def write_trained_fbb(path, words, symbol_rows, keys, lengths, centroids):
    """write_trained in FlatBufferBuilder order (src/trained_compression.cpp:92-100: the
    vectors are created in the compressor, the table last)."""
    keys = [int(k) for k in keys]
    lengths = [int(v) for v in lengths]
    codes = canonical_codes(lengths)
    code_of = {k: c for k, c in zip(keys, codes)}
    len_of = {k: v for k, v in zip(keys, lengths)}
    packed, value_offset = bytearray(), []
    for row in symbol_rows:
        value_offset.append(len(packed))
        packed += pack_row(row, code_of, len_of)
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    packed_words, word_offsets = bytearray(), []
    for i in by_word:
        word_offsets.append(len(packed_words))
        packed_words += words[i] + b"\0"
    w = _BackWriter()
    # decoder and clusterizer first (children), as save() is called before CreateTrained
    dec_keys = w.vector(np.array(keys, np.uint8), 1)
    dec_sizes = w.vector(np.array(size_offsets_from_lengths(lengths), np.uint32), 4)
    decoder = w.table([(0, "O", dec_keys), (1, "O", dec_sizes)])
    cents = w.vector(np.asarray(centroids, np.float32), 4)
    clusterizer = w.table([(0, "O", cents)])
    wo = w.vector(np.array(word_offsets, np.uint32), 4)
    vo = w.vector(np.array([value_offset[i] for i in by_word], np.uint32), 4)
    pw = w.string(bytes(packed_words))
    pv = w.vector(np.frombuffer(bytes(packed), np.uint8), 1)
    storage = w.table([(0, "O", wo), (1, "O", vo), (2, "O", pw), (3, "O", pv), (4, "O", decoder), (5, "O", clusterizer)])
    dim = len(symbol_rows[0]) if len(symbol_rows) else 0
    with open(path, "wb") as f:
        f.write(_finish_fbb(w, 3, storage, dim))
    return [words[i] for i in by_word]


def write_uniform_fbb(path, words, mins, maxs, values, levels):
    """write_uniform in FlatBufferBuilder order: every node is finished when its word is added
    (src/uniform_compression.cpp:10-34), the sorted vector and the table at finalize."""
    values = np.asarray(values, np.uint8)
    w = _BackWriter()
    nodes = {}
    for i in range(len(words)):  # insertion order
        word = w.string(words[i])
        vals = w.vector(values[i], 1)
        mn, mx = float(mins[i]), float(maxs[i])
        vec = w.table([(0, "f", mn if mn != 0.0 else None), (1, "f", mx if mx != 0.0 else None), (2, "O", vals)])
        nodes[i] = w.table([(0, "O", word), (1, "O", vec)])
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    vector = w.offset_vector([nodes[i] for i in by_word])
    storage = w.table([(0, "O", vector), (1, "B", levels or None)])
    with open(path, "wb") as f:
        f.write(_finish_fbb(w, 2, storage, values.shape[1]))
    return [words[i] for i in by_word]


def write_full_fbb(path, words, matrix):
    matrix = np.asarray(matrix, np.float32)
    w = _BackWriter()
    nodes = {}
    for i in range(len(words)):
        word = w.string(words[i])
        vals = w.vector(matrix[i], 4)
        nodes[i] = w.table([(0, "O", word), (1, "O", vals)])
    by_word = sorted(range(len(words)), key=lambda i: words[i])
    vector = w.offset_vector([nodes[i] for i in by_word])
    storage = w.table([(0, "O", vector)])
    with open(path, "wb") as f:
        f.write(_finish_fbb(w, 1, storage, matrix.shape[1]))
    return [words[i] for i in by_word]
