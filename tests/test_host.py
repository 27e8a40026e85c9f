"""Host logic without a GPU: the C-ABI library loads and exports every declared
symbol, error conventions, Builder checks, file format round trips through the
oracle, the Python-level API semantics that need no device."""
import ctypes
import os

import numpy as np
import pytest

import helpers
import membfile_py
from helpers import bits_equal


def test_cabi_exports_every_declared_symbol(abi):
    names = helpers.declared_symbols()
    assert len(names) >= 25
    for name in names:
        assert hasattr(abi.lib, name), name


def test_available_compression_strategies(abi):
    import memb_b200

    # order of src/compression_strategy.cpp:14-18
    assert memb_b200.available_compression_strategies() == ["full", "uniform", "trained"]
    raw = ctypes.string_at(abi.lib.memb_b200_available_compression_strategies(), 21)
    assert raw == b"full\0uniform\0trained\0"


def test_open_error_statuses(abi, workdir, reference_files):
    status, handle = abi.open(str(workdir / "nope.bin"))
    assert status == abi.ERR_IO and not handle  # missingFileThrows, src/tests.cpp:134-139
    bad = workdir / "bad10.bin"
    bad.write_bytes(b"0123456789")
    status, handle = abi.open(str(bad))
    assert status == abi.ERR_FORMAT and "File format verification failed" in abi.error()  # tests.cpp:141-153
    short = workdir / "short.bin"
    short.write_bytes(b"abc")
    assert abi.open(str(short))[0] == abi.ERR_FORMAT
    if abi.lib.memb_b200_device_count() == 0:
        # no CPU fallback: a good file still cannot be opened without a B200
        status, handle = abi.open(reference_files["trained"])
        assert status == abi.ERR_NO_DEVICE and not handle


def test_unsupported_storage_type(abi, workdir, reference_files):
    data = bytearray(open(reference_files["full"], "rb").read())
    root = int.from_bytes(data[0:4], "little")
    vt = root - int.from_bytes(data[root:root + 4], "little", signed=True)
    field = int.from_bytes(data[vt + 4:vt + 6], "little")
    assert data[root + field] == 1
    data[root + field] = 9
    path = workdir / "storage9.bin"
    path.write_bytes(bytes(data))
    status, _ = abi.open(str(path))
    assert status == abi.ERR_UNSUPPORTED and "is not supported" in abi.error()


def test_truncated_files_are_rejected_not_crashing(abi, synthetic, workdir):
    for storage, bits in (("trained", 4), ("uniform", 8), ("full", 8)):
        data = open(synthetic(storage, bits, rows=50, dim=8).path, "rb").read()
        for cut in (12, len(data) // 3, len(data) // 2, len(data) - 5):
            path = workdir / "cut.bin"
            path.write_bytes(data[:cut])
            status, handle = abi.open(str(path))
            assert status in (abi.ERR_FORMAT, abi.ERR_NO_DEVICE, abi.ERR_UNSUPPORTED), (storage, cut, status)
            assert not handle


def test_gpu_converter_needs_a_device(abi):
    """Builder(device=...) never falls back to the host silently."""
    import memb_b200

    if abi.lib.memb_b200_device_count() == 0:
        with pytest.raises((RuntimeError, ValueError)):
            memb_b200.Builder(3, "trained", 4, device=0)
    with pytest.raises((RuntimeError, ValueError)):
        memb_b200.Builder(3, "trained", 4, device=4096)
    memb_b200.Builder(3, "trained", 4, device=None)  # host encoder


def test_builder_errors():
    import memb_b200

    with pytest.raises(RuntimeError, match="not supported"):
        memb_b200.Builder(3, "zip", 4)
    builder = memb_b200.Builder(3, "trained", 4)
    builder.add_word("a", np.zeros(3, np.float32))
    with pytest.raises(RuntimeError):  # invalidDimensionThrows, src/tests.cpp:115-122
        builder.add_word("b", np.zeros(4, np.float32))
    with pytest.raises(RuntimeError):  # duplicateWordThrows, src/tests.cpp:124-132
        builder.add_word("a", np.zeros(3, np.float32))
    with pytest.raises(RuntimeError, match="1-dimensional"):  # python/memb_bindings.cpp:19-21
        builder.add_word("c", np.zeros((1, 3), np.float32))


@pytest.mark.parametrize("bits", [1, 2, 3, 4, 5, 6, 7, 8])
def test_trained_files_decode_to_nearest_centroid(oracle, synthetic, bits):
    syn = synthetic("trained", bits, rows=600, dim=64)
    reader = oracle.open(syn.path)
    assert reader.rows == 600 and reader.dim == 64 and reader.storage == 3
    assert reader.keys() == sorted(w.encode() for w in syn.words)
    codebook = reader.codebook()
    assert 1 <= len(codebook) <= min(1 << bits, 255)
    assert np.all(np.diff(codebook) > 0)
    got = reader.batch_embedding(syn.words)
    # every decoded value is a centroid, and it is the nearest one (kmeans predict, src/kmeans.cpp:65-79)
    assert np.isin(got, codebook).all()
    nearest = codebook[np.abs(syn.matrix[..., None] - codebook).argmin(-1)]
    assert np.mean(nearest != got) < 1e-3  # ties at split points may go either way
    keys, codes, lens = reader.prefix_code()
    assert np.all(np.diff(lens.astype(int)) >= 0) and lens.max() <= 16
    assert sum(2.0 ** -int(n) for n in lens) == 1.0 or len(lens) == 1


@pytest.mark.parametrize("bits", [1, 2, 4, 8])
def test_uniform_files_follow_the_reference_quantiser(oracle, synthetic, bits):
    syn = synthetic("uniform", bits, rows=300, dim=40)
    got = oracle.open(syn.path).batch_embedding(syn.words)
    levels = min(1 << bits, 255)
    mn = syn.matrix.min(1, keepdims=True)
    mx = syn.matrix.max(1, keepdims=True)
    # src/uniform_compression.cpp:24-30 then :68-72, all in binary32, no FMA
    q = (np.float32(levels) * (syn.matrix - mn) / (mx - mn)).astype(np.uint8)
    want = mn + (mx - mn) * q.astype(np.float32) / np.float32(levels)
    assert bits_equal(got, want.astype(np.float32))


def test_full_files_are_exact(oracle, synthetic):
    syn = synthetic("full", 8, rows=200, dim=33)
    assert bits_equal(oracle.open(syn.path).batch_embedding(syn.words), syn.matrix)


def test_python_writer_and_cpp_builder_files_read_alike(oracle, workdir, golden):
    """Files from the independent Python writer parse like the Builder's."""
    rng = np.random.default_rng(3)
    name = "gauss6_pruned"
    keys, lens = golden[name + ".keys"], golden[name + ".lens"]
    dim, rows = 37, 120
    centroids = np.sort(rng.standard_normal(int(keys.max()) + 1).astype(np.float32))
    words = [("k%03d" % i).encode() for i in rng.permutation(rows)]
    symbols = keys[rng.integers(0, len(keys), (rows, dim))]
    path = str(workdir / "py_trained.bin")
    sorted_words = membfile_py.write_trained(path, words, symbols, keys, lens, centroids,
                                             row_order=rng.permutation(rows), tail=b"\xff\xee\xdd")
    reader = oracle.open(path)
    assert reader.keys() == sorted_words and reader.dim == dim
    got = reader.batch_embedding(words)
    assert bits_equal(got, centroids[symbols])
    for r in range(0, rows, 17):
        word_row = words.index(sorted_words[r])
        assert np.array_equal(reader.decode_symbols(r), symbols[word_row])

    values = rng.integers(0, 17, (rows, dim)).astype(np.uint8)
    mins = rng.standard_normal(rows).astype(np.float32)
    maxs = mins + rng.random(rows).astype(np.float32)
    mins[::7] = 0.0  # omitted field -> default
    maxs[3::11] = 0.0
    path = str(workdir / "py_uniform.bin")
    membfile_py.write_uniform(path, words, mins, maxs, values, 16)
    got = oracle.open(path).batch_embedding(words)
    want = mins[:, None] + (maxs - mins)[:, None] * values.astype(np.float32) / np.float32(16)
    assert bits_equal(got, want.astype(np.float32))

    matrix = rng.standard_normal((rows, dim)).astype(np.float32)
    path = str(workdir / "py_full.bin")
    membfile_py.write_full(path, words, matrix)
    assert bits_equal(oracle.open(path).batch_embedding(words), matrix)


def test_product_never_touches_the_oracle():
    """No file of the product package mentions the oracle library."""
    pkg = os.path.join(helpers.ROOT, "memb_b200")
    for base, _, files in os.walk(pkg):
        if "build" in base.split(os.sep):
            continue
        for name in files:
            if name.endswith((".py", ".cpp", ".cu", ".cuh", ".h")):
                text = open(os.path.join(base, name), errors="ignore").read()
                assert "liboracle" not in text and "oracle/" not in text and "import oracle" not in text, name


# ---- Python API semantics that need no device (python/memb/readers_union.py, reader.py) ----
class FakeReader:
    pass


def _fake(dim, table):
    from memb_b200.reader import BaseReader

    class Fake(BaseReader):
        def __init__(self):
            self.dim = dim

        def keys(self):
            return sorted(table)

        def word_embedding(self, word):
            return np.asarray(table.get(word, np.zeros(dim)), np.float32)

        def batch_embedding(self, words):
            return np.stack([self.word_embedding(w) for w in words]) if words else np.zeros((0, dim), np.float32)

        def tokenizer_embedding(self, tokenizer):
            from memb_b200.reader import tokenizer_word_list

            return self.batch_embedding(tokenizer_word_list(tokenizer))

    return Fake()


def test_union_semantics_without_device():
    from memb_b200 import ReadersUnion

    a = _fake(2, {"x": [1, 2], "y": [3, 4]})
    b = _fake(2, {"y": [10, 20], "z": [5, 6]})
    c = _fake(3, {"x": [7, 8, 9]})
    with pytest.raises(AssertionError):  # readers_union.py:59-60
        ReadersUnion([a], "average")
    with pytest.raises(KeyError):  # readers_union.py:62-66
        ReadersUnion([a, b], "sum")
    with pytest.raises(AssertionError):  # readers_union.py:7-10
        ReadersUnion([a, c], "average")
    avg = ReadersUnion([a, b], "average")
    assert avg.dim == 2 and avg.keys() == ["x", "y", "z"]
    assert np.array_equal(avg["y"], np.array([6.5, 12], np.float32))
    assert np.array_equal(avg[["x", "q"]], np.array([[0.5, 1], [0, 0]], np.float32))
    cat = ReadersUnion([a, c], "concatenate")
    assert cat.dim == 5
    assert np.array_equal(cat["x"], np.array([1, 2, 7, 8, 9], np.float32))
    with pytest.raises(TypeError, match="Key type is not supported"):  # reader.py:16-17
        avg[3]


def test_tokenizer_word_list():
    from memb_b200.reader import tokenizer_word_list

    class Tok:
        word_index = {"b": 2, "a": 1, "d": 5}
        num_words = None

    assert tokenizer_word_list(Tok) == ["", "a", "b", "", "", "d"]  # reader.py:100-109
    Tok.num_words = 3
    assert tokenizer_word_list(Tok) == ["", "a", "b"]


def test_word_collections_marshal_identically():
    """list[str] (the reference's form, python/memb_bindings.cpp:54-56) and the additive
    NumPy / packed forms reach the C ABI as the same UTF-8 blob + offsets."""
    from memb_b200 import _memb

    words = ["a", "", "bcd", "é€", "\U0001F600x", "the", "w" * 40]
    blob, offsets = _memb.pack_words(words)
    assert blob == "".join(words).encode("utf-8")
    assert offsets.dtype == np.uint64
    assert offsets.tolist() == np.cumsum([0] + [len(w.encode("utf-8")) for w in words]).tolist()
    forms = [np.array(words), np.array([w.encode("utf-8") for w in words]), np.array(words, dtype=object),
             (blob, offsets), (np.frombuffer(blob, np.uint8), offsets.tolist()), np.array(words)[::-1][::-1]]
    for form in forms:
        b, o = _memb.pack_words(form)
        assert b == blob and np.array_equal(o, offsets)
    b, o = _memb.pack_words(np.array(words)[::2])  # strided view
    assert b == "".join(words[::2]).encode("utf-8")
    assert _memb.pack_words([b"raw", "str"])[0] == b"rawstr"
    assert _memb.Reader.count_words(words) == len(words)
    assert _memb.pack_words([])[1].tolist() == [0]
    with pytest.raises(TypeError):
        _memb.pack_words([1, 2])
    with pytest.raises(TypeError):
        _memb.pack_words("the")
    with pytest.raises(TypeError):
        _memb.pack_words(np.zeros(3))
    with pytest.raises(TypeError):
        _memb.pack_words(np.array([["a", "b"]]))
    with pytest.raises(ValueError):
        _memb.pack_words((b"abc", np.array([0, 2, 1], np.uint64)))
    with pytest.raises(ValueError):
        _memb.pack_words((b"abc", np.array([0, 4], np.uint64)))


def test_converter_cli_matches_the_builder(oracle, workdir):
    """tools/converter/memb_converter of the reference (flags :39-79, parsing rules :11-33):
    same flags, header / short lines skipped, bad weights reported, odd-length lines discarded,
    duplicate words reported -- and the file equals the one built directly with the Builder."""
    import io
    import subprocess
    import sys

    import memb_b200
    from memb_b200 import converter

    rng = np.random.default_rng(77)
    words = ["w%d" % i for i in range(60)] + ["é€", "the"]
    matrix = rng.standard_normal((len(words), 5)).astype(np.float32)
    text = workdir / "vectors.txt"
    with open(text, "w") as f:
        f.write("%d 5\n" % len(words))                      # word2vec header: skipped (two fields)
        for w, v in zip(words, matrix):
            f.write(w + " " + " ".join(repr(float(x)) for x in v) + "\n")
        f.write("odd 1.0 2.0 3.0\n")                         # wrong length: discarded
        f.write("bad 1.0 x 3.0 4.0 5.0\n")                   # does not parse: reported
        f.write("w3 " + " ".join(["0.5"] * 5) + "\n")        # duplicate word: reported by the builder
        f.write("lonely\n")
    log = io.StringIO()
    dest = workdir / "converted.bin"
    written = converter.convert(str(text), str(dest), "trained", 6, out=log)
    assert written == len(words)
    messages = log.getvalue()
    assert "Exception while parsing" in messages and "1 items are discarded" in messages
    assert "duplicate word w3" in messages
    direct = memb_b200.Builder(5, "trained", 6)
    for w, v in zip(words, matrix):
        direct.add_word(w, np.asarray([repr(float(x)) for x in v], dtype=np.float32))
    direct_path = workdir / "direct.bin"
    direct.save(str(direct_path))
    assert dest.read_bytes() == direct_path.read_bytes()
    reader = oracle.open(str(dest))
    assert reader.dim == 5 and reader.keys() == sorted(w.encode("utf-8") for w in words)
    got = reader.batch_embedding(words)
    assert np.abs(got - matrix).max() < 0.2  # 64 centroids over N(0,1) values
    # --max-words counts LINES (header included), like islice in the reference
    limited = workdir / "limited.bin"
    assert converter.main(["--from", str(text), "--to", str(limited), "--quantization", "full", "--max-words", "11"]) == 0
    assert oracle.open(str(limited)).rows == 10
    script = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "converter", "memb_converter")
    cli = workdir / "cli.bin"
    run = subprocess.run([sys.executable, script, "--from", str(text), "--to", str(cli), "--quantization", "uniform",
                          "--bits-per-weight", "8"], capture_output=True, text=True)
    assert run.returncode == 0, run.stderr
    assert "items are discarded" in run.stdout and oracle.open(str(cli)).rows == len(words)
    with pytest.raises(SystemExit):
        converter.parse_args(["--from", "a", "--to", "b", "--quantization", "lossless"])


def _fbb_files(workdir):
    """Files in the layout a real flatbuffers::FlatBufferBuilder emits (children before parents,
    shared vtables reached through negative soffsets, omitted defaults, root near the front)."""
    import membfile_py

    rng = np.random.default_rng(5)
    words = [b"the", b"th", b"tho", b"a", b"zz", "été".encode(), b"m", b"o0"]
    n, dim = len(words), 12
    values = rng.integers(0, 256, (n, dim)).astype(np.uint8)
    mins = [0.0, -1.0, 0.5, 0.0, 2.0, -3.0, 0.0, 1.0]
    maxs = [1.0, 0.0, 0.5, 0.0, 4.0, 3.0, 9.0, 1.5]
    matrix = rng.standard_normal((n, dim)).astype(np.float32)
    keys, lengths = [3, 1, 0, 2, 4], [1, 2, 3, 4, 4]
    centroids = rng.standard_normal(5).astype(np.float32)
    symbols = np.array(keys)[rng.integers(0, 5, (n, dim))]
    paths = {k: str(workdir / ("fbb_%s.bin" % k)) for k in ("uniform", "full", "trained")}
    membfile_py.write_uniform_fbb(paths["uniform"], words, mins, maxs, values, 255)
    membfile_py.write_full_fbb(paths["full"], words, matrix)
    membfile_py.write_trained_fbb(paths["trained"], words, symbols, keys, lengths, centroids)
    expect = {"uniform": None, "full": matrix, "trained": centroids[symbols]}
    return words, paths, expect


def test_flatbufferbuilder_layout_is_read(abi, oracle, workdir):
    """the two in-house writers lay buffers out front to back; a file written by memb_converter
    (flatc 1.9.0 FlatBufferBuilder) is back to front with shared vtables and negative soffsets.  No such
    file can be produced here (flatc is absent), so the layout is reproduced by hand."""
    words, paths, expect = _fbb_files(workdir)
    for kind, path in paths.items():
        data = open(path, "rb").read()
        root = int.from_bytes(data[0:4], "little")
        assert data[4:8] == b"memb" and root < 64  # the root table is the last thing built: lowest address
        reader = oracle.open(path)
        assert reader.keys() == sorted(words)
        got = reader.batch_embedding([w.decode() for w in words])
        if expect[kind] is not None:
            assert np.array_equal(got, expect[kind])
        status, handle = abi.open(path)  # the product's parser runs before any device is touched
        assert status in (abi.OK, abi.ERR_NO_DEVICE), abi.error()
        if status == abi.OK:
            abi.lib.memb_b200_close(handle)
    # at least one table reaches its vtable through a negative soffset (a shared vtable)
    data = open(paths["full"], "rb").read()
    negatives = 0
    for pos in range(8, len(data) - 4, 4):
        soffset = int.from_bytes(data[pos:pos + 4], "little", signed=True)
        vt = pos - soffset
        if soffset < 0 and 0 <= vt < len(data) - 4 and data[vt:vt + 4] == b"\x08\x00\x0c\x00":
            negatives += 1
    assert negatives >= len(words) - 1


def test_every_option_is_documented_in_the_header():
    """memb_b200_set_option's names (device_reader.cu) all appear in the comment of include/memb_b200.h."""
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    source = open(os.path.join(root, "memb_b200", "csrc", "device_reader.cu")).read()
    header = open(os.path.join(root, "include", "memb_b200.h")).read()
    names = set(re.findall(r'key == "([a-z0-9_]+)"', source))
    assert len(names) >= 10
    undocumented = sorted(n for n in names if '"%s"' % n not in header)
    assert not undocumented, undocumented
