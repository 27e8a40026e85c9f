"""Parity tests proper: the CUDA path, called through the C ABI (ctypes) and
through the `_memb` binding, against the CPU oracle on the same inputs.
Bar: bit-exact for row ids, decoded trained rows, full rows and union merges;
uniform dequant is allowed 1 ulp by the north star — asserted 0 ulp here."""
import numpy as np
import pytest

import helpers
import membfile_py
from conftest import TEST_VECTORS
from helpers import bits_equal

pytestmark = pytest.mark.gpu


def queries_for(syn, rng, n, miss_fraction=0.05):
    words = [syn.words[i] for i in rng.integers(0, len(syn.words), n)]
    for i in rng.integers(0, n, max(1, int(n * miss_fraction))):
        words[i] = ["", "o", "zzzzzzzz", "w", "é", "w1\0tail", "absent%d" % i][i % 7]
    return words


# ---- the reference's own cases (src/tests.cpp) through the CUDA path ----
@pytest.mark.parametrize("storage", ["full", "uniform", "trained"])
def test_reference_builder_cases_on_gpu(abi, oracle, reference_files, storage):
    reader = abi.open_checked(reference_files[storage])
    want = oracle.open(reference_files[storage])
    assert reader.dim == 3 and reader.rows == 6
    assert reader.keys() == sorted(w.encode() for w, _ in TEST_VECTORS)  # tests.cpp:36-43
    words = [w for w, _ in TEST_VECTORS] + ["o", "zzz", ""]
    got = reader.lookup_words(words)
    assert bits_equal(got, want.batch_embedding(words))
    for (word, vector), row in zip(TEST_VECTORS, got):
        assert np.allclose(row, vector, rtol=0.01, atol=0)  # tests.cpp:49
    assert np.all(got[len(TEST_VECTORS):] == 0.0)  # tests.cpp:53-56
    assert np.array_equal(reader.resolve(words), want.resolve(words))


def test_threaded_decoder_case_on_gpu(abi, oracle, reference_files):
    """threadedDecoderWorks (tests.cpp:90-113): 1025 words; any partition gives the same bits."""
    batch = [TEST_VECTORS[i % len(TEST_VECTORS)][0] for i in range(1025)]
    reader = abi.open_checked(reference_files["trained"])
    got = reader.lookup_words(batch)
    assert bits_equal(got, oracle.open(reference_files["trained"], 4).batch_embedding(batch))


def test_open_errors_on_gpu(abi, workdir):
    status, handle = abi.open(str(workdir / "missing_gpu.bin"))
    assert status == abi.ERR_IO
    bad = workdir / "bad_gpu.bin"
    bad.write_bytes(b"0123456789")
    assert abi.open(str(bad))[0] == abi.ERR_FORMAT
    import memb_b200

    with pytest.raises(RuntimeError):
        memb_b200.Reader(str(bad))
    with pytest.raises(Exception):
        memb_b200.Reader(str(workdir / "missing_gpu.bin"))


# ---- golden codec vectors (reference codec output) through the kernel ----
def test_golden_codec_vectors_on_gpu(abi, golden, workdir):
    rng = np.random.default_rng(11)
    for name in golden["names"]:
        keys, lens, symbols = golden[name + ".keys"], golden[name + ".lens"], golden[name + ".symbols"]
        centroids = rng.standard_normal(int(keys.max()) + 1).astype(np.float32)
        for dim in (len(symbols), 300, 7):
            rows = len(symbols) // dim
            sym = symbols[:rows * dim].reshape(rows, dim)
            words = [("g%04d" % i).encode() for i in range(rows)]
            path = str(workdir / ("golden_%s_%d.bin" % (name, dim)))
            membfile_py.write_trained(path, words, sym, keys, lens, centroids, tail=b"\xa5" * 3)
            reader = abi.open_checked(path)
            got = reader.decode_range(0, rows)
            assert bits_equal(got, centroids[sym]), (name, dim)
            reader.close()


# ---- trained: 1..8 bits, ids / words / range, odd sizes ----
@pytest.mark.parametrize("bits", [1, 2, 3, 4, 5, 6, 7, 8])
def test_trained_parity_bits_sweep(abi, oracle, synthetic, bits):
    syn = synthetic("trained", bits, rows=3000, dim=300)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    rng = np.random.default_rng(bits)
    assert bits_equal(reader.decode_range(0, 3000), want.decode_range(0, 3000))
    ids = rng.integers(-2, 3002, 5000).astype(np.int32)
    assert bits_equal(reader.decode_ids(ids), want.batch_ids(ids))
    words = queries_for(syn, rng, 2000)
    assert np.array_equal(reader.resolve(words), want.resolve(words))
    assert bits_equal(reader.lookup_words(words), want.batch_embedding(words))
    stats = reader.load_stats()
    assert stats.codebook_size == len(want.codebook()) and stats.max_code_len == want.prefix_code()[2].max()


@pytest.mark.parametrize("dim", [1, 3, 4, 12, 20, 31, 32, 33, 36, 44, 64, 100, 301, 768])
def test_trained_parity_dims(abi, oracle, synthetic, dim):
    syn = synthetic("trained", 4, rows=700, dim=dim, seed=dim)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    assert bits_equal(reader.decode_range(0, 700), want.decode_range(0, 700))
    assert bits_equal(reader.decode_range(13, 77), want.decode_range(13, 77))


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 1023, 1024, 1025])
def test_batch_sizes(abi, oracle, synthetic, n):
    syn = synthetic("trained", 6, rows=3000, dim=300)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path, 4)
    rng = np.random.default_rng(n)
    words = queries_for(syn, rng, n) if n else []
    got = reader.lookup_words(words)
    assert got.shape == (n, 300) and bits_equal(got, want.batch_embedding(words))
    ids = rng.integers(0, 3000, n).astype(np.int32)
    ids[: n // 2] = ids[n // 2: 2 * (n // 2)]  # duplicates in one batch
    assert bits_equal(reader.decode_ids(ids), want.batch_ids(ids))


def test_trained_edge_codes(abi, oracle, workdir):
    """Files the Builder cannot emit: one-symbol code (zero-byte rows), 16-bit codes
    (escape path beyond the device LUT), 255 symbols, rows packed in shuffled order."""
    rng = np.random.default_rng(2)
    cases = {
        "one": ([9], [0]),
        "two": ([1, 0], [1, 1]),
        "deep16": (list(range(17)), list(range(1, 17)) + [16]),
        "wide255": (list(rng.permutation(255)), [7] * 1 + [8] * 254),
    }
    for name, (keys, lens) in cases.items():
        assert sum(2.0 ** -n for n in lens) == 1.0 or len(lens) == 1
        centroids = rng.standard_normal(max(keys) + 1).astype(np.float32)
        for dim in (300, 5):
            rows = 257
            weights = np.array([2.0 ** -n for n in lens])
            weights = weights + 0.02  # long codes far more often than a Huffman source would
            sym = np.array(keys)[rng.choice(len(keys), (rows, dim), p=weights / weights.sum())]
            words = [("e%03d" % i).encode() for i in rng.permutation(rows)]
            path = str(workdir / ("edge_%s_%d.bin" % (name, dim)))
            membfile_py.write_trained(path, words, sym, keys, lens, centroids, row_order=rng.permutation(rows))
            reader, want = abi.open_checked(path), oracle.open(path)
            got = reader.lookup_words(words)
            assert bits_equal(got, centroids[sym]), (name, dim)
            assert bits_equal(got, want.batch_embedding(words)), (name, dim)
            reader.close()


def test_flatbufferbuilder_layout_on_gpu(abi, oracle, workdir):
    """Files in real FlatBufferBuilder order (tests/test_host.py::_fbb_files) through the CUDA path."""
    from test_host import _fbb_files

    words, paths, _ = _fbb_files(workdir)
    query = [w.decode() for w in words] + ["o", "", "zzz"]
    for path in paths.values():
        reader, want = abi.open_checked(path), oracle.open(path)
        assert reader.keys() == want.keys()
        assert bits_equal(reader.lookup_words(query), want.batch_embedding(query))
        assert bits_equal(reader.decode_range(0, len(words)), want.decode_range(0, len(words)))
        reader.close()


@pytest.mark.parametrize("bits", [1, 4, 6, 8])
def test_trained_kernel_variants(abi, oracle, synthetic, bits):
    """The same table through every trained path: whole-row kernel on fixed-stride records
    (default), ring kernel on records, ring kernel on the offset table; single lookups and a
    union of the table with a second one; ids with misses, ragged tiles."""
    syn, other = synthetic("trained", bits, rows=3000, dim=300), synthetic("trained", 5, rows=2500, dim=300, seed=77)
    want, want_other = oracle.open(syn.path), oracle.open(other.path)
    rng = np.random.default_rng(100 + bits)
    ids = rng.integers(-3, 3003, 4100).astype(np.int32)
    ids2 = np.stack([ids, rng.integers(-3, 2503, 4100).astype(np.int32)])
    parts = [want.batch_ids(ids2[0]), want_other.batch_ids(ids2[1])]
    try:
        for rows_kernel, fixed in ((2, 1), (1, 1), (0, 1), (0, 0), (2, 0)):
            abi.check(abi.lib.memb_b200_set_option(b"trained_rows_kernel", rows_kernel))
            abi.check(abi.lib.memb_b200_set_option(b"trained_fixed_records", fixed))
            reader, second = abi.open_checked(syn.path), abi.open_checked(other.path)
            tag = (bits, rows_kernel, fixed)
            assert bits_equal(reader.decode_range(0, 3000), want.decode_range(0, 3000)), tag
            assert bits_equal(reader.decode_range(17, 1030), want.decode_range(17, 1030)), tag
            assert bits_equal(reader.decode_ids(ids), parts[0]), tag
            assert bits_equal(helpers.union_decode_ids(abi, [reader, second], abi.AVERAGE, ids2, 300),
                              np.mean(parts, axis=0)), tag
            assert bits_equal(helpers.union_decode_ids(abi, [reader, second], abi.CONCATENATE, ids2, 600),
                              np.concatenate(parts, axis=-1)), tag
            reader.close()
            second.close()
    finally:
        abi.lib.memb_b200_set_option(b"trained_rows_kernel", 0)
        abi.lib.memb_b200_set_option(b"trained_fixed_records", 1)


def test_lut_layouts_do_not_change_results(abi, oracle, synthetic):
    """GPU analogue of trainedBuilderWorksWithIndirectDecoder (tests.cpp:76-88):
    every LUT layout / kernel variant decodes the same bits."""
    syn = synthetic("trained", 8, rows=3000, dim=300)
    want = oracle.open(syn.path).decode_range(0, 3000)
    try:
        for name, value in (("trained_lut_replicas", 1), ("trained_lut_replicas", 32), ("trained_warps", 3),
                            ("trained_generic", 1)):
            abi.check(abi.lib.memb_b200_set_option(name.encode(), value))
            reader = abi.open_checked(syn.path)
            assert bits_equal(reader.decode_range(0, 3000), want), (name, value)
            reader.close()
            abi.check(abi.lib.memb_b200_set_option(name.encode(), 0))
    finally:
        for name in ("trained_lut_replicas", "trained_warps", "trained_generic"):
            abi.lib.memb_b200_set_option(name.encode(), 0)


# ---- uniform / full ----
@pytest.mark.parametrize("bits", [1, 2, 4, 8])
def test_uniform_parity(abi, oracle, synthetic, bits):
    syn = synthetic("uniform", bits, rows=3000, dim=300, seed=5)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    rng = np.random.default_rng(bits)
    assert bits_equal(reader.decode_range(0, 3000), want.decode_range(0, 3000))  # 0 ulp (bar: 1 ulp)
    words = queries_for(syn, rng, 1500)
    assert np.array_equal(reader.resolve(words), want.resolve(words))
    assert bits_equal(reader.lookup_words(words), want.batch_embedding(words))


@pytest.mark.parametrize("bits,dim", [(8, 8), (8, 12), (4, 300), (8, 300), (2, 1024), (8, 8188)])
def test_uniform_rows_kernel_against_element_walk(abi, oracle, synthetic, bits, dim):
    """Dense uniform lookups run through uniform_rows_kernel (a warp per 32-row group, record
    offsets in shared memory a group ahead) unless levels is a power of two; option
    uniform_rows_kernel = 2 forces it, 0 sends the same call through the element walk.  Both must equal the oracle bit for bit: ragged batches, one to many groups
    per block, missing and out-of-range ids, sub-ranges."""
    rows = 700 if dim <= 1024 else 70
    syn = synthetic("uniform", bits, rows=rows, dim=dim, seed=60 + bits)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    rng = np.random.default_rng(dim)
    table = want.decode_range(0, rows)
    for option in (2, 1, 0):
        try:
            abi.check(abi.lib.memb_b200_set_option(b"uniform_rows_kernel", option))
            assert bits_equal(reader.decode_range(0, rows), table)
            assert bits_equal(reader.decode_range(rows // 3, rows // 3 + 45), table[rows // 3:rows // 3 + 45])
            for n in (1, 31, 32, 33, 1000, 40000 if dim <= 300 else 100):
                ids = rng.integers(0, rows, n).astype(np.int32)
                ids[::7] = -1
                ids[3::11] = rows + 5
                expect = np.where(((ids >= 0) & (ids < rows))[:, None], table[np.clip(ids, 0, rows - 1)], np.float32(0))
                assert bits_equal(reader.decode_ids(ids), expect)
        finally:
            abi.lib.memb_b200_set_option(b"uniform_rows_kernel", 1)


def test_uniform_omitted_fields_and_odd_dim(abi, oracle, workdir):
    rng = np.random.default_rng(4)
    rows, dim = 500, 37
    words = [("u%03d" % i).encode() for i in rng.permutation(rows)]
    values = rng.integers(0, 256, (rows, dim)).astype(np.uint8)
    mins = (rng.standard_normal(rows) * 3).astype(np.float32)
    maxs = mins + rng.random(rows).astype(np.float32) * 5
    mins[::5] = 0.0
    maxs[2::9] = 0.0
    mins[7] = maxs[7] = 1.5  # min == max
    mins[8], maxs[8] = np.float32(1e-42), np.float32(3e-42)  # denormals are kept
    path = str(workdir / "uni_edge.bin")
    membfile_py.write_uniform(path, words, mins, maxs, values, 255)
    reader, want = abi.open_checked(path), oracle.open(path)
    assert bits_equal(reader.lookup_words(words), want.batch_embedding(words))


@pytest.mark.parametrize("levels", [2, 4, 16, 255, 3, 100, 254, 1])
def test_uniform_divisor_modes_extreme_ranges(abi, oracle, workdir, levels):
    """`/ levels` runs as an exact multiply (powers of two), a constant-divisor FMA sequence
    (255) or the IEEE divide (anything else a file may hold); all must equal the reference's
    fp32 division bit for bit, including tiny, huge, zero, negative and denormal ranges where
    the FMA sequence has to hand over to the IEEE divide."""
    rng = np.random.default_rng(levels)
    rows, dim = 640, 300
    words = [("x%03d" % i).encode() for i in rng.permutation(rows)]
    values = rng.integers(0, 256, (rows, dim)).astype(np.uint8)
    scale = np.float32(10.0) ** rng.integers(-44, 38, rows).astype(np.float32)
    mins = (rng.standard_normal(rows).astype(np.float32) * scale).astype(np.float32)
    maxs = (mins + rng.random(rows).astype(np.float32) * scale).astype(np.float32)
    maxs[::7] = mins[::7]                      # zero range
    maxs[3::11] = -maxs[3::11]                 # max < min: negative range (and -0 products)
    mins[5::13] = np.float32(-0.0)
    maxs[6::17] = np.float32(3.0e38)           # products overflow to inf
    # ranges on either side of the per-row gate of the 255 path (1e-20, 1e20 / 256) and of the
    # per-element one (x = range * k around 1e-20 and 1e20)
    edge = np.array([1e-20, 1.0000001e-20, 9.999999e-21, 3.9e17, 3.8999999e17, 3.9000001e17, 3.90625e17, 3.92e17,
                     1e20 / 255, 1e20, 4e-23, 1e-22], np.float32)
    mins[100:100 + len(edge)] = np.float32(0.0)
    maxs[100:100 + len(edge)] = edge
    mins[120:120 + len(edge)] = -edge / 2
    maxs[120:120 + len(edge)] = edge / 2
    path = str(workdir / ("uni_div_%d.bin" % levels))
    membfile_py.write_uniform(path, words, mins, maxs, values, levels)
    reader, want = abi.open_checked(path), oracle.open(path)
    with np.errstate(all="ignore"):
        assert bits_equal(reader.lookup_words(words), want.batch_embedding(words))
        assert bits_equal(reader.decode_range(0, rows), want.decode_range(0, rows))


@pytest.mark.parametrize("dim", [3, 300])
def test_full_parity(abi, oracle, synthetic, dim):
    syn = synthetic("full", 8, rows=2000, dim=dim)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    rng = np.random.default_rng(1)
    words = queries_for(syn, rng, 1500)
    got = reader.lookup_words(words)
    assert bits_equal(got, want.batch_embedding(words))
    assert bits_equal(reader.decode_range(0, 2000), syn.matrix[np.argsort([w.encode() for w in syn.words])])


# ---- word -> row id ----
def test_resolve_semantics(abi, oracle, workdir):
    """strcmp semantics: cut at NUL, prefix / extension of keys, non-ASCII byte order,
    empty string as a key, word beyond the last key."""
    words = ["", "a", "ab", "abc", "b", "th", "the", "tho", "z", "zz", "é", "éa", "ж", "\U0001f600",
             "x" * 300]
    rng = np.random.default_rng(0)
    matrix = rng.standard_normal((len(words), 8)).astype(np.float32)
    path = str(workdir / "resolve.bin")
    helpers.build_file(path, "full", 8, words, matrix)
    reader, want = abi.open_checked(path), oracle.open(path)
    assert reader.keys() == sorted(w.encode("utf-8") for w in words)
    queries = words + ["t", "thee", "zzz", "ab\0c", "\0", "abd", "è", "\xff", "x" * 299, "x" * 301, b"\xc3",
                       b"\xff\xfe"]
    ids = reader.resolve(queries)
    assert np.array_equal(ids, want.resolve(queries))
    assert ids[queries.index("ab\0c")] == reader.keys().index(b"ab")
    assert ids[queries.index("\0")] == 0 and ids[queries.index("zzz")] == -1
    assert bits_equal(reader.lookup_words(queries), want.batch_embedding(queries))


def test_resolve_large_vocabulary(abi, oracle, synthetic):
    syn = synthetic("full", 8, rows=60000, dim=4, seed=9)
    reader, want = abi.open_checked(syn.path), oracle.open(syn.path)
    rng = np.random.default_rng(3)
    ids = reader.resolve(syn.words)
    order = np.argsort(np.array([w.encode() for w in syn.words], dtype=object))
    expect = np.empty(len(order), np.int32)
    expect[order] = np.arange(len(order))
    assert np.array_equal(ids, expect)
    queries = queries_for(syn, rng, 20000, 0.2)
    assert np.array_equal(reader.resolve(queries), want.resolve(queries))


# ---- ReadersUnion fused merges ----
def union_case(synthetic):
    a = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    b = synthetic("trained", 6, rows=2500, dim=300, seed=22)
    c = synthetic("uniform", 8, rows=2000, dim=300, seed=23)
    d = synthetic("full", 8, rows=1500, dim=300, seed=24)
    return [a, b, c, d]


@pytest.mark.parametrize("count", [5, 6, 7, 9])
def test_union_average_small_divisors(abi, oracle, synthetic, count):
    """np.mean over R = 5, 6, 7 readers takes the constant-divisor sequence of divideByReaders
    (verified exhaustively on the host, scripts/divconst_exhaustive.c), R = 9 the IEEE divide;
    sums that are exactly zero (words missing everywhere) and mixed storages included."""
    kinds = [("trained", 4), ("trained", 5), ("uniform", 8), ("full", 8), ("trained", 3), ("uniform", 4), ("trained", 6),
             ("full", 8), ("trained", 2)]
    syns = [synthetic(k, b, rows=600 + 50 * i, dim=60, seed=70 + i) for i, (k, b) in enumerate(kinds[:count])]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    rng = np.random.default_rng(count)
    words = sum([queries_for(s, rng, 150) for s in syns], [])
    want = np.mean([o.batch_embedding(words) for o in oracles], axis=0)
    assert bits_equal(helpers.union_lookup_words(abi, readers, abi.AVERAGE, words, 60), want)


@pytest.mark.parametrize("count", [2, 3, 4])
def test_union_average_and_concatenate(abi, oracle, synthetic, count):
    syns = union_case(synthetic)[:count]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    rng = np.random.default_rng(count)
    words = []
    for s in syns:
        words += queries_for(s, rng, 700)
    rng.shuffle(words)
    parts = [o.batch_embedding(words) for o in oracles]
    # the reference's own merge expressions (python/memb/readers_union.py:18,32)
    want_avg = np.mean(parts, axis=0)
    want_cat = np.concatenate(parts, axis=-1)
    got_avg = helpers.union_lookup_words(abi, readers, abi.AVERAGE, words, 300)
    got_cat = helpers.union_lookup_words(abi, readers, abi.CONCATENATE, words, 300 * count)
    assert want_avg.dtype == np.float32 and bits_equal(got_avg, want_avg)
    assert bits_equal(got_cat, want_cat)
    ids = np.stack([o.resolve(words) for o in oracles])
    assert bits_equal(helpers.union_decode_ids(abi, readers, abi.AVERAGE, ids, 300), want_avg)
    assert bits_equal(helpers.union_decode_ids(abi, readers, abi.CONCATENATE, ids, 300 * count), want_cat)


def test_union_mixed_dims_concatenate(abi, oracle, synthetic):
    syns = [synthetic("trained", 5, rows=900, dim=100, seed=31), synthetic("uniform", 4, rows=800, dim=33, seed=32),
            synthetic("trained", 3, rows=700, dim=7, seed=33)]
    readers = [abi.open_checked(s.path) for s in syns]
    rng = np.random.default_rng(8)
    words = sum([queries_for(s, rng, 300) for s in syns], [])
    want = np.concatenate([oracle.open(s.path).batch_embedding(words) for s in syns], axis=-1)
    assert bits_equal(helpers.union_lookup_words(abi, readers, abi.CONCATENATE, words, 140), want)
    status = abi.lib.memb_b200_union_lookup_words((helpers.c_void_p * 3)(*[r.h for r in readers]), 3, abi.AVERAGE,
                                                  b"", None, 0, None)
    assert status == abi.ERR_ARGUMENT  # dims differ (readers_union.py:7-10)


# ---- the Python drop-in API on top ----
def test_python_api(oracle, synthetic):
    import memb_b200
    from memb import Reader, ReadersUnion  # drop-in alias package

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    other = synthetic("trained", 6, rows=2500, dim=300, seed=22)
    want, want2 = oracle.open(syn.path), oracle.open(other.path)
    reader = Reader(syn.path)
    assert reader.dim == 300
    assert reader.keys() == [k.decode() for k in want.keys()]
    word = syn.words[5]
    assert reader[word].shape == (300,) and bits_equal(reader[word], want.word_embedding(word))
    assert np.all(reader["definitely missing"] == 0)
    words = syn.words[:50] + ["nope", ""]
    assert bits_equal(reader[words], want.batch_embedding(words))
    with pytest.raises(TypeError):
        reader[0]

    class Tok:
        word_index = {w: i + 1 for i, w in enumerate(syn.words[:40])}
        num_words = 30

    emb = reader.tokenizer_embedding(Tok)
    assert emb.shape == (30, 300) and np.all(emb[0] == 0)
    assert bits_equal(emb[1:], want.batch_embedding(syn.words[:29]))
    assert bits_equal(reader._all_rows(None), want.decode_range(0, 3000))

    union = ReadersUnion([reader, Reader(other.path)], "average")
    parts = [want.batch_embedding(words), want2.batch_embedding(words)]
    assert bits_equal(union[words], np.mean(parts, axis=0))
    assert bits_equal(union[word], np.mean([p[5] for p in parts], axis=0))
    cat = ReadersUnion([reader, Reader(other.path)], "concatenate")
    assert cat.dim == 600 and bits_equal(cat[words], np.concatenate(parts, axis=-1))
    assert union.keys() == sorted(set(syn.words) | set(other.words))
    assert memb_b200._memb.launch_count() > 0


def test_to_keyed_vectors_and_nested_unions(oracle, synthetic, monkeypatch):
    """to_keyed_vectors (python/memb/reader.py:19-30) with a stub of gensim's KeyedVectors: the
    vectors handed over are the rows of keys(), in that order (a Reader takes them from one
    decode_range(0, V), a union from one batch over its keys).  A union nested in a union merges
    on the GPU with np.mean's arithmetic; a foreign reader object still works (host merge)."""
    import sys
    import types
    from memb_b200 import Reader, ReadersUnion

    class KeyedVectors:
        def __init__(self, vector_size):
            self.vector_size, self.words, self.vectors = vector_size, None, None

        def add(self, words, vectors):
            self.words, self.vectors = list(words), np.asarray(vectors)

    gensim, models = types.ModuleType("gensim"), types.ModuleType("gensim.models")
    models.KeyedVectors = KeyedVectors
    gensim.models = models
    monkeypatch.setitem(sys.modules, "gensim", gensim)
    monkeypatch.setitem(sys.modules, "gensim.models", models)

    a, b, c = (synthetic("trained", 4, rows=3000, dim=300, seed=21), synthetic("trained", 6, rows=2500, dim=300, seed=22),
               synthetic("uniform", 8, rows=1500, dim=300, seed=23))
    oa, ob, oc = oracle.open(a.path), oracle.open(b.path), oracle.open(c.path)
    ra, rb, rc = Reader(a.path), Reader(b.path), Reader(c.path)
    kv = ra.to_keyed_vectors()
    assert kv.vector_size == 300 and kv.words == [k.decode() for k in oa.keys()]
    assert bits_equal(kv.vectors, oa.decode_range(0, 3000))
    union = ReadersUnion([ra, rb], "average")
    kv = union.to_keyed_vectors()
    assert kv.words == sorted(set(a.words) | set(b.words))
    assert bits_equal(kv.vectors, np.mean([oa.batch_embedding(kv.words), ob.batch_embedding(kv.words)], axis=0))

    rng = np.random.default_rng(12)
    words = queries_for(a, rng, 400) + queries_for(c, rng, 300)
    inner = np.mean([oa.batch_embedding(words), ob.batch_embedding(words)], axis=0)
    nested = ReadersUnion([union, rc], "average")
    assert bits_equal(nested[words], np.mean([inner, oc.batch_embedding(words)], axis=0))
    assert bits_equal(nested.batch_embedding(words, as_tensor=True).cpu().numpy(), nested[words])
    nested_cat = ReadersUnion([rc, union], "concatenate")
    assert nested_cat.dim == 600 and bits_equal(nested_cat[words], np.concatenate([oc.batch_embedding(words), inner], axis=-1))

    class Foreign:  # any object with the reader interface (python/memb/readers_union.py:90-96 duck-types)
        dim = 300

        def keys(self):
            return ["x"]

        def batch_embedding(self, words):
            return np.full((len(words), 300), 2.0, np.float32)

    mixed = ReadersUnion([ra, Foreign()], "average")
    assert bits_equal(mixed[words], np.mean([oa.batch_embedding(words), np.full((len(words), 300), 2.0, np.float32)], axis=0))


def test_device_results_are_ordered_on_the_callers_stream(oracle, synthetic):
    """as_tensor=True allocates from torch's caching allocator, which hands out a block as soon as
    the tensor that owned it is freed -- kernels queued earlier on the stream may still write it.
    The lookup must therefore be queued on the caller's stream (memb_b200_lookup_words_on)."""
    import torch
    from memb_b200 import Reader, ReadersUnion

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    other = synthetic("trained", 6, rows=2500, dim=300, seed=22)
    reader, reader2 = Reader(syn.path), Reader(other.path)
    rng = np.random.default_rng(3)
    words = queries_for(syn, rng, 60000)
    want = oracle.open(syn.path).batch_embedding(words)
    want_avg = np.mean([want, oracle.open(other.path).batch_embedding(words)], axis=0)
    union = ReadersUnion([reader, reader2], "average")
    for lookup, expect in ((reader.batch_embedding, want), (union.batch_embedding, want_avg)):
        for _ in range(3):
            scratch = torch.zeros((len(words), 300), dtype=torch.float32, device="cuda")
            for _ in range(400):  # tens of milliseconds of kernels that write `scratch`
                scratch.add_(1.0)
            del scratch  # the block is free for the allocator, the kernels are still queued
            got = lookup(words, as_tensor=True)
            assert bits_equal(got.cpu().numpy(), expect)


# ---- GPU converter ----
@pytest.mark.parametrize("bits,rows,dim", [(1, 2000, 300), (2, 2000, 300), (3, 1500, 64), (4, 12000, 300), (5, 2000, 301),
                                           (6, 3000, 100), (7, 2000, 300), (8, 11000, 300), (4, 1, 1), (8, 40, 3)])
def test_gpu_converter_matches_cpu_builder(oracle, workdir, bits, rows, dim):
    """Builder(..., device=0) runs predict / row sizes / bit packing on the GPU (device_builder.cu); the file
    must be BYTE-identical to the host encoder's -- same symbols, same first-occurrence order (Huffman
    tie-breaks), same packed rows -- and decode to the same rows.  More than 10 000 rows: values beyond the
    k-means sample; NaN / inf / huge values: the lower_bound's comparison semantics."""
    import memb_b200

    rng = np.random.default_rng(1000 * bits + dim)
    matrix = rng.standard_normal((rows, dim), dtype=np.float32)
    if rows > 100:
        matrix[rows // 2, 0] = np.nan
        matrix[rows // 2, 1 % dim] = np.inf
        matrix[rows // 3, 0] = -3e38
    words = ["g%x" % i for i in rng.permutation(rows)]
    paths = {}
    for name, device in (("cpu", None), ("gpu", 0)):
        builder = memb_b200.Builder(dim, "trained", bits, device=device)
        builder.add_words(words, matrix)
        paths[name] = str(workdir / ("conv_%s_%d_%d_%d.bin" % (name, bits, rows, dim)))
        builder.save(paths[name])
    cpu, gpu = open(paths["cpu"], "rb").read(), open(paths["gpu"], "rb").read()
    assert cpu == gpu
    reader = memb_b200.Reader(paths["gpu"])
    assert bits_equal(reader.decode_range(0, rows), oracle.open(paths["cpu"]).decode_range(0, rows))


# ---- additive word-collection forms and device-resident results ----
def test_word_collection_forms_and_device_tensors(oracle, synthetic):
    import torch
    from memb_b200 import Reader, ReadersUnion, _memb

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    other = synthetic("uniform", 8, rows=2500, dim=300, seed=23)
    want, want2 = oracle.open(syn.path), oracle.open(other.path)
    reader, reader2 = Reader(syn.path), Reader(other.path)
    rng = np.random.default_rng(5)
    words = queries_for(syn, rng, 700)
    words = [w for w in words if "\0" not in w]  # NumPy 'S'/'U' arrays cannot hold an embedded NUL
    expect = want.batch_embedding(words)
    packed = _memb.pack_words(words)
    for form in (np.array(words), np.array([w.encode() for w in words]), np.array(words, dtype=object), packed):
        assert bits_equal(reader.batch_embedding(form), expect)
        assert np.array_equal(reader.resolve(form), reader.resolve(words))
    sliced = (packed[0], packed[1][100:201])  # offsets need not start at 0
    assert bits_equal(reader.batch_embedding(sliced), expect[100:200])
    with pytest.raises(ValueError):
        reader.batch_embedding((packed[0][:10], packed[1]))
    with pytest.raises(TypeError):
        reader.batch_embedding(np.zeros(3))

    got = reader.batch_embedding(words, as_tensor=True)
    assert got.is_cuda and got.dtype == torch.float32 and bits_equal(got.cpu().numpy(), expect)
    assert bits_equal(reader.batch_embedding(packed, as_tensor=True).cpu().numpy(), expect)
    assert reader.batch_embedding([], as_tensor=True).shape == (0, 300)
    assert bits_equal(reader.word_embedding(words[3], as_tensor=True).cpu().numpy(), expect[3])
    ids = reader.resolve(words)
    on_device = torch.as_tensor(ids, device="cuda")
    assert bits_equal(reader.batch_embedding_ids(on_device, as_tensor=True).cpu().numpy(), expect)
    assert bits_equal(reader.batch_embedding_ids(ids, as_tensor=True).cpu().numpy(), expect)
    assert bits_equal(reader.decode_range(7, 1500, as_tensor=True).cpu().numpy(), want.decode_range(7, 1500))
    assert bits_equal(reader.to_tensor().cpu().numpy(), want.decode_range(0, 3000))

    class Tok:
        word_index = {w: i + 1 for i, w in enumerate(syn.words[:40])}
        num_words = None

    emb = reader.tokenizer_embedding(Tok, as_tensor=True)
    assert emb.shape == (41, 300) and bits_equal(emb[1:].cpu().numpy(), want.batch_embedding(syn.words[:40]))

    parts = [expect, want2.batch_embedding(words)]
    for mode, merged in (("average", np.mean(parts, axis=0)), ("concatenate", np.concatenate(parts, axis=-1))):
        union = ReadersUnion([reader, reader2], mode)
        assert bits_equal(union.batch_embedding(np.array(words)), merged)
        got = union.batch_embedding(packed, as_tensor=True)
        assert got.is_cuda and bits_equal(got.cpu().numpy(), merged)


def test_pinned_result_buffers_are_recycled_and_optional(oracle, synthetic):
    import gc
    from memb_b200 import Reader, _memb

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    want = oracle.open(syn.path)
    reader = Reader(syn.path)
    ids = np.random.default_rng(9).integers(-1, 3000, 2000).astype(np.int32)  # 2.4 MB result
    expect = want.batch_ids(ids)
    try:
        _memb.set_pinned_result_limit(1 << 30)
        first = reader.batch_embedding_ids(ids)
        assert bits_equal(first, expect) and first.flags.writeable and first.flags.c_contiguous
        first[0, 0] = 1.0  # an ordinary array to its owner
        kept = reader.batch_embedding_ids(ids)  # `first` still alive: must not share its buffer
        assert kept.ctypes.data != first.ctypes.data and bits_equal(kept, expect)
        address = first.ctypes.data
        del first
        gc.collect()
        assert _memb.pinned_result_state()[1] > 0
        again = reader.decode_range(0, 2000)
        assert again.ctypes.data == address and bits_equal(again, want.decode_range(0, 2000))
        assert bits_equal(kept, expect)
        del again, kept
        gc.collect()
        _memb.set_pinned_result_limit(0)
        assert _memb.pinned_result_state() == (0, 0)
        pageable = reader.batch_embedding_ids(ids)
        assert pageable.flags.owndata and bits_equal(pageable, expect)
    finally:
        _memb.set_pinned_result_limit(8 << 30)


@pytest.mark.parametrize("n", [1, 31, 33, 1000])
def test_union_concatenate_shared_sectors(abi, oracle, synthetic, n):
    """Column blocks that begin / end in the middle of a 32 B sector (dims 12, 20, 300): the
    shared sectors are written by the stitch pass from the readers' side buffers."""
    syns = [synthetic("trained", 4, rows=900, dim=12, seed=41), synthetic("uniform", 8, rows=800, dim=20, seed=42),
            synthetic("full", 4, rows=700, dim=300, seed=43), synthetic("trained", 2, rows=600, dim=300, seed=44),
            synthetic("trained", 6, rows=500, dim=8, seed=45)]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    rng = np.random.default_rng(n)
    words = sum([queries_for(s, rng, max(1, n // 5 + 1)) for s in syns], [])[:n]
    assert len(words) == n
    want = np.concatenate([o.batch_embedding(words) for o in oracles], axis=-1)
    assert want.shape == (n, 640)  # 2560 B rows: boundaries at 48, 128, 1328, 2528 B
    assert bits_equal(helpers.union_lookup_words(abi, readers, abi.CONCATENATE, words, 640), want)
    ids = np.stack([o.resolve(words) for o in oracles])
    assert bits_equal(helpers.union_decode_ids(abi, readers, abi.CONCATENATE, ids, 640), want)


def test_concurrent_calls_on_one_reader(oracle, synthetic):
    """memb::Reader is const and re-entrant (src/reader.h:11-40, per-call DecodeState in
    huffman_table_decoder.h:96-100): host-buffer calls from several Python threads (the GIL is
    released inside the binding) and device-pointer calls on several CUDA streams must all give
    the oracle's rows."""
    import threading
    import torch
    from memb_b200 import Reader, ReadersUnion

    syn = synthetic("trained", 4, rows=3000, dim=300, seed=21)
    other = synthetic("trained", 6, rows=2500, dim=300, seed=22)
    want, want2 = oracle.open(syn.path), oracle.open(other.path)
    reader, reader2 = Reader(syn.path), Reader(other.path)
    union = ReadersUnion([reader, reader2], "concatenate")
    keys = reader.keys()  # row id -> word
    failures = []

    def worker(seed):
        try:
            rng = np.random.default_rng(seed)
            stream = torch.cuda.Stream()
            for it in range(6):
                n = int(rng.integers(1, 4000))
                ids = rng.integers(-1, 3000, n).astype(np.int32)
                expect = want.batch_ids(ids)
                kind = (seed + it) % 4
                if kind == 0:
                    got = reader.batch_embedding_ids(ids)
                elif kind == 1:
                    with torch.cuda.stream(stream):
                        got = reader.batch_embedding_ids(torch.from_numpy(ids).cuda(), as_tensor=True)
                    stream.synchronize()
                    got = got.cpu().numpy()
                elif kind == 2:
                    words = [keys[i] if i >= 0 else "missing word" for i in ids]
                    got = reader.batch_embedding(words)
                else:
                    words = [keys[i] if i >= 0 else "missing word" for i in ids]
                    got = union.batch_embedding(words)
                    expect = np.concatenate([expect, want2.batch_embedding(words)], axis=-1)
                if not bits_equal(got, expect):
                    failures.append((seed, it, kind))
        except Exception as exc:  # surfaced below: an exception in a thread would otherwise pass silently
            failures.append((seed, repr(exc)))

    threads = [threading.Thread(target=worker, args=(s,)) for s in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not failures, failures


def test_host_outputs_in_several_passes(abi, oracle, synthetic):
    """Host outputs larger than the staging buffer (4 GiB of fp32 by default) are decoded in
    passes; forced down to 100 rows here, for ids, words, ranges and both union modes."""
    syns = union_case(synthetic)[:2]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    rng = np.random.default_rng(3)
    words = queries_for(syns[0], rng, 777)
    parts = [o.batch_embedding(words) for o in oracles]
    ids = rng.integers(-1, readers[0].rows, 1234).astype(np.int32)
    try:
        abi.check(abi.lib.memb_b200_set_option(b"host_pass_rows", 100))
        assert bits_equal(readers[0].lookup_words(words), parts[0])
        assert bits_equal(readers[0].decode_ids(ids), oracles[0].batch_ids(ids))
        assert bits_equal(readers[0].decode_range(5, 905), oracles[0].decode_range(5, 905))
        assert bits_equal(helpers.union_lookup_words(abi, readers, abi.AVERAGE, words, 300), np.mean(parts, axis=0))
        assert bits_equal(helpers.union_lookup_words(abi, readers, abi.CONCATENATE, words, 600),
                          np.concatenate(parts, axis=-1))
    finally:
        abi.lib.memb_b200_set_option(b"host_pass_rows", 0)


@pytest.mark.parametrize("bits0,bits1,dim,n", [(8, 4, 100, 999), (3, 5, 64, 1024), (4, 4, 12, 33), (1, 8, 300, 1500),
                                               (6, 4, 300, 700)])
def test_union_average_through_symbol_words(abi, oracle, synthetic, bits0, bits1, dim, n):
    """average of trained readers: reader 0 leaves symbol words, reader 1 gathers both codebooks
    (OutputSpec::sym); every codebook size / field format, ragged tiles, missing words on either
    side, three readers (the third adds to the fp32 sum), and the plain path as a cross-check."""
    syns = [synthetic("trained", bits0, rows=900, dim=dim, seed=51), synthetic("trained", bits1, rows=800, dim=dim, seed=52),
            synthetic("trained", 4, rows=700, dim=dim, seed=53)]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    rng = np.random.default_rng(n)
    words = sum([queries_for(s, rng, n // 3 + 1) for s in syns], [])[:n]
    parts = [o.batch_embedding(words) for o in oracles]
    for count in (2, 3):
        want = np.mean(parts[:count], axis=0)
        # default: the multi-launch ring path -- symbol words for the first two readers
        assert bits_equal(helpers.union_lookup_words(abi, readers[:count], abi.AVERAGE, words, dim), want)
        try:
            # two trained readers in ONE launch of trained_rows_kernel (three: falls back to the ring path)
            abi.check(abi.lib.memb_b200_set_option(b"trained_rows_kernel", 1))
            assert bits_equal(helpers.union_lookup_words(abi, readers[:count], abi.AVERAGE, words, dim), want)
            abi.check(abi.lib.memb_b200_set_option(b"trained_rows_kernel", 0))
            # symbol words with the two-gather flush instead of the pair table
            for flavour in (2, 1, 0):  # ONE table copy, predicate-free middle chunks; pair table through the generic flush loop; no pair table
                abi.check(abi.lib.memb_b200_set_option(b"union_pair_table", flavour))
                assert bits_equal(helpers.union_lookup_words(abi, readers[:count], abi.AVERAGE, words, dim), want)
            abi.check(abi.lib.memb_b200_set_option(b"union_pair_table", 3))
            # plain read-modify-write of the output
            abi.check(abi.lib.memb_b200_set_option(b"union_symbol_average", 0))
            assert bits_equal(helpers.union_lookup_words(abi, readers[:count], abi.AVERAGE, words, dim), want)
        finally:
            abi.lib.memb_b200_set_option(b"union_symbol_average", 1)
            abi.lib.memb_b200_set_option(b"union_pair_table", 3)
            abi.lib.memb_b200_set_option(b"trained_rows_kernel", 0)


@pytest.mark.parametrize("mode_name", ["average", "concatenate"])
def test_union_distinct_tuples_path(abi, oracle, synthetic, mode_name):
    """Batches that repeat their words: the distinct id tuples are decoded once and copied out
    (union_dedup; kernels.cuh 'repeated queries').  Same bytes as decoding every query: forced on small
    batches (all equal, all distinct, ragged sizes, missing words, three readers, mixed storages, arbitrary
    id tuples through union_decode_ids), and the automatic choice on a 150 k-word batch over a 900-word
    vocabulary."""
    dim = 64
    syns = [synthetic("trained", 4, rows=900, dim=dim, seed=61), synthetic("trained", 5, rows=800, dim=dim, seed=62),
            synthetic("uniform", 8, rows=700, dim=dim, seed=63)]
    readers = [abi.open_checked(s.path) for s in syns]
    oracles = [oracle.open(s.path) for s in syns]
    mode = abi.AVERAGE if mode_name == "average" else abi.CONCATENATE
    rng = np.random.default_rng(5)

    def merged(parts):
        return np.mean(parts, axis=0) if mode_name == "average" else np.concatenate(parts, axis=-1)

    def check_words(words, count):
        want = merged([o.batch_embedding(words) for o in oracles[:count]])
        got = helpers.union_lookup_words(abi, readers[:count], mode, words, want.shape[1])
        assert bits_equal(got, want)

    try:
        abi.check(abi.lib.memb_b200_set_option(b"union_dedup", 2))  # always
        launches = abi.lib.memb_b200_launch_count()
        for count in (2, 3):
            zipf = sum([queries_for(s, rng, 40) for s in syns], [])
            check_words([zipf[i] for i in rng.zipf(1.3, 1537) % len(zipf)], count)   # heavy repeats
            check_words([syns[0].words[7]] * 333, count)                              # one word only
            check_words(["absent"] * 65, count)                                       # missing everywhere
            check_words(sorted(set(queries_for(syns[1], rng, 500))), count)           # no repeats at all
            check_words(queries_for(syns[0], rng, 2), count)
        assert abi.lib.memb_b200_launch_count() > launches
        # two SMALL codebooks (16 and 8 centroids): the temporary holds symbol words and pair_gather_kernel merges them
        # on the way out (average only; union_dedup_symbols = 0: the fp32 temporary for the same batch)
        small = synthetic("trained", 3, rows=600, dim=dim, seed=64)
        pair_readers, pair_oracles = [readers[0], abi.open_checked(small.path)], [oracles[0], oracle.open(small.path)]
        for symbols in (1, 0):
            abi.check(abi.lib.memb_b200_set_option(b"union_dedup_symbols", symbols))
            for words in ([w for i in rng.zipf(1.2, 2111) for w in [(syns[0].words + small.words + ["absent"])[i % 1501]]],
                          [small.words[3]] * 40, queries_for(small, rng, 33)):
                want = merged([o.batch_embedding(words) for o in pair_oracles])
                assert bits_equal(helpers.union_lookup_words(abi, pair_readers, mode, words, want.shape[1]), want)
        pair_readers[1].close()
        # id tuples that no word would produce (the ids entry point takes any): equal tuples share a row, others do not
        ids = np.stack([rng.integers(-2, 12, 4000), rng.integers(-1, 9, 4000)]).astype(np.int32)
        parts = [o.batch_ids(ids[k]) for k, o in enumerate(oracles[:2])]
        assert bits_equal(helpers.union_decode_ids(abi, readers[:2], mode, ids, merged(parts).shape[1]), merged(parts))
        # automatic: a sample decides
        abi.check(abi.lib.memb_b200_set_option(b"union_dedup", 1))
        words = queries_for(syns[0], rng, 150_000)
        before = abi.lib.memb_b200_launch_count()
        check_words(words, 2)
        with_distinct_path = abi.lib.memb_b200_launch_count() - before
        abi.check(abi.lib.memb_b200_set_option(b"union_dedup", 0))
        before = abi.lib.memb_b200_launch_count()
        check_words(words, 2)
        if mode_name == "average":  # (claim x 2, follow, copy-out on top; concatenate: the copy-out costs more than it saves)
            assert with_distinct_path > abi.lib.memb_b200_launch_count() - before
        for flavour in (2, 0):  # forced; off
            abi.check(abi.lib.memb_b200_set_option(b"union_dedup", flavour))
            abi.check(abi.lib.memb_b200_set_option(b"union_dedup_keys", 0))  # generic claim for two readers too
            check_words(words[:5000], 2)
    finally:
        abi.lib.memb_b200_set_option(b"union_dedup", 1)
        abi.lib.memb_b200_set_option(b"union_dedup_keys", 1)
        abi.lib.memb_b200_set_option(b"union_dedup_symbols", 1)
        for r in readers:
            r.close()
