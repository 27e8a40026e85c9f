"""Pins the CPU oracle (oracle/oracle.cpp) before anything is compared against it:
the reference's own known-answer tests and test vectors, the committed golden
vectors produced by the reference's codec sources, and — where oracle/_ref was
built — the reference's headers themselves.  CPU only."""
import numpy as np
import pytest

from conftest import TEST_VECTORS
from helpers import bits_equal
import membfile_py


# ---- src/bit_stream_tests.cpp:31-59 : bit order known-answer test ----
KAT_CODES = [(1023, 14), (33, 6), (0, 4), (1234, 11), (7, 2)]


def kat_expected():
    # prettyBitString(code, n) keeps the LOW n bits (src/bit_stream_tests.cpp:7-17)
    bits = "".join(format(code & ((1 << n) - 1), "0%db" % n) for code, n in KAT_CODES)
    bits += "0" * (8 - len(bits) % 8)
    return np.array([int(bits[i:i + 8], 2) for i in range(0, len(bits), 8)], np.uint8)


def test_bitstream_known_answer(oracle, ref_codec):
    codes = [c for c, _ in KAT_CODES]
    lens = [n for _, n in KAT_CODES]
    want = kat_expected()
    got = oracle.bitstream_pack(codes, lens)
    # the reference test pads with 8 - total % 8 zero bits (37 bits -> 40)
    assert np.array_equal(got, want)
    if ref_codec is not None:
        assert np.array_equal(ref_codec.bitstream_pack(codes, lens), want)


# ---- golden vectors from the reference's codec (tests/golden/make_golden.py) ----
def test_oracle_matches_golden_codec_vectors(oracle, golden):
    for name in golden["names"]:
        keys, so = golden[name + ".keys"], golden[name + ".size_offsets"]
        codes, lens = oracle.canonical_codes(keys, so)
        assert np.array_equal(codes, golden[name + ".codes"]), name
        assert np.array_equal(lens, golden[name + ".lens"]), name
        symbols = golden[name + ".symbols"]
        idx = {int(k): i for i, k in enumerate(keys)}
        order = np.array([idx[int(s)] for s in symbols])
        assert np.array_equal(oracle.bitstream_pack(codes[order], lens[order]), golden[name + ".data"]), name
        for width in (1, 3, 10, 12, 16):
            got = oracle.decode_stream(keys, so, width, golden[name + ".data"], len(symbols))
            assert np.array_equal(got, symbols), (name, width)


def test_python_writer_codes_match_golden(golden):
    for name in golden["names"]:
        lens = golden[name + ".lens"].tolist()
        assert membfile_py.canonical_codes(lens) == golden[name + ".codes"].tolist(), name
        assert membfile_py.size_offsets_from_lengths(lens) == golden[name + ".size_offsets"].tolist(), name


def test_restatement_equals_reference_headers_fuzz(oracle, ref_codec):
    """Random prefix codes x random streams x table widths: restated decoder ==
    reference HuffmanTableDecoder (only where oracle/_ref exists)."""
    if ref_codec is None:
        pytest.skip("oracle/_ref not built on this machine")
    rng = np.random.default_rng(5)
    for trial in range(300):
        nsym = int(rng.integers(1, 256))
        # random Kraft-complete lengths: split leaves until nsym leaves exist
        leaves = [0]
        while len(leaves) < nsym:
            candidates = [i for i, d in enumerate(leaves) if d < 16]
            if not candidates:
                break
            i = candidates[int(rng.integers(len(candidates)))]
            d = leaves.pop(i)
            leaves += [d + 1, d + 1]
        lens = sorted(leaves)
        keys = rng.permutation(256)[:len(lens)].astype(np.uint8)
        so = np.array(membfile_py.size_offsets_from_lengths(lens), np.uint32)
        codes_o, lens_o = oracle.canonical_codes(keys, so)
        codes_r, lens_r = ref_codec.canonical_codes(keys, so)
        assert np.array_equal(codes_o, codes_r) and np.array_equal(lens_o, lens_r)
        count = 200
        idx = rng.integers(0, len(lens), count)
        data = ref_codec.bitstream_pack(codes_r[idx], lens_r[idx])
        assert np.array_equal(oracle.bitstream_pack(codes_o[idx], lens_o[idx]), data)
        data = np.concatenate([data, rng.integers(0, 256, 5, dtype=np.uint8)])
        width = int(rng.choice([1, 2, 5, 8, 10, 12]))
        want = ref_codec.decode_stream(keys, so, width, data, count)
        assert np.array_equal(want, keys[idx])
        assert np.array_equal(oracle.decode_stream(keys, so, width, data, count), want)


# ---- src/tests.cpp:29-57 builderTestImpl, for the three storages ----
@pytest.mark.parametrize("storage", ["full", "uniform", "trained"])
def test_reference_builder_cases(oracle, ref_oracle, reference_files, storage):
    for lib in (oracle, ref_oracle):
        if lib is None:
            continue
        reader = lib.open(reference_files[storage], 1)
        assert reader.dim == 3
        assert reader.keys() == sorted(w.encode() for w, _ in TEST_VECTORS)  # tests.cpp:36-43
        for word, vector in TEST_VECTORS:
            got = reader.word_embedding(word)
            for g, w in zip(got, vector):
                # BOOST_CHECK_CLOSE_FRACTION(.., 0.01), tests.cpp:49
                assert abs(g - w) <= 0.01 * min(abs(g), abs(w)) or g == w, (storage, word, got)
        assert np.all(reader.word_embedding("o") == 0.0)  # tests.cpp:53-56
        # beyond the last key: UB in the reference, zeros by intent
        assert np.all(reader.word_embedding("zzz") == 0.0)
        assert np.all(reader.word_embedding("") == 0.0)


def test_indirect_decoder_same_result(oracle, ref_oracle, reference_files):
    """trainedBuilderWorksWithIndirectDecoder, src/tests.cpp:76-88: table width 1."""
    words = [w for w, _ in TEST_VECTORS] + ["o"]
    for lib in (oracle, ref_oracle):
        if lib is None:
            continue
        reader = lib.open(reference_files["trained"], 1)
        base = reader.batch_embedding(words)
        for width in (1, 2, 4):
            reader.set_table_bits(width)
            assert bits_equal(reader.batch_embedding(words), base)


def test_threaded_equals_serial(oracle, reference_files):
    """threadedDecoderWorks, src/tests.cpp:90-113: 1025 words, 1 vs 4 threads."""
    batch = [TEST_VECTORS[i % len(TEST_VECTORS)][0] for i in range(1025)]
    serial = oracle.open(reference_files["trained"], 1).batch_embedding(batch)
    threaded = oracle.open(reference_files["trained"], 4).batch_embedding(batch)
    assert bits_equal(serial, threaded)


def test_missing_and_invalid_files(oracle, workdir):
    """missingFileThrows / invalidFileThrows, src/tests.cpp:134-153."""
    with pytest.raises(RuntimeError):
        oracle.open(str(workdir / "missing.bin"))
    bad = workdir / "bad.bin"
    bad.write_bytes(b"0123456789")
    with pytest.raises(RuntimeError, match="File format verification failed"):
        oracle.open(str(bad))


def test_oracle_and_reference_headers_agree_on_synthetic(oracle, ref_oracle, synthetic):
    if ref_oracle is None:
        pytest.skip("oracle/_ref not built on this machine")
    for bits in (1, 4, 8):
        syn = synthetic("trained", bits, rows=400)
        a = oracle.open(syn.path).decode_range(0, 400)
        b = ref_oracle.open(syn.path).decode_range(0, 400)
        assert ref_oracle.uses_reference_codec and not oracle.uses_reference_codec
        assert bits_equal(a, b)


def test_vectorised_packer_matches_the_bit_order_kat():
    """tests/membfile_py.pack_rows_numpy (used for the bench's 16-bit-code table) packs exactly like pack_row,
    which the bit-order KAT above pins (src/bit_stream_tests.cpp:31-59): MSB-first, zero padded to a byte."""
    import membfile_py

    rng = np.random.default_rng(3)
    keys, lengths = list(range(17)), list(range(1, 17)) + [16]
    weights = np.array([2.0 ** -n for n in lengths]) + 0.01
    symbols = rng.choice(17, (257, 301), p=weights / weights.sum()).astype(np.uint8)
    packed, offsets = membfile_py.pack_rows_numpy(symbols, keys, lengths, chunk_rows=50)
    codes = membfile_py.canonical_codes(lengths)
    code_of, len_of = dict(zip(keys, codes)), dict(zip(keys, lengths))
    for r in range(len(symbols)):
        end = int(offsets[r + 1]) if r + 1 < len(symbols) else len(packed)
        assert bytes(packed[int(offsets[r]):end]) == membfile_py.pack_row(symbols[r], code_of, len_of), r


def test_builder_kmeans_on_the_references_own_case(abi):
    """kMeansWork, src/kmeans_tests.cpp:9-38: three separated 1-D clusters, KMeansClusterizer(3).fit(data),
    predict(data) must give exactly the expected cluster ids.  Replayed against the Builder's k-means
    (memb_b200/csrc/builder.cpp, restated from src/kmeans.cpp:26-112) through memb_b200_kmeans_fit_predict --
    the one reference-held vector that pins the write path's clustering."""
    import ctypes

    data, expected = [], []
    for i in range(8):   # MIDDLE_CLUSTER_SIZE
        data.append(-0.5 + i * 0.125)
        expected.append(1)
    for i in range(16):  # FIRST_CLUSTER_SIZE
        data.append(-9 + i * 0.125)
        expected.append(0)
    for i in range(4):   # LAST_CLUSTER_SIZE
        data.append(11.75 + i * 0.125)
        expected.append(2)
    values = np.array(data, np.float32)
    ids = np.zeros(len(values), np.uint8)
    centroids = np.zeros(3, np.float32)
    count = ctypes.c_uint32(0)
    fn = abi.lib.memb_b200_kmeans_fit_predict
    fn.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p,
                   ctypes.POINTER(ctypes.c_uint32)]
    abi.check(fn(values.ctypes.data, len(values), 3, ids.ctypes.data, centroids.ctypes.data, ctypes.byref(count)))
    assert ids.tolist() == expected
    assert count.value == 3 and np.all(np.diff(centroids) > 0)
    # the centroids are the cluster means (running-mean update, src/kmeans.cpp:86-102), to float rounding
    means = [np.mean(values[8:24]), np.mean(values[:8]), np.mean(values[24:])]
    assert np.allclose(centroids, means, rtol=1e-5)
