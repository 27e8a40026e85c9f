"""pytest configuration: `gpu` marker, in-tree native build, shared fixtures."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for path in (ROOT, os.path.join(ROOT, "tests")):
    if path not in sys.path:
        sys.path.insert(0, path)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")
    # build everything in-tree once (no-op when the .so files are up to date)
    import build_native

    build_native.build_all()


def _has_gpu():
    try:
        import memb_b200

        return memb_b200._memb.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no sm_100 GPU visible")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def _oracle_libs():
    from helpers import load_oracle, load_ref_oracle

    return load_oracle(), load_ref_oracle()


@pytest.fixture
def oracle(request, _oracle_libs):
    """The checker.  CPU tests get the restatement (oracle/liboracle.so; tests/test_oracle_pins.py ties it to the
    reference's headers); GPU parity tests get the oracle whose Huffman decode IS the reference's own
    huffman_table_decoder.h / bit_stream_reader.h (oracle/_ref/liboracle_ref.so) whenever that library travelled."""
    restated, reference_headers = _oracle_libs
    if reference_headers is not None and request.node.get_closest_marker("gpu"):
        return reference_headers
    return restated


@pytest.fixture(scope="session")
def ref_oracle():
    from helpers import load_ref_oracle

    return load_ref_oracle()


@pytest.fixture(scope="session")
def ref_codec():
    from helpers import load_ref_codec

    return load_ref_codec()


@pytest.fixture(scope="session")
def abi():
    from helpers import CAbi

    return CAbi()


@pytest.fixture(scope="session")
def golden():
    from helpers import GOLDEN

    return np.load(os.path.join(GOLDEN, "codec_vectors.npz"))


@pytest.fixture(scope="session")
def workdir(tmp_path_factory):
    return tmp_path_factory.mktemp("memb")


# The reference's own test vectors, src/tests.cpp:20-27.
TEST_VECTORS = [
    ("the", [0.0, 1.0, 2.0]),
    ("of", [0.0, -1.0, 2.0]),
    ("th", [2.0, 0.0, 1.0]),
    ("a", [1.0, 0.0, -2.0]),
    ("tho", [2.0, 0.0, -1.0]),
    ("abc", [-2.0, 0.0, 1.0]),
]


@pytest.fixture(scope="session")
def reference_files(workdir):
    """The files of builderTestImpl (src/tests.cpp:29-43): 6 words x dim 3, 8 bits,
    one per storage type, written through the Builder API in insertion order."""
    import memb_b200

    paths = {}
    for storage in ("full", "uniform", "trained"):
        builder = memb_b200.Builder(3, storage, 8)
        for word, vector in TEST_VECTORS:
            builder.add_word(word, np.array(vector, np.float32))
        path = str(workdir / ("ref_%s.bin" % storage))
        builder.save(path)
        paths[storage] = path
    return paths


class Synthetic:
    def __init__(self, path, words, matrix, storage, bits):
        self.path, self.words, self.matrix, self.storage, self.bits = path, words, matrix, storage, bits


@pytest.fixture(scope="session")
def synthetic(workdir):
    """Factory of cached synthetic files: synthetic(storage, bits, rows, dim, seed, scale)."""
    from helpers import build_file, synthetic_words

    cache = {}

    def make(storage, bits, rows=3000, dim=300, seed=1, scale=1.0):
        key = (storage, bits, rows, dim, seed, scale)
        if key not in cache:
            rng = np.random.default_rng(seed)
            words = synthetic_words(rows, seed)
            matrix = (rng.standard_normal((rows, dim), dtype=np.float32) * np.float32(scale)).astype(np.float32)
            path = str(workdir / ("syn_%s_%d_%d_%d_%d.bin" % (storage, bits, rows, dim, seed)))
            build_file(path, storage, bits, words, matrix)
            cache[key] = Synthetic(path, words, matrix, storage, bits)
        return cache[key]

    return make
