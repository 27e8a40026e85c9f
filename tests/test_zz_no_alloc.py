"""memb_b200_lookup_words_noalloc: the words lookup with its id scratch in a static __device__
array (no cudaMalloc in the call).  No counterpart in the reference.
Collected last on purpose (a failure here hides nothing)."""
import numpy as np
import pytest

import helpers
from helpers import bits_equal, c_u64, c_void_p

pytestmark = pytest.mark.gpu

MAX_WORDS = 1 << 20  # MEMB_B200_NOALLOC_MAX_WORDS, include/memb_b200.h


def _queries(syn, rng, n):
    words = [syn.words[i] for i in rng.integers(0, len(syn.words), n)]
    for i in rng.integers(0, n, max(1, n // 20)):
        words[i] = ["", "o", "zzzzzzzz", "absent%d" % i][i % 4]
    return words


def _noalloc_lookup(abi, reader, words, dim, torch):
    """The new entry point through the C ABI: every buffer is the caller's, on the device."""
    blob, offsets = helpers.marshal(words)
    d_blob = torch.frombuffer(bytearray(blob) or bytearray(1), dtype=torch.uint8).cuda()
    d_offsets = torch.from_numpy(offsets.astype(np.int64)).cuda()
    out = torch.full((len(words), dim), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    stream = torch.cuda.current_stream().cuda_stream
    abi.check(abi.lib.memb_b200_lookup_words_noalloc(reader.h, d_blob.data_ptr(), d_offsets.data_ptr(), len(words),
                                                    out.data_ptr(), stream))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return out.cpu().numpy(), free_before - free_after


@pytest.mark.parametrize("storage,bits", [("trained", 4), ("uniform", 8)])
def test_lookup_words_noalloc_matches_lookup_words(abi, synthetic, storage, bits):
    import torch

    abi.lib.memb_b200_lookup_words_noalloc.argtypes = [c_void_p, c_void_p, c_void_p, c_u64, c_void_p, c_void_p]
    syn = synthetic(storage, bits, rows=3000, dim=300)
    reader = abi.open_checked(syn.path)
    rng = np.random.default_rng(77)
    # two calls in a row share the static scratch: the second is SHORTER, so stale ids of the first
    # (beyond its n) must not matter; a third repeats the first
    first, second = _queries(syn, rng, 5000), _queries(syn, rng, 1237)
    _noalloc_lookup(abi, reader, first[:64], 300, torch)  # warm-up: kernels are loaded lazily on first launch
    for words in (first, second, first):
        want = reader.lookup_words(words)  # the existing entry point (host buffers, cudaMalloc'ed scratch)
        got, device_bytes_taken = _noalloc_lookup(abi, reader, words, 300, torch)
        assert bits_equal(got, want)  # bit-exact, like every parity test of this path
        assert device_bytes_taken <= 0  # the call took no device memory
    # beyond the static scratch: rejected with a clear error, nothing launched
    launches = abi.lib.memb_b200_launch_count()
    status = abi.lib.memb_b200_lookup_words_noalloc(reader.h, None, None, MAX_WORDS + 1, None, None)
    assert status == abi.ERR_ARGUMENT and "MEMB_B200_NOALLOC_MAX_WORDS" in abi.error()
    assert abi.lib.memb_b200_launch_count() == launches
    assert abi.lib.memb_b200_lookup_words_noalloc(reader.h, None, None, 0, None, None) == abi.OK  # empty batch
    reader.close()
