"""Generates tests/golden/codec_vectors.npz with the REFERENCE's own codec.

Run in the build container only (needs /root/reference via oracle/_ref, built by
`make -C oracle ref`):   python tests/golden/make_golden.py

For a set of prefix codes (as the serialised keys + size_offsets of
src/flatbuffers/huffman_decoder.fbs) it records what the reference's sources do:
  - createCanonicalPrefixCodes (src/prefix_code.cpp:5-22)        -> codes, lens
  - BitStream::push            (src/bit_stream.h:18-34)          -> packed bytes
  - HuffmanTableDecoder::next  (src/huffman_table_decoder.h:96-118), table widths
    L = 1, 5, 10, 12 (all must agree)                            -> decoded symbols
The committed .npz is what the CPU oracle and the CUDA kernels are pinned to on
machines where /root/reference does not exist.
"""
import heapq
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from helpers import load_ref_codec  # noqa: E402
from membfile_py import size_offsets_from_lengths  # noqa: E402


def huffman_lengths(freqs):
    """Code length per symbol of a Huffman tree over freqs (dict symbol -> weight)."""
    if len(freqs) == 1:
        return {next(iter(freqs)): 0}
    heap = [(w, i, [s]) for i, (s, w) in enumerate(freqs.items())]
    heapq.heapify(heap)
    depth = {s: 0 for s in freqs}
    tick = len(heap)
    while len(heap) > 1:
        w1, _, s1 = heapq.heappop(heap)
        w2, _, s2 = heapq.heappop(heap)
        for s in s1 + s2:
            depth[s] += 1
        heapq.heappush(heap, (w1 + w2, tick, s1 + s2))
        tick += 1
    return depth


def cases(rng):
    fib = [1, 1]
    while len(fib) < 17:
        fib.append(fib[-1] + fib[-2])
    yield "one_symbol", {7: 1.0}
    yield "two_symbols", {0: 3.0, 1: 1.0}
    yield "sixteen_flat", {s: 1.0 for s in range(16)}
    yield "gauss4", {s: float(np.exp(-0.5 * ((s - 7.5) / 3.0) ** 2)) for s in range(16)}
    yield "gauss6_pruned", {s: float(np.exp(-0.5 * ((s - 22) / 8.0) ** 2)) + 1e-3 for s in range(44)}
    yield "gauss8", {s: float(np.exp(-0.5 * ((s - 86) / 30.0) ** 2)) + 1e-4 for s in range(173)}
    yield "fib_maxlen16", {s: float(fib[s]) for s in range(17)}
    yield "all255_skewed", {s: float(1.0 / (1 + s) ** 1.6) for s in range(255)}
    yield "random40", {int(s): float(w) for s, w in zip(rng.permutation(200)[:40], rng.random(40) ** 4 + 1e-3)}


def main():
    ref = load_ref_codec()
    if ref is None:
        raise SystemExit("oracle/_ref/libmemb_ref_codec.so missing: run `make -C oracle ref` where /root/reference exists")
    rng = np.random.default_rng(20240607)
    out = {}
    names = []
    for name, freqs in cases(rng):
        depth = huffman_lengths(freqs)
        assert max(depth.values()) <= 16, name
        order = sorted(freqs, key=lambda s: (depth[s], s))
        keys = np.array(order, np.uint8)
        lens = np.array([depth[s] for s in order], np.uint8)
        so = np.array(size_offsets_from_lengths(lens.tolist()), np.uint32)
        codes, ref_lens = ref.canonical_codes(keys, so)
        assert np.array_equal(ref_lens, lens), name
        weights = np.array([freqs[s] for s in order], np.float64)
        count = 1500
        idx = rng.choice(len(order), size=count, p=weights / weights.sum())
        # make sure every symbol (hence the longest codes) occurs
        idx[:len(order)] = np.arange(len(order))
        rng.shuffle(idx)
        symbols = keys[idx]
        data = ref.bitstream_pack(codes[idx], lens[idx])
        padded = np.concatenate([data, rng.integers(0, 256, 9, dtype=np.uint8)])  # junk after the row
        decoded = None
        for width in (1, 5, 10, 12):
            got = ref.decode_stream(keys, so, width, padded, count)
            assert np.array_equal(got, symbols), (name, width)
            # and with the row being the very end of the buffer (zero padding past the end)
            got = ref.decode_stream(keys, so, width, data, count)
            assert np.array_equal(got, symbols), (name, width, "tail")
            decoded = got
        names.append(name)
        out[name + ".keys"] = keys
        out[name + ".size_offsets"] = so
        out[name + ".codes"] = codes
        out[name + ".lens"] = lens
        out[name + ".data"] = data
        out[name + ".symbols"] = decoded
    out["names"] = np.array(names)
    path = os.path.join(HERE, "codec_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(names), "codes")


if __name__ == "__main__":
    main()
