"""K2 / K3 probe: time (or, under ncu, just launch) the uniform / full kernels on the bench's
synthetic tables.  usage: python scripts/k2_probe.py [--bits 8] [--storage uniform] [--reps 20]
Prints one line per case: ms, rows/s, fraction of MEASURED_PEAKS.json hbm_gbs."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--bits", type=int, default=8)
    p.add_argument("--storage", default="uniform")
    p.add_argument("--rows", type=int, default=2_000_000)
    p.add_argument("--reps", type=int, default=20)
    p.add_argument("--rows-kernel", type=int, default=1, help="0: element walk, 1: auto, 2: uniform_rows_kernel")
    a = p.parse_args()
    import torch

    import memb_b200

    from memb_b200 import _memb
    _memb.set_option("uniform_rows_kernel", a.rows_kernel)
    sys.argv = [sys.argv[0]]
    args = bench.parse_args()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    peak = bench.measured_peak()[0]
    seed = 5 if a.storage == "uniform" else 9
    bits = a.bits if a.storage == "uniform" else 32
    bench.build_table(args, a.storage, bits, seed=seed, rows=a.rows)
    rd = memb_b200.Reader(bench.table_path(args, a.storage, bits, seed, a.rows), device=0)
    n, dim = 1_000_000, args.dim
    ids = torch.from_numpy(np.random.Generator(np.random.PCG64(41)).integers(0, a.rows, n).astype(np.int32)).to(device)
    buf = torch.empty((max(n, a.rows), dim), dtype=torch.float32, device=device)
    stream = torch.cuda.current_stream().cuda_stream
    per_row = 4 * dim + dim + 8 if a.storage == "uniform" else 8 * dim
    ms = bench._time_launches(lambda: rd._impl.decode_ids_into(ids.data_ptr(), True, n, buf.data_ptr(), True, stream), a.reps, 2)
    print("%s %d random: %.4f ms  %.3f G rows/s  frac %.3f" % (a.storage, bits, ms, n / ms / 1e6, n * (per_row + 4) / (ms / 1e3) / 1e9 / peak))
    ms = bench._time_launches(lambda: rd._impl.decode_range_into(0, a.rows, buf.data_ptr(), True, stream), a.reps, 2)
    print("%s %d full:   %.4f ms  %.3f G rows/s  frac %.3f" % (a.storage, bits, ms, a.rows / ms / 1e6, a.rows * per_row / (ms / 1e3) / 1e9 / peak))


if __name__ == "__main__":
    main()
