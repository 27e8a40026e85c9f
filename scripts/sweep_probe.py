"""Development aid: time the trained decode of the configs[3] sweep tables under memb_b200 options.
usage: python scripts/sweep_probe.py --bits 5 8 --option trained_lut_replicas=1 [--deep]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    p = argparse.ArgumentParser()
    p.add_argument("--bits", type=int, nargs="*", default=[5, 6, 7, 8])
    p.add_argument("--deep", action="store_true")
    p.add_argument("--option", action="append", default=[])
    p.add_argument("--rows", type=int, default=2_000_000)
    a = p.parse_args()
    import torch

    import memb_b200
    from memb_b200 import _memb

    args = argparse.Namespace(rows=a.rows, dim=300, bits=4, batch=1_000_000,
                              cache_dir=os.environ.get("MEMB_B200_BENCH_DIR", "/tmp/memb_b200_bench"))
    for item in a.option:
        name, _, value = item.partition("=")
        _memb.set_option(name, int(value))
    device = torch.device("cuda", 0)
    timer = bench.Timer(torch)
    peak = bench.measured_peak()[0]
    stream = torch.cuda.current_stream().cuda_stream
    buf = torch.empty((a.rows, 300), dtype=torch.float32, device=device)
    paths = []
    for bits in a.bits:
        bench.build_table(args, "trained", bits, seed=5, rows=a.rows)
        paths.append(("trained_%dbit" % bits, bench.table_path(args, "trained", bits, 5, a.rows)))
    if a.deep:
        paths.append(("deep16", bench.build_deep_table(args)))
    for label, path in paths:
        rd = memb_b200.Reader(path, device=0)
        st = rd.load_stats()
        rows = rd.rows
        ids_np = np.random.Generator(np.random.PCG64(41)).integers(0, rows, args.batch).astype(np.int32)
        ids = torch.from_numpy(ids_np).to(device)
        n = args.batch
        ms_ids = timer.ms(lambda: rd._impl.decode_ids_into(ids.data_ptr(), True, n, buf.data_ptr(), True, stream), reps=10)
        torch.cuda.synchronize()
        oracle = bench.CpuOracle(path, threads=0)
        ok = bench.same_bits(oracle.rows_of(ids_np[:20000], 300), buf[:20000].cpu().numpy())
        ms_all = timer.ms(lambda: rd._impl.decode_range_into(0, rows, buf.data_ptr(), True, stream), reps=10)
        per_row = 1204 + st["payload_bytes"] / rows
        print(json.dumps({"table": label, "options": a.option, "lut_bits": st["lut_bits"], "replicas": st["lut_replicas"],
                          "max_len": st["max_code_len"], "random_ms": round(ms_ids, 4),
                          "random_frac": round(n * (per_row + 4) / (ms_ids / 1e3) / 1e9 / peak, 3),
                          "full_ms": round(ms_all, 4), "full_frac": round(rows * per_row / (ms_all / 1e3) / 1e9 / peak, 3),
                          "parity": ok}), flush=True)
        del rd


if __name__ == "__main__":
    main()
