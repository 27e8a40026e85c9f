"""Development aid: configs[4] (bench.sharded_union) at one GPU under memb_b200 options.
usage: python scripts/union_probe.py [--words N] NAME=VALUE[,NAME=VALUE] ..."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
import memb_b200  # noqa: E402
from memb_b200 import _memb  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--words", type=int, default=1_000_000)
p.add_argument("settings", nargs="*", default=["trained_rows_kernel=1", "trained_rows_kernel=0"])
a = p.parse_args()
args = argparse.Namespace(rows=2_200_000, dim=300, bits=4, batch=1_000_000, union_words=a.words,
                          cache_dir=os.environ.get("MEMB_B200_BENCH_DIR", "/tmp/memb_b200_bench"))
device = torch.device("cuda", 0)
torch.cuda.set_device(0)
ctx = {"torch": torch, "dist": None, "memb_b200": memb_b200, "_memb": _memb, "rank": 0, "world": 1, "local": 0,
       "device": device, "barrier": torch.cuda.synchronize, "max_over_ranks": lambda ms: ms,
       "all_ranks_true": bool, "timer": bench.Timer(torch), "peak": bench.measured_peak()[0]}
for setting in a.settings:
    for item in setting.split(","):
        name, _, value = item.partition("=")
        _memb.set_option(name, int(value))
    r = bench.sharded_union(args, ctx)
    print(json.dumps({"setting": setting, "words": a.words, "parity": r["parity"],
                      **{m: {"ms_decode": round(r[m]["ms_decode"], 4), "ms_resolve_and_decode": round(r[m]["ms_resolve_and_decode"], 4),
                             "frac_decode": round(r[m]["roofline_frac_decode_per_gpu"], 3),
                             "every_query_ms_decode": round(r[m]["every_query_decoded"]["ms_decode"], 4),
                             "launches": r[m]["launches_per_call"]} for m in ("average", "concatenate")}}), flush=True)
