"""Development aid: whole-matrix equality of the range decode with bulk-async staging on and off."""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import memb_b200
from memb_b200 import _memb

args = argparse.Namespace(rows=int(os.environ.get("PROBE_ROWS", "2200000")), dim=300, bits=4, batch=1000, cache_dir="/tmp/memb_b200_bench")
bench.build_table(args)
reader = memb_b200.Reader(bench.table_path(args), device=0)
rows = reader.rows
_memb.set_option("trained_tma", 0)
ref = reader.decode_range(0, rows, as_tensor=True)
for opt in (1, 2):
    _memb.set_option("trained_tma", opt)
    for trial in range(3):
        got = reader.decode_range(0, rows, as_tensor=True)
        bad = (got.view(torch.int32) != ref.view(torch.int32))
        n = int(bad.sum().item())
        print("tma", opt, "trial", trial, "mismatching elements", n, flush=True)
        if n:
            rows_bad = bad.any(dim=1).nonzero().flatten()
            print("   bad rows", rows_bad.numel(), "first", rows_bad[:10].tolist(), "tiles", sorted(set((rows_bad[:2000] // 32).tolist()))[:10])
            r0 = int(rows_bad[0]); cols = bad[r0].nonzero().flatten()
            print("   row", r0, "bad cols", cols[:12].tolist(), "count", cols.numel())
ids = torch.arange(rows, dtype=torch.int32, device="cuda")
got = reader.batch_embedding_ids(ids, as_tensor=True)
print("ids tma=2 mismatches", int((got.view(torch.int32) != ref.view(torch.int32)).sum().item()))
_memb.set_option("trained_tma", 1)
got = reader.decode_range(0, rows, as_tensor=True)
bad = (got.view(torch.int32) != ref.view(torch.int32))
rows_bad = bad.any(dim=1).nonzero().flatten()
for r0 in rows_bad[:3].tolist():
    print("row", r0, "got", [round(float(x), 4) for x in got[r0, :8]], "ref", [round(float(x), 4) for x in ref[r0, :8]])
    # is the wrong prefix the prefix of another row (stale buffer)?  look for rows whose first 6 values equal got's
    key = got[r0, :6]
    hits = (ref[:, :6] == key).all(dim=1).nonzero().flatten()
    print("   rows of the table starting with got's first 6 values:", hits[:8].tolist(), "delta", [(h - r0) for h in hits[:8].tolist()])
