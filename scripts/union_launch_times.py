"""Development aid: one ReadersUnion average + concatenate call on the bench tables, meant to be run
under `ncu --metrics gpu__time_duration.sum` to see the per-launch times."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
import memb_b200  # noqa: E402
from memb_b200 import _memb  # noqa: E402

sys.argv = [sys.argv[0]]
args = bench.parse_args()
bench.build_table(args)
bench.build_table(args, "trained", args.bits, seed=7)
a = memb_b200.Reader(bench.table_path(args))
b = memb_b200.Reader(bench.table_path(args, "trained", args.bits, 7))
n = args.batch
perm = np.random.Generator(np.random.PCG64(88)).permutation(args.rows)
zipf = perm[bench.zipf_ranks(args.rows, n, 8)].astype(np.int32)
ids = torch.from_numpy(np.concatenate([zipf, zipf])).cuda()
for mode, cols in (("average", 300), ("concatenate", 600)):
    out = torch.empty((n, cols), dtype=torch.float32, device="cuda")
    for _ in range(2):
        _memb.union_decode_ids_into([a._impl, b._impl], mode, ids.data_ptr(), True, n, out.data_ptr(), True, 0)
    torch.cuda.synchronize()
