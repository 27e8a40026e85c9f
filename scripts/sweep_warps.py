"""Development aid: decode throughput of the headline workload vs resident warps per SM."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, memb_b200
from memb_b200 import _memb

args = bench.parse_args()
bench.build_table(args)
dev = torch.device("cuda", 0)
ids = torch.from_numpy(bench.batch_ids(args, 0)).to(dev)
out = torch.empty((args.batch, args.dim), dtype=torch.float32, device=dev)
for warps in [int(w) for w in os.environ.get("WARPS", "4,8,12,16,20,24").split(",")]:
    _memb.set_option("trained_warps", warps)
    r = memb_b200.Reader(bench.table_path(args))
    impl = r._impl
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        impl.decode_ids_into(ids.data_ptr(), True, args.batch, out.data_ptr(), True, s)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        impl.decode_ids_into(ids.data_ptr(), True, args.batch, out.data_ptr(), True, s)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print("warps %2d: %.4f ms  %.3f Grows/s" % (warps, ms, args.batch / ms / 1e6), flush=True)
    del r
