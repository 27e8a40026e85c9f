"""Development aid: GatheredMatrix (multicast decode -> all ranks) on a small table.
timeout 120 torchrun --nproc-per-node N scripts/fused_gather_probe.py"""
import argparse
import faulthandler
import os
import sys
import time

faulthandler.dump_traceback_later(60, exit=True)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import bench
import memb_b200
from memb_b200.sharding import GatheredMatrix

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rows = int(os.environ.get("PROBE_ROWS", "200000"))
args = argparse.Namespace(rows=rows, dim=300, bits=4, batch=1000, cache_dir="/tmp/memb_b200_bench")
if rank == 0:
    bench.build_table(args)
dist.barrier()
reader = memb_b200.Reader(bench.table_path(args), device=local)
print(rank, "opened", flush=True)
g = GatheredMatrix(reader)
print(rank, "rendezvous done, mc", hex(g.handle.multicast_ptr), flush=True)
faulthandler.dump_traceback_later(40, exit=True)
g.handle.barrier(channel=0)
torch.cuda.synchronize()
print(rank, "barrier 0 ok", flush=True)
g.handle.barrier(channel=1)
torch.cuda.synchronize()
print(rank, "barrier 1 ok", flush=True)
from memb_b200.sharding import shard_range
lo, hi = shard_range(reader.rows, rank, world)
# plain decode into the LOCAL symmetric buffer slot first
reader._impl.decode_range_into(lo, hi, g.buffer[rank * g.job:].data_ptr(), True, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(rank, "local decode into symmetric memory ok", flush=True)
# a torch write through the multicast pointer? (not exposed) -> our kernel
reader._impl.decode_range_multicast(lo, hi, g.handle.multicast_ptr + rank * g.job * 300 * 4, torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(rank, "multicast decode ok", flush=True)
stream = torch.cuda.current_stream().cuda_stream
slot = g.handle.multicast_ptr + rank * g.job * 300 * 4
faulthandler.dump_traceback_later(25, exit=True)
reader._impl.decode_range_multicast(lo, hi, slot, stream)
dist.barrier()
torch.cuda.synchronize()
print(rank, "K + nccl barrier ok", flush=True)
faulthandler.dump_traceback_later(40, exit=True)
out = g.fill()
torch.cuda.synchronize()
print(rank, "filled", flush=True)
whole = reader.decode_range(0, reader.rows, as_tensor=True)
for r in range(world):
    a_, b_ = r * g.job, min(reader.rows, (r + 1) * g.job)
    bad = int((out[a_:b_].view(torch.int32) != whole[a_:b_].view(torch.int32)).sum().item())
    print(rank, "slot", r, "mismatching elements", bad, "of", (b_ - a_) * 300, flush=True)
time.sleep(0.5)
print(rank, "equal after 0.5 s", bool(torch.equal(out.view(torch.int32), whole.view(torch.int32))), flush=True)
# unicast P2P: the ordinary decode with the PEER's buffer as output
peer = (rank + 1) % world
peer_slot = g.handle.buffer_ptrs[peer] + rank * g.job * 300 * 4
g.buffer.zero_()
dist.barrier()
torch.cuda.synchronize()
stream = torch.cuda.current_stream().cuda_stream
reader._impl.decode_range_into(lo, hi, peer_slot, True, stream)
reader._impl.decode_range_into(lo, hi, g.buffer[rank * g.job:].data_ptr(), True, stream)
torch.cuda.synchronize()
dist.barrier()
torch.cuda.synchronize()
print(rank, "p2p equal", bool(torch.equal(out.view(torch.int32), whole.view(torch.int32))), flush=True)
tm = bench.Timer(torch)
dist.barrier()
ms_p2p = tm.ms(lambda: reader._impl.decode_range_into(lo, hi, peer_slot, True, stream), reps=5, warm=2)
ms_loc = tm.ms(lambda: reader._impl.decode_range_into(lo, hi, g.buffer[rank * g.job:].data_ptr(), True, stream), reps=5, warm=2)
print(rank, "decode into peer ms", ms_p2p, "GB/s", (hi - lo) * 1200 / (ms_p2p / 1e3) / 1e9, "local ms", ms_loc, flush=True)
print(rank, "equal", bool(torch.equal(out.view(torch.int32), whole.view(torch.int32))), flush=True)
timer = bench.Timer(torch)
dist.barrier()
ms = timer.ms(g.fill, reps=5, warm=2)
print(rank, "fused ms", ms, "GB/s in", rows * 300 * 4 * (world - 1) / world / (ms / 1e3) / 1e9, flush=True)
dist.barrier()
dist.destroy_process_group()
