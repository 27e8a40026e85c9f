"""bench.py — the headline measurement of the batched embedding lookup.

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): GloVe-840B-shaped synthetic table, 2.2 M x 300,
`trained` 4-bit (k-means codebook + Huffman), one step = one batched lookup of 1 M
random row ids (uniform with replacement, seed 4) -> (1 M, 300) fp32.  At N>1 every
rank holds a replica of the compressed table and decodes its own 1 M-id batch
(queries shard, no data-path collective): weak scaling.

One JSON line on stdout (rank 0).  `value` = rows/s with ids and output resident in
HBM; `e2e` = the same lookup through the C ABI with HOST buffers (pinned ids in,
pinned fp32 rows out), copies inside the timed region.  `roofline` is for the
dominant kernel (trained_decode_kernel); `cpu_baseline` times the CPU oracle (whose
Huffman decode is the reference's own header when oracle/_ref was built) on the host
cores.  The oracle is only ever the thing compared against / the CPU baseline here.
"""
import argparse
import ctypes
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

_JSON_OUT = sys.stdout
METRIC = "decoded rows/sec (fp32 output GB/s vs HBM peak) of the batched embedding lookup"
UNIT = "rows/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=200)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--rows", type=int, default=2_200_000, help="vocabulary rows of the synthetic table")
    p.add_argument("--dim", type=int, default=300)
    p.add_argument("--bits", type=int, default=4)
    p.add_argument("--batch", type=int, default=1_000_000, help="row ids per step and per GPU")
    p.add_argument("--e2e-steps", type=int, default=0, help="steps of the host-buffer leg (0 = auto)")
    p.add_argument("--no-extra", action="store_true", help="skip the secondary measurements")
    p.add_argument("--ids", default="random", choices=["random", "sequential", "blocks32", "misses5"],
                   help="experiment: id distribution of the batch (the headline is `random`; misses5 = random "
                        "with 5 %% of the ids -1, i.e. out-of-vocabulary words)")
    p.add_argument("--trained-warps", type=int, default=0, help="experiment: cap resident warps per SM (0 = auto)")
    p.add_argument("--quick", action="store_true",
                   help="experiment: device-timed leg + a 20 k-row parity sample only (no e2e, no CPU baseline, no extras)")
    p.add_argument("--option", action="append", default=[], metavar="NAME=VALUE",
                   help="experiment: memb_b200_set_option before the table is opened (e.g. trained_rows_kernel=0)")
    p.add_argument("--cache-dir", default=os.environ.get("MEMB_B200_BENCH_DIR", "/tmp/memb_b200_bench"))
    return p.parse_args()


# ------------------------------------------------------------- workload ----
def table_path(args, storage="trained", bits=None, seed=3, rows=None):
    bits = args.bits if bits is None else bits
    rows = args.rows if rows is None else rows
    tag = "" if seed == 3 else "_s%d" % seed
    return os.path.join(args.cache_dir, "glove_shape_%d_x_%d_%s%d%s.bin" % (rows, args.dim, storage, bits, tag))


def build_table(args, storage="trained", bits=None, seed=3, rows=None):
    """Synthetic GloVe-840B-shaped file through the repo's Builder (the memb_converter
    equivalent): N(0,1) float32 vectors (PCG64 seed `seed` per chunk), unique words in
    shuffled insertion order.  Cached on disk; returns build seconds (0 when cached)."""
    bits = args.bits if bits is None else bits
    rows = args.rows if rows is None else rows
    path = table_path(args, storage, bits, seed, rows)
    if os.path.exists(path):
        return 0.0
    import memb_b200
    from concurrent.futures import ThreadPoolExecutor

    t0 = time.time()
    os.makedirs(args.cache_dir, exist_ok=True)
    dim = args.dim
    matrix = np.empty((rows, dim), np.float32)
    chunks = max(1, min(32, os.cpu_count() or 1))
    bounds = np.linspace(0, rows, chunks + 1).astype(np.int64)

    def fill(i):
        rng = np.random.Generator(np.random.PCG64([seed, i]))
        rng.standard_normal(out=matrix[bounds[i]:bounds[i + 1]], dtype=np.float32)

    with ThreadPoolExecutor(chunks) as pool:
        list(pool.map(fill, range(chunks)))
    order = np.random.Generator(np.random.PCG64(30 + seed)).permutation(rows)
    words = ["w%x" % i for i in order]
    builder = memb_b200.Builder(dim, storage, bits)
    builder.add_words(words, matrix)
    tmp = path + ".tmp%d" % os.getpid()
    builder.save(tmp)
    os.replace(tmp, path)
    return time.time() - t0


def batch_ids(args, rank):
    rng = np.random.Generator(np.random.PCG64([4, rank]))
    if getattr(args, "ids", "random") == "sequential":
        return (np.arange(args.batch, dtype=np.int64) % args.rows).astype(np.int32)
    if getattr(args, "ids", "random") == "blocks32":
        starts = rng.integers(0, args.rows // 32, (args.batch + 31) // 32, dtype=np.int64) * 32
        return (starts[:, None] + np.arange(32)[None, :]).reshape(-1)[:args.batch].astype(np.int32)
    ids = rng.integers(0, args.rows, args.batch, dtype=np.int64).astype(np.int32)
    if getattr(args, "ids", "random") == "misses5":
        ids[rng.random(args.batch) < 0.05] = -1
    return ids


# ------------------------------------------------------------- oracle (CPU baseline only) ----
class CpuOracle:
    """ctypes view of oracle/_ref/liboracle_ref.so (reference Huffman headers) or, when
    that was not built, oracle/liboracle.so (the restatement)."""

    def __init__(self, path, threads=0):
        ref = os.path.join(ROOT, "oracle", "_ref", "liboracle_ref.so")
        port = os.path.join(ROOT, "oracle", "liboracle.so")
        self.uses_reference_codec = os.path.exists(ref)
        lib = ctypes.CDLL(ref if self.uses_reference_codec else port)
        lib.oracle_open.restype = ctypes.c_void_p
        lib.oracle_open.argtypes = [ctypes.c_char_p, ctypes.c_uint64]
        lib.oracle_threads.restype = ctypes.c_uint64
        lib.oracle_threads.argtypes = [ctypes.c_void_p]
        lib.oracle_batch_ids.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_close.argtypes = [ctypes.c_void_p]
        self.lib = lib
        self.h = lib.oracle_open(path.encode(), threads)
        if not self.h:
            raise RuntimeError("oracle cannot open " + path)
        self.threads = int(lib.oracle_threads(self.h))

    def batch_ids(self, ids, out):
        self.lib.oracle_batch_ids(self.h, ids.ctypes.data, len(ids), out.ctypes.data)

    def describe(self):
        codec = ("reference huffman_table_decoder.h/bit_stream_reader.h compiled from /root/reference (oracle/_ref)"
                 if self.uses_reference_codec else "restated codec (oracle/oracle.cpp)")
        return "CPU oracle Reader: %s + restated lookup/threading of src/reader.cpp:59-86" % codec


def time_cpu(oracle, ids, dim, steps, warmup, budget_s=25.0):
    """Seconds per pass over `ids` (best and mean over <= steps passes within budget)."""
    out = np.empty((len(ids), dim), np.float32)
    for _ in range(warmup):
        oracle.batch_ids(ids, out)
    times, start = [], time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        oracle.batch_ids(ids, out)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - start > budget_s:
            break
    return times, out


# ------------------------------------------------------------- clocks ----
class ClockSampler:
    """Samples SM clock and throttle reasons DURING the timed region (NVML, ~every 5 ms)."""
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, device_index):
        self.samples, self.reasons, self.power = [], set(), []
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            index = device_index
            if visible:
                entry = visible.split(",")[device_index]
                self.handle = (pynvml.nvmlDeviceGetHandleByUUID(entry) if entry.startswith("GPU-")
                               else pynvml.nvmlDeviceGetHandleByIndex(int(entry)))
            else:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
        except Exception as exc:  # noqa: BLE001
            self.nvml = None
            self.error = str(exc)

    def _loop(self):
        nv = self.nvml
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
                self.power.append(nv.nvmlDeviceGetPowerUsage(self.handle) / 1000.0)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.005)

    def start(self):
        if self.nvml:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        if self._thread:
            self._stop.set()
            self._thread.join()
        if not self.nvml or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0,
                    "note": getattr(self, "error", "no samples")}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples),
                "power_w_max": max(self.power) if self.power else None}


def workload_config(args, extra=None):
    config = {
        "workload": "configs[1]: GloVe-840B-shaped synthetic %d x %d, trained %d-bit; random batched lookup of %d "
                    "row ids per step per GPU" % (args.rows, args.dim, args.bits, args.batch),
        "rows": args.rows, "dim": args.dim, "bits_per_weight": args.bits, "batch_ids": args.batch,
        "ids": "uniform with replacement, PCG64 seed [4, rank]",
        "l2": "inputs larger than L2: every step writes %.0f MB and gathers from a %s compressed table; no flush "
              "needed" % (args.batch * args.dim * 4 / 1e6, "~290 MB" if args.rows >= 2_000_000 else "small"),
        "parallelism": "replicated table, queries sharded by rank, no collective",
    }
    config.update(extra or {})
    return config


# ------------------------------------------------------------- reference arm ----
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 measures the CPU implementation
    build_table(args)
    oracle = CpuOracle(table_path(args), threads=0)
    ids = batch_ids(args, 0)
    # bounded sample: at most ~25 s per phase; a step is the full batch unless it is too slow
    sample = len(ids)
    t0 = time.perf_counter()
    probe = np.empty((min(sample, 20000), args.dim), np.float32)
    oracle.batch_ids(ids[:len(probe)], probe)
    per_row = (time.perf_counter() - t0) / len(probe)
    budget_rows = int(60.0 / max(per_row, 1e-12) / max(args.steps + args.warmup, 1))
    if budget_rows < sample:
        sample = max(1024, budget_rows)
    times, _ = time_cpu(oracle, ids[:sample], args.dim, args.steps, args.warmup, budget_s=120.0)
    ms = 1e3 * float(np.mean(times))
    value = sample / (ms / 1e3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8 symbols -> f32 codebook gather", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": oracle.threads, "kind": "port",
                         "codec": "reference headers" if oracle.uses_reference_codec else "restated",
                         "sample": "%d of the %d ids of one step, %d timed passes; %s; host has %d logical cpus"
                                   % (sample, len(ids), len(times), oracle.describe(), os.cpu_count() or 0)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "output_gb_per_s": value * args.dim * 4 / 1e9,
    }
    print(json.dumps(line), file=_JSON_OUT, flush=True)


# ------------------------------------------------------------- B200 arm ----
class near_gpu:
    """While active, the process runs on the CPUs NVML names as local to the GPU, so the
    pinned host buffers allocated inside land on that GPU's NUMA node (host-buffer copies of
    N ranks otherwise cross the socket interconnect).  Leaves the affinity as it was."""

    def __init__(self, device_index):
        self.device_index = device_index
        self.cpus = None
        self.saved = None

    def __enter__(self):
        try:
            import pynvml
            import torch

            pynvml.nvmlInit()
            uuid = str(torch.cuda.get_device_properties(self.device_index).uuid)
            try:
                handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
            except Exception:
                handle = pynvml.nvmlDeviceGetHandleByIndex(self.device_index)
            allowed = os.sched_getaffinity(0)
            words = (max(allowed) + 64) // 64
            mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
            local = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (int(w) >> b) & 1}
            cpus = local & allowed
            if cpus and cpus != allowed:
                self.saved = allowed
                os.sched_setaffinity(0, cpus)
            self.cpus = sorted(cpus)
        except Exception as exc:  # no NVML, no affinity call: run unbound
            print("near_gpu: not bound (%s)" % exc, file=sys.stderr)
        return self

    def __exit__(self, *exc):
        if self.saved is not None:
            os.sched_setaffinity(0, self.saved)
        return False


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy, kernel timed alone)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)"


def ncu_traffic(args):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the decode kernel from the
    committed `ncu --set full` capture (profiles/), valid for the default workload only."""
    try:
        with open(os.path.join(ROOT, "profiles", "k1_ncu_latest.json")) as f:
            rec = json.load(f)
        if (rec["rows"], rec["dim"], rec["bits"], rec["batch"]) == (args.rows, args.dim, args.bits, args.batch):
            return {"bytes_per_launch": rec["dram_bytes_read"] + rec["dram_bytes_write"], "unit": "B",
                    "dram_bytes_read": rec["dram_bytes_read"], "dram_bytes_write": rec["dram_bytes_write"],
                    "source": rec["source"]}
    except Exception:  # noqa: BLE001
        pass
    return None


def run_b200(args):
    import torch
    import torch.distributed as dist

    import memb_b200
    from memb_b200 import _memb

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d: launch with torchrun --nproc-per-node %d" %
                         (args.gpus, world, args.gpus))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU path in memb_b200")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    build_s = build_table(args) if rank == 0 else 0.0
    barrier()

    if args.trained_warps:
        _memb.set_option("trained_warps", args.trained_warps)
    for item in args.option:
        name, _, value = item.partition("=")
        _memb.set_option(name, int(value))
    t0 = time.perf_counter()
    reader = memb_b200.Reader(table_path(args), device=local)
    open_s = time.perf_counter() - t0
    stats = reader.load_stats()
    impl = reader._impl
    rows, dim, n = reader.rows, reader.dim, args.batch

    with near_gpu(local) as binding:
        ids_host = torch.from_numpy(batch_ids(args, rank)).pin_memory()
    ids_dev = ids_host.to(device)
    out_dev = torch.empty((n, dim), dtype=torch.float32, device=device)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device():
        impl.decode_ids_into(ids_dev.data_ptr(), True, n, out_dev.data_ptr(), True, stream)

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    launches0 = _memb.launch_count()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler.start()
    start.record()
    for _ in range(args.steps):
        step_device()
    stop.record()
    barrier()
    clocks = sampler.stop()
    launches = _memb.launch_count() - launches0
    ms_total = max_over_ranks(start.elapsed_time(stop))
    ms_step = ms_total / args.steps
    value = world * n / (ms_step / 1e3)

    if args.quick:
        if rank == 0:
            oracle = CpuOracle(table_path(args), threads=0)
            sample = ids_host.numpy()[:20000]
            want = np.empty((len(sample), dim), np.float32)
            oracle.batch_ids(sample, want)
            step_device()
            torch.cuda.synchronize()
            ok = bool(np.array_equal(want.view(np.uint32), out_dev[:len(sample)].cpu().numpy().view(np.uint32)))
            payload_per_row = stats["payload_bytes"] / rows
            frac = n * (4 * dim + payload_per_row + 8) / (ms_step / 1e3) / 1e9 / measured_peak()[0]
            print(json.dumps({"quick": True, "ids": args.ids, "options": args.option, "ms_per_step": ms_step,
                              "rows_per_s": value, "roofline_frac": frac, "parity_sample": ok, "clocks": clocks,
                              "gpu_launches": int(launches)}), file=_JSON_OUT, flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- e2e: the C-ABI call with HOST buffers (pinned), H2D ids + D2H rows inside the timed region ----
    with near_gpu(local):
        out_host = torch.empty((n, dim), dtype=torch.float32).pin_memory()
    e2e_steps = args.e2e_steps or max(3, min(args.steps, 20))

    def step_host():
        impl.decode_ids_into(ids_host.data_ptr(), False, n, out_host.data_ptr(), False, 0)

    step_host()
    step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()  # returns when out_host is complete
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / e2e_steps
    e2e_value = world * n / (e2e_ms / 1e3)

    # ---- all ranks: configs[2], full-vocabulary decode sharded by row range (+ optional all-gather) ----
    sharded = None if args.no_extra else sharded_full_vocab(reader, device, world, rank, barrier, max_over_ranks)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0: parity spot check, roofline, cpu baseline, extras ----
    peak, peak_source = measured_peak()
    payload_per_row = stats["payload_bytes"] / rows
    bytes_per_row = 4 * dim + payload_per_row + 4 + 4  # fp32 out + compressed row + value_offset + id
    achieved = n * bytes_per_row / (ms_step / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": "trained_decode_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": ncu_traffic(args), "peak_source": peak_source,
                "algorithmic_bytes_per_row": bytes_per_row,
                "write_only_frac": n * 4 * dim / (ms_step / 1e3) / 1e9 / peak,
                "launch_ms": ms_step}

    oracle = CpuOracle(table_path(args), threads=0)
    sample = min(n, 200_000) if (os.cpu_count() or 1) < 32 else n
    cpu_times, cpu_out = time_cpu(oracle, ids_host.numpy()[:sample], dim, steps=5, warmup=1, budget_s=20.0)
    parity = bool(np.array_equal(cpu_out.view(np.uint32), out_host.numpy()[:sample].view(np.uint32)))
    cpu_value = sample / float(np.min(cpu_times))
    cpu_baseline = {"value": cpu_value, "unit": UNIT, "cores": oracle.threads, "kind": "port",
                    "codec": "reference headers" if oracle.uses_reference_codec else "restated",
                    "sample": "first %d ids of the step's batch, best of %d passes; %s; host has %d logical cpus"
                              % (sample, len(cpu_times), oracle.describe(), os.cpu_count() or 0)}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 symbols -> f32 codebook gather", "data": "synthetic",
        "config": workload_config(args),
        "output_gb_per_s": value * dim * 4 / 1e9,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": n * 4, "d2h_bytes_per_step": n * dim * 4,
                "ms_per_step": e2e_ms, "steps": e2e_steps,
                "api": "memb_b200_decode_ids(host ids -> host rows) via _memb.Reader.decode_ids_into",
                "host_buffers": "pinned, allocated on the GPU's NUMA node (%s local cpus of %d allowed)"
                                % (len(binding.cpus) if binding.cpus is not None else "?", len(os.sched_getaffinity(0)))},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "parity_vs_oracle": parity,
        "load": {"table_build_s": build_s, "open_s": open_s, "parse_ms": stats["parse_ms"],
                 "upload_ms": stats["upload_ms"], "prepare_ms": stats["prepare_ms"],
                 "device_bytes": stats["device_bytes"], "payload_bytes": stats["payload_bytes"],
                 "max_code_len": stats["max_code_len"], "lut_bits": stats["lut_bits"],
                 "lut_replicas": stats["lut_replicas"], "codebook_size": stats["codebook_size"]},
    }
    if not args.no_extra:
        line["extra"] = extras(args, reader, device, peak)
        bytes_per_row = 4 * dim + payload_per_row + 4
        sharded["roofline_frac_per_gpu"] = (sharded["rows_per_s"] / world) * bytes_per_row / 1e9 / peak
        line["extra"]["full_vocab_sharded"] = sharded
        line["extra"].update(extra_configs(args, reader, device, peak))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    print(json.dumps(line), file=_JSON_OUT, flush=True)
    if not parity:
        raise SystemExit("bench: GPU rows differ from the oracle")


def sharded_full_vocab(reader, device, world, rank, barrier, max_over_ranks):
    """configs[2]: every rank decodes its row range of the whole vocabulary (strong scaling:
    total work fixed); then one NCCL all-gather so that every device holds the dense matrix.
    Decode and gather are timed separately (max over ranks, CUDA events)."""
    import torch

    from memb_b200.sharding import all_gather_rows, decode_range_sharded, shard_range

    rows, dim = reader.rows, reader.dim
    lo, hi = shard_range(rows, rank, world)
    out = torch.empty((hi - lo, dim), dtype=torch.float32, device=device)
    for _ in range(3):
        decode_range_sharded(reader, out=out)
    barrier()
    reps = 20
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(reps):
        decode_range_sharded(reader, out=out)
    stop.record()
    barrier()
    decode_ms = max_over_ranks(start.elapsed_time(stop)) / reps
    result = {"rows": rows, "gpus": world, "decode_ms": decode_ms, "rows_per_s": rows / (decode_ms / 1e3),
              "scaling": "strong", "gather_ms": None}
    if world > 1:
        gathered = all_gather_rows(out, rows)
        barrier()
        start.record()
        for _ in range(5):
            gathered = all_gather_rows(out, rows)
        stop.record()
        barrier()
        result["gather_ms"] = max_over_ranks(start.elapsed_time(stop)) / 5
        result["gather_gb_per_s_per_gpu"] = rows * dim * 4 * (world - 1) / world / (result["gather_ms"] / 1e3) / 1e9
        del gathered
    del out
    return result


def extras(args, reader, device, peak):
    """Secondary measurements (rank 0, one GPU): full-vocabulary decode (configs[2] at
    G=1) and the words entry point end to end."""
    import torch

    impl = reader._impl
    rows, dim = reader.rows, reader.dim
    out = torch.empty((rows, dim), dtype=torch.float32, device=device)
    stream = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        impl.decode_range_into(0, rows, out.data_ptr(), True, stream)
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    torch.cuda.synchronize()
    start.record()
    for _ in range(reps):
        impl.decode_range_into(0, rows, out.data_ptr(), True, stream)
    stop.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(stop) / reps
    stats = reader.load_stats()
    bytes_per_row = 4 * dim + stats["payload_bytes"] / rows + 4
    result = {"full_vocab_decode": {"rows": rows, "ms": ms, "rows_per_s": rows / (ms / 1e3),
                                    "roofline_frac": rows * bytes_per_row / (ms / 1e3) / 1e9 / peak}}
    del out
    # words entry point: 1M words -> rows, host buffers, through the binding's list[str] API
    n = min(args.batch, 1_000_000)
    keys = reader.keys()
    rng = np.random.Generator(np.random.PCG64(44))
    words = [keys[i] for i in rng.integers(0, rows, n)]
    reader.batch_embedding(words[:1000])
    t0 = time.perf_counter()
    got = reader.batch_embedding(words)
    first = time.perf_counter() - t0
    del got
    best = float("inf")
    for _ in range(3):
        t0 = time.perf_counter()
        got = reader.batch_embedding(words)
        best = min(best, time.perf_counter() - t0)
        del got
    result["words_api_e2e"] = {"words": n, "s": best, "rows_per_s": n / best, "first_call_s": first,
                               "note": "Reader.batch_embedding(list[str]) -> NumPy array: str marshalling, H2D, resolve "
                                       "kernel, decode, D2H; the first call also pins its result buffer, later calls "
                                       "reuse it"}
    # the same words with the result left on the GPU (as_tensor=True), and pre-packed
    # (UTF-8 blob + offsets: no Python objects per word); resolve alone = K0 + its copies
    from memb_b200 import _memb
    packed = _memb.pack_words(words)

    def wall(fn, reps=5):
        fn()
        torch.cuda.synchronize()
        best = float("inf")
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best

    for name, fn in (("words_list_to_device_tensor", lambda: reader.batch_embedding(words, as_tensor=True)),
                     ("words_packed_to_device_tensor", lambda: reader.batch_embedding(packed, as_tensor=True)),
                     ("words_packed_to_numpy", lambda: reader.batch_embedding(packed)),
                     ("resolve_packed_to_host_ids", lambda: reader.resolve(packed))):
        dt = wall(fn)
        result[name] = {"words": n, "s": dt, "rows_per_s": n / dt}
    result["words_query_bytes"] = len(packed[0])
    return result


def _time_launches(fn, reps=20, warm=3):
    import torch

    for _ in range(warm):
        fn()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    start.record()
    for _ in range(reps):
        fn()
    stop.record()
    torch.cuda.synchronize()
    return start.elapsed_time(stop) / reps


def zipf_ids(rows, n, seed):
    """Zipf(s = 1) over a random permutation of the rows: rank k drawn with p ~ 1/k
    (inverse CDF of the continuous approximation), BASELINE.json configs[4]."""
    rng = np.random.Generator(np.random.PCG64(seed))
    ranks = np.minimum(np.exp(rng.random(n) * np.log(rows + 1.0)).astype(np.int64), rows) - 1
    perm = rng.permutation(rows)
    return perm[ranks].astype(np.int32)


def extra_configs(args, reader, device, peak):
    """BASELINE.json configs[3] and configs[4] on one GPU (rank 0): the uniform storage at 8 bits
    (payload is one byte per weight whatever the bit width) and the fused ReadersUnion
    of two trained 4-bit readers, average and concatenate, Zipf-distributed lookups."""
    import torch

    import memb_b200
    from memb_b200 import _memb

    out = {}
    stream = torch.cuda.current_stream().cuda_stream
    n, dim = args.batch, args.dim
    ids = torch.from_numpy(batch_ids(args, 0)).to(device)

    # ---- configs[3]: uniform, 2 M x dim; the payload is one byte per weight at every bit width, so the
    # sweep only changes the divisor: levels = 16 (a power of two: exact multiply) and 255 (IEEE divide) ----
    rows_u = min(args.rows, 2_000_000)
    ids_u = torch.from_numpy(np.random.Generator(np.random.PCG64(41)).integers(0, rows_u, n).astype(np.int32)).to(device)
    buf = torch.empty((max(n, rows_u), dim), dtype=torch.float32, device=device)
    per_row = 4 * dim + dim + 8
    for bits in (4, 8):
        build_table(args, "uniform", bits, seed=5, rows=rows_u)
        uni = memb_b200.Reader(table_path(args, "uniform", bits, 5, rows_u), device=device.index)
        ms_ids = _time_launches(lambda: uni._impl.decode_ids_into(ids_u.data_ptr(), True, n, buf.data_ptr(), True, stream))
        ms_all = _time_launches(lambda: uni._impl.decode_range_into(0, rows_u, buf.data_ptr(), True, stream))
        out["uniform_%dbit" % bits] = {
            "rows": rows_u, "kernel": "uniform_decode_kernel" if bits in (1, 2, 4) else "uniform_rows_kernel",
            "random_ids": {"n": n, "ms": ms_ids, "rows_per_s": n / (ms_ids / 1e3),
                           "roofline_frac": n * (per_row + 4) / (ms_ids / 1e3) / 1e9 / peak},
            "full_vocab": {"ms": ms_all, "rows_per_s": rows_u / (ms_all / 1e3),
                           "roofline_frac": rows_u * per_row / (ms_all / 1e3) / 1e9 / peak},
            "algorithmic_bytes_per_row": per_row}
        del uni

    # ---- the `full` storage (K3, src/full_compression.cpp:41-42): 1 M x dim fp32 rows, a pure gather ----
    rows_f = min(args.rows, 1_000_000)
    build_table(args, "full", 32, seed=9, rows=rows_f)
    full = memb_b200.Reader(table_path(args, "full", 32, 9, rows_f), device=device.index)
    ids_f = torch.from_numpy(np.random.Generator(np.random.PCG64(42)).integers(0, rows_f, n).astype(np.int32)).to(device)
    ms_ids = _time_launches(lambda: full._impl.decode_ids_into(ids_f.data_ptr(), True, n, buf.data_ptr(), True, stream))
    ms_all = _time_launches(lambda: full._impl.decode_range_into(0, rows_f, buf.data_ptr(), True, stream))
    out["full_fp32"] = {
        "rows": rows_f, "kernel": "full_gather_kernel",
        "random_ids": {"n": n, "ms": ms_ids, "rows_per_s": n / (ms_ids / 1e3),
                       "roofline_frac": n * (8 * dim + 4) / (ms_ids / 1e3) / 1e9 / peak},
        "full_vocab": {"ms": ms_all, "rows_per_s": rows_f / (ms_all / 1e3),
                       "roofline_frac": rows_f * 8 * dim / (ms_all / 1e3) / 1e9 / peak},
        "algorithmic_bytes_per_row": 8 * dim}
    del full, ids_f
    del buf

    # ---- configs[4]: ReadersUnion of two trained readers, Zipf lookups ----
    build_table(args, "trained", args.bits, seed=7)
    second = memb_b200.Reader(table_path(args, "trained", args.bits, 7), device=device.index)
    readers = [reader._impl, second._impl]
    zipf = zipf_ids(args.rows, n, 8)
    # row ids per reader, reader-major; the two vocabularies share their words here, so the same
    # id addresses the same word in both
    ids2 = torch.from_numpy(np.concatenate([zipf, zipf])).to(device)
    payload = reader.load_stats()["payload_bytes"] / reader.rows + second.load_stats()["payload_bytes"] / second.rows
    union = {}
    for mode, cols in (("average", dim), ("concatenate", 2 * dim)):
        buf = torch.empty((n, cols), dtype=torch.float32, device=device)
        ms = _time_launches(lambda: _memb.union_decode_ids_into(readers, mode, ids2.data_ptr(), True, n,
                                                                buf.data_ptr(), True, stream))
        bytes_row = 4 * cols + payload + 2 * (4 + 4)
        union[mode] = {"n": n, "ms": ms, "rows_per_s": n / (ms / 1e3), "out_columns": cols,
                       "algorithmic_bytes_per_row": bytes_row,
                       "roofline_frac": n * bytes_row / (ms / 1e3) / 1e9 / peak}
        del buf
    union["ids"] = "Zipf(s=1) over a permutation of the rows, PCG64 seed 8"
    union["launches_per_call"] = {"average": 2, "concatenate": 3}  # concatenate: + the shared-sector stitch pass
    out["readers_union_2x_trained"] = union
    del second, ids, ids2
    return out


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version
    # banner, for one) is sent to stderr instead
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
