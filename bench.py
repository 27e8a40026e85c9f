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
pinned fp32 rows out), copies inside the timed region, beside the box's own D2H
ceiling.  `roofline` is for the dominant kernel; `cpu_baseline` times the CPU oracle
(whose Huffman decode is the reference's own header when oracle/_ref was built) on the
host cores.  First-class fields besides the contract's: `strong_scaling` (configs[2]:
full-vocabulary decode sharded over the N GPUs + in-place all-gather, checked bit for
bit on every rank), `readers_union` (configs[4]: two readers with 90 % overlapping
vocabularies, Zipf words sharded over the N GPUs, one fused launch), `parity` (every
rank checks its own rows against the oracle; any mismatch -> exit code != 0).  At one
GPU `extra` adds the words entry points, K0 alone, configs[0] and the configs[3] sweep.
The oracle is only ever the thing compared against / the CPU baseline here, and the
reference arm never imports the product (`--prepare` builds its table in a child process).
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
    p.add_argument("--no-sweep", action="store_true", help="skip the configs[3] sweep (trained 1..8 bits, 16-bit codes, uniform)")
    p.add_argument("--prepare", action="store_true", help="build the headline table into the cache and exit (the reference "
                                                          "arm runs this in a child process so that it never maps the product)")
    p.add_argument("--union-words", type=int, default=4_000_000, help="configs[4]: Zipf words per step, all GPUs together")
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
def table_path(args, storage="trained", bits=None, seed=3, rows=None, first_word=0):
    bits = args.bits if bits is None else bits
    rows = args.rows if rows is None else rows
    tag = ("" if seed == 3 else "_s%d" % seed) + ("_w%d" % first_word if first_word else "")
    return os.path.join(args.cache_dir, "glove_shape_%d_x_%d_%s%d%s.bin" % (rows, args.dim, storage, bits, tag))


def build_table(args, storage="trained", bits=None, seed=3, rows=None, first_word=0):
    """Synthetic GloVe-840B-shaped file through the repo's Builder (the memb_converter
    equivalent): N(0,1) float32 vectors (PCG64 seed `seed` per chunk), unique words
    "w<hex id>" (ids first_word .. first_word + rows - 1) in shuffled insertion order.
    Cached on disk; returns build seconds (0 when cached)."""
    bits = args.bits if bits is None else bits
    rows = args.rows if rows is None else rows
    path = table_path(args, storage, bits, seed, rows, first_word)
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
    words = ["w%x" % (first_word + i) for i in order]
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


def build_deep_table(args, rows=1_000_000):
    """configs[3], adversarial file: 17 symbols with code lengths 1, 2, .., 15, 16, 16 (the longest a
    memb file can hold, src/prefix_code.h:10-13), symbols drawn with the code's own probabilities
    2^-len.  No Builder emits this (k-means prunes rare clusters), so the rows are packed with the
    test suite's independent Python writer (tests/membfile_py.py)."""
    path = os.path.join(args.cache_dir, "deep16_%d_x_%d.bin" % (rows, args.dim))
    if os.path.exists(path):
        return path
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import membfile_py

    os.makedirs(args.cache_dir, exist_ok=True)
    keys, lengths = list(range(17)), list(range(1, 17)) + [16]
    rng = np.random.Generator(np.random.PCG64(16))
    weights = np.array([2.0 ** -n for n in lengths])
    symbols = rng.choice(17, (rows, args.dim), p=weights / weights.sum()).astype(np.uint8)
    packed, offsets = membfile_py.pack_rows_numpy(symbols, keys, lengths)
    words = [b"d%07x" % i for i in range(rows)]  # fixed width: already sorted
    tmp = path + ".tmp%d" % os.getpid()
    membfile_py.write_trained_packed(tmp, words, packed, offsets, keys, lengths,
                                     rng.standard_normal(17).astype(np.float32), args.dim)
    os.replace(tmp, path)
    return path


def build_c1_table(args):
    """configs[0]: 10 k words x 300 as word2vec TEXT -> the memb_converter equivalent
    (memb_b200.converter, same flags as tools/converter/memb_converter) with
    --quantization trained --bits-per-weight 6."""
    path = os.path.join(args.cache_dir, "c1_10000_x_%d_trained6.bin" % args.dim)
    if os.path.exists(path):
        return path
    from memb_b200.converter import convert

    os.makedirs(args.cache_dir, exist_ok=True)
    rng = np.random.Generator(np.random.PCG64(1))
    matrix = rng.standard_normal((10_000, args.dim), dtype=np.float32)
    text = path + ".txt"
    with open(text, "w") as f:
        f.write("10000 %d\n" % args.dim)
        for i, row in enumerate(matrix):
            f.write("c%x " % i + " ".join("%.6g" % v for v in row) + "\n")
    convert(text, path + ".tmp", "trained", 6, out=sys.stderr)
    os.replace(path + ".tmp", path)
    os.remove(text)
    return path


# ------------------------------------------------------------- oracle (checker / CPU baseline only) ----
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
        lib.oracle_batch_embedding.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_uint64,
                                               ctypes.c_void_p]
        lib.oracle_decode_range.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_void_p]
        lib.oracle_close.argtypes = [ctypes.c_void_p]
        self.lib = lib
        self.h = lib.oracle_open(path.encode(), threads)
        if not self.h:
            raise RuntimeError("oracle cannot open " + path)
        self.threads = int(lib.oracle_threads(self.h))

    def close(self):
        if self.h:
            self.lib.oracle_close(self.h)
            self.h = None

    def batch_ids(self, ids, out):
        self.lib.oracle_batch_ids(self.h, ids.ctypes.data, len(ids), out.ctypes.data)

    def batch_words(self, blob, offsets, out):
        """the reference's own call: lower_bound + strcmp per word, then decode (src/reader.cpp:59-86)"""
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self.lib.oracle_batch_embedding(self.h, bytes(blob), offsets.ctypes.data, len(offsets) - 1, out.ctypes.data)

    def rows_of(self, ids, dim):
        out = np.empty((len(ids), dim), np.float32)
        self.batch_ids(np.ascontiguousarray(ids, np.int32), out)
        return out

    def describe(self):
        codec = ("reference huffman_table_decoder.h/bit_stream_reader.h compiled from /root/reference (oracle/_ref)"
                 if self.uses_reference_codec else "restated codec (oracle/oracle.cpp)")
        return "CPU oracle Reader: %s + restated lookup/threading of src/reader.cpp:59-86" % codec


def time_cpu(call, steps, warmup, budget_s=25.0):
    """Seconds of every pass of `call` (<= steps passes within the budget)."""
    for _ in range(warmup):
        call()
    times, start = [], time.perf_counter()
    for _ in range(steps):
        t0 = time.perf_counter()
        call()
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - start > budget_s:
            break
    return times


def same_bits(a, b):
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and bool(np.array_equal(a.view(np.uint32), b.view(np.uint32)))


def pack_word_list(words):
    """list[str] -> (utf-8 blob, uint64 offsets[n+1]) without the product (reference arm)."""
    encoded = [w.encode("utf-8") for w in words]
    offsets = np.zeros(len(encoded) + 1, np.uint64)
    offsets[1:] = np.cumsum([len(e) for e in encoded])
    return b"".join(encoded), offsets


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
              "needed" % (args.batch * args.dim * 4 / 1e6, "~300 MB" if args.rows >= 2_000_000 else "small"),
        "parallelism": "replicated table, queries sharded by rank, no collective",
    }
    config.update(extra or {})
    return config


# ------------------------------------------------------------- reference arm ----
def run_reference(args):
    """The reference's CPU implementation of the path on the host cores: the oracle Reader
    (reference Huffman headers when oracle/_ref was built).  This process never imports the
    product: a missing table is built by a child process (`bench.py --prepare`)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 measures the CPU implementation
    if not os.path.exists(table_path(args)):
        import subprocess

        child_env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
        subprocess.run([sys.executable, os.path.abspath(__file__), "--prepare", "--rows", str(args.rows), "--dim",
                        str(args.dim), "--bits", str(args.bits), "--cache-dir", args.cache_dir],
                       check=True, stdout=sys.stderr, env=child_env)
    oracle = CpuOracle(table_path(args), threads=0)
    ids = batch_ids(args, 0)
    dim = args.dim
    # bounded sample: a step is the full batch unless the whole run would take minutes
    sample = len(ids)
    t0 = time.perf_counter()
    oracle.rows_of(ids[:20000], dim)
    per_row = (time.perf_counter() - t0) / 20000
    budget_rows = int(60.0 / max(per_row, 1e-12) / max(args.steps + args.warmup, 1))
    if budget_rows < sample:
        sample = max(1024, budget_rows)
    out = np.empty((sample, dim), np.float32)
    times = time_cpu(lambda: oracle.batch_ids(ids[:sample], out), args.steps, args.warmup, budget_s=120.0)
    ms = 1e3 * float(np.mean(times))
    value = sample / (ms / 1e3)
    # the same rows asked for by WORD, the reference's own entry point: ~21 lower_bound probes per word
    # (src/trained_compression.cpp:115-127) on top of the decode
    n_words = min(sample, 200_000)
    blob, offsets = pack_word_list(["w%x" % i for i in ids[:n_words]])
    out_w = np.empty((n_words, dim), np.float32)
    times_w = time_cpu(lambda: oracle.batch_words(blob, offsets, out_w), 5, 1, budget_s=20.0)
    words_value = n_words / float(np.mean(times_w))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8 symbols -> f32 codebook gather", "data": "synthetic",
        "config": workload_config(args),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": oracle.threads, "kind": "port",
                         "codec": "reference headers" if oracle.uses_reference_codec else "restated",
                         "sample": "%d of the %d ids of one step, %d timed passes; %s; host has %d logical cpus"
                                   % (sample, len(ids), len(times), oracle.describe(), os.cpu_count() or 0),
                         "note": "decodes ROW IDS like the b200 arm's `value`: the reference's per-word lower_bound + "
                                 "strcmp (src/trained_compression.cpp:115-127) is skipped, which favours the CPU; "
                                 "`by_word_value` is the same lookup through the word entry point",
                         "by_word_value": words_value, "by_word_sample": n_words},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "output_gb_per_s": value * args.dim * 4 / 1e9,
    }
    if "memb_b200" in sys.modules or "torch" in sys.modules:
        raise SystemExit("reference arm imported the product")
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


def ncu_traffic(args, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the headline kernel from the
    committed `ncu --set full` capture (profiles/k1_ncu_latest.json), valid for the default workload only."""
    try:
        with open(os.path.join(ROOT, "profiles", "k1_ncu_latest.json")) as f:
            rec = json.load(f)
        if ((rec["rows"], rec["dim"], rec["bits"], rec["batch"]) == (args.rows, args.dim, args.bits, args.batch)
                and rec.get("kernel", kernel) == kernel):
            return rec
    except Exception:  # noqa: BLE001
        pass
    return None


class Timer:
    """CUDA-event timing on torch's current stream (the stream every launch below is queued on)."""

    def __init__(self, torch):
        self.torch = torch
        self.start, self.stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def ms(self, fn, reps=20, warm=3):
        for _ in range(warm):
            fn()
        self.torch.cuda.synchronize()
        self.start.record()
        for _ in range(reps):
            fn()
        self.stop.record()
        self.torch.cuda.synchronize()
        return self.start.elapsed_time(self.stop) / reps


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

    def all_ranks_true(flag):
        if world == 1:
            return bool(flag)
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    ctx = {"torch": torch, "dist": dist, "memb_b200": memb_b200, "_memb": _memb, "rank": rank, "world": world,
           "local": local, "device": device, "barrier": barrier, "max_over_ranks": max_over_ranks,
           "all_ranks_true": all_ranks_true, "timer": Timer(torch), "peak": measured_peak()[0]}

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

    # ---- parity: EVERY rank checks a sample of its own batch against the oracle ----
    oracle = CpuOracle(table_path(args), threads=0)
    sample_n = min(n, 20_000 if args.quick else 50_000)
    want = oracle.rows_of(ids_host.numpy()[:sample_n], dim)
    parity_device = all_ranks_true(same_bits(want, out_dev[:sample_n].cpu().numpy()))

    payload_per_row = stats["payload_bytes"] / rows
    bytes_per_row = 4 * dim + payload_per_row + 4 + 4  # fp32 out + compressed row + value_offset + id
    peak, peak_source = measured_peak()
    if args.quick:
        if rank == 0:
            frac = n * bytes_per_row / (ms_step / 1e3) / 1e9 / peak
            whole = torch.empty((rows, dim), dtype=torch.float32, device=device)
            full_ms = ctx["timer"].ms(lambda: impl.decode_range_into(0, rows, whole.data_ptr(), True, stream), reps=10)
            odd_ms = ctx["timer"].ms(lambda: impl.decode_range_into(7, rows, whole.data_ptr(), True, stream), reps=10)
            want_tail = np.empty((5000, dim), np.float32)
            oracle.lib.oracle_decode_range(oracle.h, rows - 5000, rows, want_tail.ctypes.data)
            parity_device = parity_device and same_bits(want_tail, whole[rows - 5007:rows - 7].cpu().numpy())
            print(json.dumps({"quick": True, "full_vocab_ms": full_ms, "full_vocab_from_row_7_ms": odd_ms,
                              "full_vocab_frac": rows * (bytes_per_row - 4) / (full_ms / 1e3) / 1e9 / peak, "ids": args.ids, "options": args.option, "ms_per_step": ms_step,
                              "rows_per_s": value, "roofline_frac": frac, "parity_sample": parity_device, "clocks": clocks,
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
    parity_host = all_ranks_true(same_bits(want, out_host.numpy()[:sample_n]))
    # the box's own ceiling for this leg: the same 1.2 GB device -> pinned host copy, nothing else, on all
    # N GPUs at once (every rank copies into its own pinned buffer)
    barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        out_host.copy_(out_dev, non_blocking=True)
        torch.cuda.synchronize()
    d2h_ms = max_over_ranks((time.perf_counter() - t0) * 1e3) / 5
    d2h_ceiling = {"ms_per_copy": d2h_ms, "gb_per_s_per_gpu": n * dim * 4 / (d2h_ms / 1e3) / 1e9,
                   "e2e_frac_of_ceiling": d2h_ms / e2e_ms,
                   "how": "cudaMemcpyAsync device -> pinned host of one step's result, %d GPUs at once, max over ranks" % world}
    del out_host

    # ---- all ranks: configs[2] strong scaling and configs[4] sharded union ----
    strong = strong_scaling(args, ctx, reader, oracle)
    union = None if args.no_extra else sharded_union(args, ctx)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0: roofline, cpu baseline, extras ----
    kernel = "trained_decode_kernel" if "trained_rows_kernel=2" not in args.option else "trained_rows_kernel"
    achieved = n * bytes_per_row / (ms_step / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": None, "peak_source": peak_source,
                "algorithmic_bytes_per_row": bytes_per_row, "algorithmic_bytes_per_launch": n * bytes_per_row,
                "write_only_frac": n * 4 * dim / (ms_step / 1e3) / 1e9 / peak,
                "launch_ms": ms_step}
    traffic = ncu_traffic(args, kernel)
    if traffic:
        roofline["traffic"] = traffic["dram_bytes_read"] + traffic["dram_bytes_write"]
        roofline["traffic_over_algorithmic"] = roofline["traffic"] / (n * bytes_per_row)
        roofline["traffic_source"] = traffic["source"]

    cpu_n = min(n, 200_000) if (os.cpu_count() or 1) < 32 else n
    cpu_out = np.empty((cpu_n, dim), np.float32)
    cpu_ids = ids_host.numpy()[:cpu_n]
    cpu_times = time_cpu(lambda: oracle.batch_ids(cpu_ids, cpu_out), steps=5, warmup=1, budget_s=20.0)
    cpu_value = cpu_n / float(np.min(cpu_times))
    cpu_baseline = {"value": cpu_value, "unit": UNIT, "cores": oracle.threads, "kind": "port",
                    "codec": "reference headers" if oracle.uses_reference_codec else "restated",
                    "sample": "first %d ids of the step's batch, best of %d passes; %s; host has %d logical cpus"
                              % (cpu_n, len(cpu_times), oracle.describe(), os.cpu_count() or 0)}
    parity = parity_device and parity_host and strong["parity_sharded"] and (union is None or union["parity"])

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
                                % (len(binding.cpus) if binding.cpus is not None else "?", len(os.sched_getaffinity(0))),
                "d2h_ceiling": d2h_ceiling},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
        "parity_vs_oracle": parity,
        "parity": {"device_rows_every_rank": parity_device, "host_rows_every_rank": parity_host,
                   "sharded_full_vocab_every_rank": strong["parity_sharded"],
                   "union_every_rank": None if union is None else union["parity"], "rows_checked_per_rank": sample_n},
        "strong_scaling": strong,
        "load": {"table_build_s": build_s, "open_s": open_s, "parse_ms": stats["parse_ms"],
                 "upload_ms": stats["upload_ms"], "prepare_ms": stats["prepare_ms"],
                 "device_bytes": stats["device_bytes"], "payload_bytes": stats["payload_bytes"],
                 "max_code_len": stats["max_code_len"], "lut_bits": stats["lut_bits"],
                 "lut_replicas": stats["lut_replicas"], "codebook_size": stats["codebook_size"]},
    }
    if union is not None:
        line["readers_union"] = union
    if not args.no_extra and world == 1:
        extra = extras(args, ctx, reader)
        if not args.no_sweep:
            extra["configs3_sweep"] = sweep(args, ctx)
        line["extra"] = extra
        extras_ok = (extra["configs0_small_batch"]["parity"] and extra["gpu_converter"]["files_byte_identical"] and
                     extra["resolve_kernel_device_only"]["ids_match_host_resolve"] and
                     all(t["parity"] for t in extra.get("configs3_sweep", {}).values()))
        line["parity"]["extras"] = bool(extras_ok)
        parity = parity and bool(extras_ok)
        line["parity_vs_oracle"] = parity
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    print(json.dumps(line), file=_JSON_OUT, flush=True)
    if not parity:
        raise SystemExit("bench: GPU rows differ from the oracle: %s" % json.dumps(line["parity"]))


def strong_scaling(args, ctx, reader, oracle):
    """configs[2]: full-vocabulary decode (the tokenizer_embedding / to_keyed_vectors path) sharded by
    row range -- total work fixed, so this is STRONG scaling -- then every rank must hold the dense
    matrix: each rank decodes straight into its slot of the gathered buffer and one in-place NCCL
    all-gather follows.  Decode, gather and decode + gather are timed separately (CUDA events, max
    over ranks).  Every rank compares the gathered matrix bit for bit with a whole decode of its own,
    and a sample of it with the oracle."""
    torch, dist, timer = ctx["torch"], ctx["dist"], ctx["timer"]
    world, rank, device, peak = ctx["world"], ctx["rank"], ctx["device"], ctx["peak"]
    from memb_b200.sharding import decode_range_gathered, shard_range

    rows, dim = reader.rows, reader.dim
    impl = reader._impl
    stream = torch.cuda.current_stream().cuda_stream
    job = (rows + world - 1) // world
    gathered = torch.empty((world * job, dim), dtype=torch.float32, device=device)
    lo, hi = shard_range(rows, rank, world)
    slot = gathered[rank * job:(rank + 1) * job]

    def decode_shard():
        impl.decode_range_into(lo, hi, slot.data_ptr(), True, stream)

    def gather():
        dist.all_gather_into_tensor(gathered, slot)

    def both():
        decode_range_gathered(reader, out=gathered)

    ctx["barrier"]()
    decode_ms = ctx["max_over_ranks"](timer.ms(decode_shard, reps=20))
    stats = reader.load_stats()
    bytes_per_row = 4 * dim + stats["payload_bytes"] / rows + 4
    result = {"config": "configs[2]: full-vocabulary decode %d x %d -> dense fp32, sharded by row range" % (rows, dim),
              "rows": rows, "gpus": world, "scaling": "strong", "decode_ms": decode_ms,
              "rows_per_s": rows / (decode_ms / 1e3),
              "roofline_frac_per_gpu": (rows / world) * bytes_per_row / (decode_ms / 1e3) / 1e9 / peak,
              "gather_ms": None, "decode_plus_gather_ms": decode_ms}
    whole = torch.empty((rows, dim), dtype=torch.float32, device=device)
    if world > 1:
        ctx["barrier"]()
        result["gather_ms"] = ctx["max_over_ranks"](timer.ms(gather, reps=5, warm=2))
        result["gather_gb_per_s_per_gpu"] = rows * dim * 4 * (world - 1) / world / (result["gather_ms"] / 1e3) / 1e9
        ctx["barrier"]()
        result["decode_plus_gather_ms"] = ctx["max_over_ranks"](timer.ms(both, reps=5, warm=2))
        result["gather"] = "in place: ncclAllGather of the slot each rank decoded into (no send copy)"
        # fused: the decode kernel's stores go to an NVLS multicast address (multimem.st), the NVSwitch
        # replicates them into every rank's buffer while the decode runs -- no collective at all
        try:
            from memb_b200.sharding import GatheredMatrix

            fused = GatheredMatrix(reader)
            ctx["barrier"]()
            result["fused_decode_gather_ms"] = ctx["max_over_ranks"](timer.ms(fused.fill, reps=5, warm=2))
            result["fused_gb_per_s_into_each_gpu"] = rows * dim * 4 * (world - 1) / world / (result["fused_decode_gather_ms"] / 1e3) / 1e9
            result["fused"] = "decode kernel storing to torch symmetric memory's NVLS multicast address (the switch replicates every store) + a one-element all-reduce as the barrier"
            fused_view = fused.fill()
            torch.cuda.synchronize()
            result["fused_available"] = True
        except Exception as exc:  # noqa: BLE001  (no NVLS / symmetric memory on this box)
            fused_view = None
            result["fused_available"] = False
            result["fused_error"] = str(exc)[:200]
    # one GPU decoding everything alone, timed in the same run on rank 0's GPU while the others wait
    ctx["barrier"]()
    alone_ms = timer.ms(lambda: impl.decode_range_into(0, rows, whole.data_ptr(), True, stream), reps=10) if rank == 0 else 0.0
    alone_ms = ctx["max_over_ranks"](alone_ms)
    result["decode_ms_one_gpu"] = alone_ms
    result["efficiency_vs_n1"] = alone_ms / (world * decode_ms)
    # parity on every rank: gathered == whole decode (bitwise), and a sample of the whole decode == oracle
    both()
    impl.decode_range_into(0, rows, whole.data_ptr(), True, stream)
    torch.cuda.synchronize()
    same = bool(torch.equal(gathered[:rows].view(torch.int32), whole.view(torch.int32)))
    if world > 1 and fused_view is not None:
        result["parity_fused"] = bool(torch.equal(fused_view.view(torch.int32), whole.view(torch.int32)))
        same = same and result["parity_fused"]
    begin = (rows // max(world, 1)) * rank
    count = min(20_000, rows - begin)
    want = np.empty((count, dim), np.float32)
    oracle.lib.oracle_decode_range(oracle.h, begin, begin + count, want.ctypes.data)
    same = same and same_bits(want, whole[begin:begin + count].cpu().numpy())
    result["parity_sharded"] = ctx["all_ranks_true"](same)
    del gathered, whole
    return result


def zipf_ranks(count, n, seed):
    """Zipf(s = 1) ranks 0..count-1: rank k drawn with p ~ 1/(k+1) (inverse CDF of the continuous
    approximation), BASELINE.json configs[4]."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return np.minimum(np.exp(rng.random(n) * np.log(count + 1.0)).astype(np.int64), count) - 1


def sharded_union(args, ctx):
    """configs[4]: ReadersUnion of two 2 M x dim trained readers whose vocabularies overlap 90 %
    (reader A: words 0 .. 2 M - 1, reader B: words 0.2 M .. 2.2 M - 1, so each misses 9 % of the union
    vocabulary -- the zero row that still counts in np.mean's divisor), Zipf(1.0) WORDS over the union
    vocabulary, average and concatenate, queries sharded over the ranks (no collective).  Timed on the
    device: K0 for both readers + the ONE fused decode launch (CUDA events, queries resident in HBM),
    and end to end through the words entry point (host blob in, device tensor out)."""
    torch, timer, _memb, memb_b200 = ctx["torch"], ctx["timer"], ctx["_memb"], ctx["memb_b200"]
    world, rank, device, peak = ctx["world"], ctx["rank"], ctx["device"], ctx["peak"]
    from memb_b200.sharding import shard_range

    rows_each = min(args.rows, 2_000_000)
    shift = rows_each // 10
    vocab = rows_each + shift
    specs = [dict(storage="trained", bits=args.bits, seed=6, rows=rows_each, first_word=0),
             dict(storage="trained", bits=args.bits, seed=7, rows=rows_each, first_word=shift)]
    build_s = 0.0
    if rank == 0:
        build_s = sum(build_table(args, **spec) for spec in specs)
    ctx["barrier"]()
    readers = [memb_b200.Reader(table_path(args, **spec), device=ctx["local"]) for spec in specs]
    impls = [r._impl for r in readers]
    dim = readers[0].dim
    total = args.union_words
    ranks = zipf_ranks(vocab, total, 8)
    perm = np.random.Generator(np.random.PCG64(88)).permutation(vocab)
    lo, hi = shard_range(total, rank, world)
    word_ids = perm[ranks[lo:hi]]
    n = hi - lo
    distinct_fraction = float(len(np.unique(word_ids))) / max(n, 1)
    width = len("w%x" % (vocab - 1))
    table = np.array(["w%x" % i for i in range(vocab)], dtype="S%d" % width)
    blob, offsets = _memb.pack_words(table[word_ids])
    del table
    d_blob = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
    d_offsets = torch.from_numpy(offsets.astype(np.int64)).to(device)
    d_ids = torch.empty((2, n), dtype=torch.int32, device=device)
    stream = torch.cuda.current_stream().cuda_stream
    payload = sum(r.load_stats()["payload_bytes"] / r.rows for r in readers)
    query_bytes = len(blob) / max(n, 1)
    result = {"config": "configs[4]: union of two %d x %d trained %d-bit readers, 90 %% overlapping vocabularies, %d "
                        "Zipf(1.0) words per step over the %d-word union vocabulary, sharded over %d GPU(s)"
                        % (rows_each, dim, args.bits, total, vocab, world),
              "words": total, "gpus": world, "table_build_s": build_s, "parity": True}
    # oracle for this rank's first words
    check = min(n, max(1, 100_000 // world))
    oracles = [CpuOracle(table_path(args, **spec), threads=0) for spec in specs]
    parts = []
    for o in oracles:
        part = np.empty((check, dim), np.float32)
        o.batch_words(blob, offsets[:check + 1], part)
        parts.append(part)
        o.close()
    missing = float(np.mean([np.all(p == 0, axis=1).mean() for p in parts]))
    for mode, cols in (("average", dim), ("concatenate", 2 * dim)):
        out = torch.empty((n, cols), dtype=torch.float32, device=device)
        want = np.mean(parts, axis=0) if mode == "average" else np.concatenate(parts, axis=-1)

        def device_step():
            for k, impl in enumerate(impls):
                impl.resolve_into(d_blob.data_ptr(), d_offsets.data_ptr(), n, d_ids[k].data_ptr(), stream)
            _memb.union_decode_ids_into(impls, mode, d_ids.data_ptr(), True, n, out.data_ptr(), True, stream)

        def decode_step():
            _memb.union_decode_ids_into(impls, mode, d_ids.data_ptr(), True, n, out.data_ptr(), True, stream)

        def measure():
            launches0 = _memb.launch_count()
            device_step()
            per_call = _memb.launch_count() - launches0
            ctx["barrier"]()
            both = ctx["max_over_ranks"](timer.ms(device_step, reps=10))
            dec = ctx["max_over_ranks"](timer.ms(decode_step, reps=10))
            torch.cuda.synchronize()
            return both, dec, int(per_call), same_bits(want, out[:check].cpu().numpy())

        # Every query decoded (union_dedup = 0) -- the kernels' own numbers: the fused ONE-launch kernel
        # (trained_rows_kernel, option trained_rows_kernel = 1) beside one ring-kernel launch per reader with fused
        # merge epilogues; both measured, the faster is the default.
        _memb.set_option("union_dedup", 0)
        _memb.set_option("trained_rows_kernel", 1)
        one_ms, one_decode, one_launches, one_ok = measure()
        _memb.set_option("trained_rows_kernel", 0)
        all_ms, all_decode, all_launches, all_ok = measure()
        # The default (union_dedup = 1): a batch that repeats its words (this one is Zipf) has its DISTINCT id tuples
        # decoded once and the rows copied out, when the cost model says the copy-out is cheaper than the decodes
        # it saves (average: yes; concatenate of two readers: no -- same launches as above).
        _memb.set_option("union_dedup", 1)
        ms, decode_only, launches_per_call, ok = measure()
        ok = ok and one_ok and all_ok
        # end to end: host blob -> device tensor through the words entry point (H2D + 2 x K0 + fused decode)
        packed = (blob, offsets)
        _memb.union_batch_embedding_into(impls, mode, packed, out.data_ptr(), n, True, stream)
        ctx["barrier"]()
        t0 = time.perf_counter()
        for _ in range(3):
            _memb.union_batch_embedding_into(impls, mode, packed, out.data_ptr(), n, True, stream)
        torch.cuda.synchronize()
        e2e_ms = ctx["max_over_ranks"]((time.perf_counter() - t0) * 1e3) / 3
        ok = ok and same_bits(want, out[:check].cpu().numpy())
        bytes_row = 4 * cols + payload + 2 * (4 + 4)      # fp32 out + both compressed rows + both offsets + both ids
        bytes_row_k0 = 2 * (query_bytes + 8 + 8 + query_bytes + 4)  # per reader: query + offset + hash slot + key compared + id written
        result[mode] = {"ms_resolve_and_decode": ms, "ms_decode": decode_only, "union_rows_per_s": total / (ms / 1e3),
                        "launches_per_call": {"resolve": 2, "decode": int(launches_per_call) - 2},
                        "algorithmic_bytes_per_row_decode": bytes_row,
                        "roofline_frac_decode_per_gpu": n * bytes_row / (decode_only / 1e3) / 1e9 / peak,
                        "roofline_frac_resolve_and_decode_per_gpu": n * (bytes_row + bytes_row_k0) / (ms / 1e3) / 1e9 / peak,
                        "words_api_host_blob_to_device_tensor_ms": e2e_ms, "out_columns": cols, "parity": ok,
                        "distinct_tuples_path": {"taken": int(launches_per_call) > int(all_launches),
                                                 "distinct_fraction_of_batch": distinct_fraction,
                                                 "note": "option union_dedup (default 1): claim kernel over the id tuples, decode of the "
                                                         "distinct ones, copy-out; two 4-byte host round trips decide"},
                        "every_query_decoded": {"ms_decode": all_decode, "ms_resolve_and_decode": all_ms,
                                                "launches_per_call": {"resolve": 2, "decode": int(all_launches) - 2},
                                                "roofline_frac_decode_per_gpu": n * bytes_row / (all_decode / 1e3) / 1e9 / peak,
                                                "parity": all_ok},
                        "one_launch_kernel": {"kernel": "trained_rows_kernel (option trained_rows_kernel=1)", "ms_decode": one_decode,
                                              "ms_resolve_and_decode": one_ms, "decode_launches": one_launches - 2,
                                              "roofline_frac_decode_per_gpu": n * bytes_row / (one_decode / 1e3) / 1e9 / peak,
                                              "parity": one_ok}}
        result["parity"] = result["parity"] and ok
        del out
    result["parity"] = ctx["all_ranks_true"](result["parity"])
    result["checked_rows_per_rank"] = check
    result["rows_missing_in_one_reader_frac"] = missing
    result["ids"] = "Zipf(s=1) ranks (PCG64 seed 8) over a permutation (seed 88) of the union vocabulary"
    del readers, impls, d_blob, d_offsets, d_ids
    torch.cuda.empty_cache()
    return result


def extras(args, ctx, reader):
    """Secondary measurements at one GPU: the words entry points end to end, K0 alone on the device,
    configs[0] (small table, 1 k words: latency of a call), the uniform and full storages."""
    torch, timer, _memb, memb_b200, device, peak = (ctx["torch"], ctx["timer"], ctx["_memb"], ctx["memb_b200"],
                                                    ctx["device"], ctx["peak"])
    impl = reader._impl
    rows, dim = reader.rows, reader.dim
    stream = torch.cuda.current_stream().cuda_stream
    result = {}
    # words entry point: 1M words -> rows, host buffers, through the binding's list[str] API
    n = min(args.batch, 1_000_000)
    keys = reader.keys()
    rng = np.random.Generator(np.random.PCG64(44))
    words = [keys[i] for i in rng.integers(0, rows, n)]
    del keys
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
    # K0 alone, everything resident in HBM (CUDA events): query bytes + offset + hash slot + key bytes + id
    d_blob = torch.frombuffer(bytearray(packed[0]), dtype=torch.uint8).to(device)
    d_offsets = torch.from_numpy(packed[1].astype(np.int64)).to(device)
    d_ids = torch.empty(n, dtype=torch.int32, device=device)
    ms = timer.ms(lambda: impl.resolve_into(d_blob.data_ptr(), d_offsets.data_ptr(), n, d_ids.data_ptr(), stream))
    k0_bytes = 2 * len(packed[0]) + n * (8 + 8 + 4)
    result["resolve_kernel_device_only"] = {"kernel": "resolve_words_kernel", "words": n, "ms": ms, "words_per_s": n / (ms / 1e3),
                                            "algorithmic_bytes": k0_bytes, "gb_per_s": k0_bytes / (ms / 1e3) / 1e9,
                                            "ids_match_host_resolve": bool(np.array_equal(d_ids.cpu().numpy(), reader.resolve(packed))),
                                            "note": "latency-bound random probes (slot 8 B, then the key text); not a bandwidth kernel"}
    del d_blob, d_offsets, d_ids, words, packed

    # ---- the write path: Builder.save of a trained 4-bit table, host threads vs GPU converter ----
    conv_rows = min(args.rows, 500_000)
    rng = np.random.Generator(np.random.PCG64(90))
    matrix = rng.standard_normal((conv_rows, dim), dtype=np.float32)
    conv_words = ["k%x" % i for i in rng.permutation(conv_rows)]
    conv = {}
    for name, dev in (("host_threads", None), ("gpu", device.index)):
        builder = memb_b200.Builder(dim, "trained", args.bits, device=dev)
        builder.add_words(conv_words, matrix)
        path = os.path.join(args.cache_dir, "converter_probe_%s.bin" % name)
        t0 = time.perf_counter()
        builder.save(path)
        conv[name] = {"save_s": time.perf_counter() - t0, "path": path}
        del builder
    with open(conv["host_threads"]["path"], "rb") as fa, open(conv["gpu"]["path"], "rb") as fb:
        same_file = fa.read() == fb.read()
    for entry in conv.values():
        os.remove(entry.pop("path"))
    result["gpu_converter"] = {"rows": conv_rows, "dim": dim, "bits": args.bits, "host_threads_save_s": conv["host_threads"]["save_s"],
                               "gpu_save_s": conv["gpu"]["save_s"], "files_byte_identical": same_file,
                               "host_cpus": os.cpu_count(),
                               "note": "save() = k-means fit on the 10 000-row sample (host, sequential in both) + predict / "
                                       "Huffman-encode all rows (host threads or GPU, incl. the 0.6 GB H2D from pageable memory) "
                                       "+ word sort + file write"}
    del matrix, conv_words

    # ---- configs[0]: 10 k x dim, trained 6 bit through the converter; 1 k words incl. 5 % misses ----
    c1_path = build_c1_table(args)
    small = memb_b200.Reader(c1_path, device=device.index)
    rng = np.random.Generator(np.random.PCG64(2))
    picks = rng.choice(10_000, 1000, replace=False)
    c1_words = ["c%x" % i for i in picks]
    for i in rng.choice(1000, 50, replace=False):
        c1_words[i] = "missing%d" % i
    blob, offsets = pack_word_list(c1_words)
    cpu = CpuOracle(c1_path, threads=0)  # n = 1000 < 1024: the Reader decodes serially whatever the thread count (src/reader.cpp:9,61)
    want = np.empty((1000, dim), np.float32)
    cpu_times = time_cpu(lambda: cpu.batch_words(blob, offsets, want), 200, 5, budget_s=5.0)
    gpu_times = []
    got = small.batch_embedding(c1_words)
    for _ in range(200):
        t0 = time.perf_counter()
        got = small.batch_embedding(c1_words)
        gpu_times.append(time.perf_counter() - t0)
    packed_small = _memb.pack_words(c1_words)
    dev_times = []
    for _ in range(200):
        t0 = time.perf_counter()
        t = small.batch_embedding(packed_small, as_tensor=True)
        torch.cuda.synchronize()
        dev_times.append(time.perf_counter() - t0)
    result["configs0_small_batch"] = {
        "config": "configs[0]: 10 k x %d word2vec text -> converter --quantization trained --bits-per-weight 6; one "
                  "batch_embedding of 1 k words (5 %% missing)" % dim,
        "cpu_reader_median_ms": 1e3 * float(np.median(cpu_times)), "cpu_threads": 1,
        "cpu_note": "below THREADED_DECODER_THRESHOLD = 1024 the reference Reader is serial (src/reader.cpp:9,61)",
        "gpu_list_of_str_to_numpy_median_ms": 1e3 * float(np.median(gpu_times)),
        "gpu_packed_to_device_tensor_median_ms": 1e3 * float(np.median(dev_times)),
        "gpu_note": "latency of ONE call: marshalling + 2 H2D + K0 + K1 + D2H (1.2 MB) + stream sync; launch-bound",
        "parity": same_bits(want, got) and same_bits(want, t.cpu().numpy())}
    cpu.close()
    del small
    return result


def sweep(args, ctx):
    """configs[3] as SURVEY D2 re-scoped it: the uniform storage has no Huffman code (one byte per weight
    at every width; the sweep only changes the divisor), so the code-length / LUT-depth sweep is run on
    the TRAINED storage, 1 .. 8 bits, plus one adversarial file with 16-bit codes.  Per table: full
    decode and 1 M random ids, device-timed, roofline fraction on the table's own algorithmic bytes,
    and a parity sample against the oracle."""
    torch, timer, memb_b200, device, peak = ctx["torch"], ctx["timer"], ctx["memb_b200"], ctx["device"], ctx["peak"]
    stream = torch.cuda.current_stream().cuda_stream
    n, dim = args.batch, args.dim
    rows_t = min(args.rows, 2_000_000)
    buf = torch.empty((max(n, rows_t), dim), dtype=torch.float32, device=device)
    out = {}

    def measure(path, label, per_row_fixed, kernel):
        rd = memb_b200.Reader(path, device=device.index)
        st = rd.load_stats()
        rows = rd.rows
        ids_np = np.random.Generator(np.random.PCG64(41)).integers(0, rows, n).astype(np.int32)
        ids = torch.from_numpy(ids_np).to(device)
        ms_ids = timer.ms(lambda: rd._impl.decode_ids_into(ids.data_ptr(), True, n, buf.data_ptr(), True, stream), reps=10)
        torch.cuda.synchronize()
        oracle = CpuOracle(path, threads=0)
        ok = same_bits(oracle.rows_of(ids_np[:20_000], dim), buf[:20_000].cpu().numpy())
        ms_all = timer.ms(lambda: rd._impl.decode_range_into(0, rows, buf.data_ptr(), True, stream), reps=10)
        torch.cuda.synchronize()
        want = np.empty((10_000, dim), np.float32)
        oracle.lib.oracle_decode_range(oracle.h, rows - 10_000, rows, want.ctypes.data)
        ok = ok and same_bits(want, buf[rows - 10_000:rows].cpu().numpy())
        oracle.close()
        per_row = per_row_fixed + st["payload_bytes"] / rows
        out[label] = {"rows": rows, "kernel": kernel, "payload_bytes_per_row": st["payload_bytes"] / rows,
                      "max_code_len": st["max_code_len"], "lut_bits": st["lut_bits"], "lut_replicas": st["lut_replicas"],
                      "second_level_table": bool(st["max_code_len"] > st["lut_bits"]), "codebook_size": st["codebook_size"],
                      "random_ids": {"n": n, "ms": ms_ids, "rows_per_s": n / (ms_ids / 1e3),
                                     "roofline_frac": n * (per_row + 4) / (ms_ids / 1e3) / 1e9 / peak},
                      "full_vocab": {"ms": ms_all, "rows_per_s": rows / (ms_all / 1e3),
                                     "roofline_frac": rows * per_row / (ms_all / 1e3) / 1e9 / peak},
                      "algorithmic_bytes_per_row": per_row, "parity": ok}
        del rd, ids

    for bits in range(1, 9):
        build_table(args, "trained", bits, seed=5, rows=rows_t)
        measure(table_path(args, "trained", bits, 5, rows_t), "trained_%dbit" % bits, 4 * dim + 4, "trained_decode_kernel")
    measure(build_deep_table(args), "trained_16bit_codes", 4 * dim + 4, "trained_decode_kernel")
    for bits in (1, 2, 4, 8):
        build_table(args, "uniform", bits, seed=5, rows=rows_t)
        measure(table_path(args, "uniform", bits, 5, rows_t), "uniform_%dbit" % bits, 4 * dim,
                "uniform_decode_kernel" if bits in (1, 2, 4) else "uniform_rows_kernel")
    rows_f = min(args.rows, 1_000_000)
    build_table(args, "full", 32, seed=9, rows=rows_f)
    measure(table_path(args, "full", 32, 9, rows_f), "full_fp32", 4 * dim, "full_gather_kernel")
    del buf
    torch.cuda.empty_cache()
    return out


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything a library prints there (NCCL's version
    # banner, for one) is sent to stderr instead
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.prepare:
        build_table(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
