"""Summarise an .ncu-rep here (no GPU needed): headline metrics, stall mix, hottest SASS lines.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [--top N] [--range A B]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 30
rng = None
if "--range" in sys.argv:
    i = sys.argv.index("--range")
    rng = (int(sys.argv[i + 1]), int(sys.argv[i + 2]))

raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.max", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_lsu.sum",
        "smsp__inst_executed_pipe_alu.sum", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"]
for r in rows[2:3]:
    print("kernel:", r[hdr.index("Kernel Name")][:90] if "Kernel Name" in hdr else "")
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print("  %-70s %-8s %s" % (k, units[i], r[i]))
    for i, h in enumerate(hdr):
        if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio"):
            v = float(r[i].replace(",", ""))
            if v > 0.05:
                print("  stall %-40s %.3f" % (h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], v))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if len(r) < 10 or r[0] in ("Address", "Kernel Name"):
        break
    body.append(r)
tot = sum(int(r[idx["# Samples"]]) for r in body)
print("instructions", len(body), "samples", tot, "executed",
      sum(int(r[idx["Instructions Executed"]]) for r in body))
for s in ["stall_long_sb", "stall_short_sb", "stall_wait", "stall_math", "stall_selected", "stall_not_selected",
          "stall_mio", "stall_lg", "stall_dispatch", "stall_no_inst", "stall_barrier", "stall_branch_resolving"]:
    print("  %-24s %6d" % (s, sum(int(r[idx[s]]) for r in body)))
print("per 100-instruction region: executed, samples")
for k in range(0, len(body), 100):
    e = sum(int(r[idx["Instructions Executed"]]) for r in body[k:k + 100])
    s = sum(int(r[idx["# Samples"]]) for r in body[k:k + 100])
    print("  %5d %12d %7d" % (k, e, s))
sel = range(*rng) if rng else sorted(sorted(range(len(body)), key=lambda i: -int(body[i][idx["# Samples"]]))[:top_n])
for i in sel:
    r = body[i]
    print(i, r[idx["Source"]].strip()[:76].ljust(76), r[idx["# Samples"]], "exec", r[idx["Instructions Executed"]],
          "L", r[idx["stall_long_sb"]], "S", r[idx["stall_short_sb"]], "W", r[idx["stall_wait"]],
          "M", r[idx["stall_mio"]], "G", r[idx["stall_lg"]])
