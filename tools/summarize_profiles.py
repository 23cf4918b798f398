#!/usr/bin/env python
"""Turn the scratch captures of tools/profile_r2.sh (gpurun_out/<TAG>_*) into the committed summaries under profiles/:
  launches_<TAG>.csv / launches_<TAG>_summary.txt   per-launch times (ncu --metrics gpu__time_duration.sum), shares per kernel
  <kernel>_<TAG>_metrics.txt                        key counters of one `ncu --set full` capture per hot kernel
  mb_encode_<TAG>_source.txt                        stall samples / instructions per source line of the wavefront kernel
  traffic.json                                      dram bytes per launch per kernel (bench.py's roofline.traffic)
usage: summarize_profiles.py TAG"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
tag = sys.argv[1]
CMD = "python bench.py --steps 1 --warmup 1 --streams 2 --frames-per-step 6 --no-cpu-baseline --pool 24"

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes.sum.per_second",
    "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_alu.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.sum", "sm__cycles_elapsed.max",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def launches():
    src = os.path.join(G, f"{tag}_launches.csv")
    shutil.copy(src, os.path.join(P, f"launches_{tag}.csv"))
    rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    d = collections.defaultdict(list)
    for r in rows[1:]:
        d[r[ki].split("(")[0]].append(float(r[vi].replace(",", "")))
    tot = sum(sum(v) for v in d.values())
    with open(os.path.join(P, f"launches_{tag}_summary.txt"), "w") as f:
        f.write(f"ncu --metrics gpu__time_duration.sum --clock-control none -c 400  {CMD}\n")
        f.write("(serialised, cold-cache launch times; compare SHARES with bench.py's CUDA-event stage times)\n")
        f.write(f"{'kernel':14s} {'launches':>8s} {'avg_us':>10s} {'share':>8s}\n")
        for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
            f.write(f"{k:14s} {len(v):8d} {sum(v) / len(v) / 1e3:10.1f} {100 * sum(v) / tot:7.2f}%\n")


def kernel(kn, traffic):
    rep = os.path.join(G, f"{tag}_{kn}.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, units = rows[0], rows[1]
    short = kn[2:] if kn.startswith("k_") else kn
    with open(os.path.join(P, f"{short}_{tag}_metrics.txt"), "w") as f:
        f.write(f"ncu --set full --clock-control none --import-source on -k regex:{kn} -s 8 -c 2  {CMD}\n")
        f.write("(1080p P frames of two concurrent streams; one column per captured launch)\n")
        for m in METRICS:
            if m in h:
                i = h.index(m)
                f.write(f"{m:82s} {' '.join(f'{r[i]:>16s}' for r in rows[2:])} {units[i]}\n")
    rd, wr = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    n = len(rows) - 2
    traffic[kn] = {"dram_bytes_read": sum(to_bytes(r[rd], units[rd]) for r in rows[2:]) / n,
                   "dram_bytes_write": sum(to_bytes(r[wr], units[wr]) for r in rows[2:]) / n,
                   "launches_captured": n, "capture": f"profiles/{short}_{tag}_metrics.txt"}
    if kn == "k_mb_encode":
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
        tmp = os.path.join(G, f"{tag}_mb_source.csv")
        open(tmp, "w").write(src)
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_lines.py"), tmp, "40"], capture_output=True, text=True).stdout
        open(os.path.join(P, f"mb_encode_{tag}_source.txt"), "w").write("\n".join(l[:200] for l in out.splitlines()) + "\n")


if __name__ == "__main__":
    os.makedirs(P, exist_ok=True)
    launches()
    traffic = {}
    for kn in ("k_mb_encode", "k_deblock", "k_p_code", "k_me_l1", "k_me_l0", "k_cavlc_size", "k_cavlc_write", "k_hpel", "k_ingest", "k_border", "k_pack"):
        kernel(kn, traffic)
    json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
    for f in (f"{tag}_cycles.err",):
        if os.path.exists(os.path.join(G, f)):
            lines = [l for l in open(os.path.join(G, f)) if "cycles" in l]
            open(os.path.join(P, f"wavefront_cycles_{tag}.txt"), "w").write(
                "CC_DEBUG_CYCLES=1 python bench.py --steps 2 --warmup 1 --no-cpu-baseline   (clock64 counters of k_mb_encode, 8 streams; one pair of lines per stream)\n" + "".join(lines))
    ins = os.path.join(G, f"{tag}_insts.csv")
    if os.path.exists(ins):  # one stream alone: per-launch duration / instructions / utilisation
        rows = list(csv.reader(l for l in open(ins) if l.startswith('"')))
        h = rows[0]
        ki, mi, vi, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
        d = collections.OrderedDict()
        for r in rows[1:]:
            d.setdefault((r[ii], r[ki].split("(")[0]), {})[r[mi]] = r[vi]
        with open(os.path.join(P, f"insts_{tag}.txt"), "w") as f:
            f.write("one stream alone on the GPU (bench.py --streams 1 --pool 24): consecutive launches under ncu (serialised, cold cache)\n")
            for (i, k), m in d.items():
                f.write(f"{i:>4} {k:24s} dur_us={float(m['gpu__time_duration.sum'].replace(',', '')) / 1000:8.1f} inst={float(m['smsp__inst_executed.sum'].replace(',', '')) / 1e6:7.2f}M "
                        f"sm%={m['sm__throughput.avg.pct_of_peak_sustained_elapsed'][:5]} issue%={m['smsp__issue_active.avg.pct_of_peak_sustained_active'][:5]}\n")
    print("profiles written for", tag)
