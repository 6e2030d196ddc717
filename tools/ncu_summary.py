"""Developer helper: key metrics + per-segment instruction/stall shares from an .ncu-rep."""
import csv
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum",
        "l1tex__data_pipe_lsu_wavefronts.sum", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
for vals in rows[2:]:
    d = dict(zip(hdr, vals))
    print("==", d.get("Kernel Name", "")[:80])
    for k in KEYS:
        if k in d:
            print(f"  {k:75s} {d[k]} {units[hdr.index(k)]}")
    for h in hdr:
        if "issue_stalled" in h and "per_issue_active" in h and float(d[h] or 0) > 0.15:
            print(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:28s} {float(d[h]):.2f}")
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    hi = [i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r][0]
    hdr = rows[hi]
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[hdr.index("Instructions Executed")].isdigit()]
    ia, isrc, iss = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
    tot = sum(int(r[ia]) for r in data); tots = sum(int(r[iss]) for r in data)
    segs, cur = [], None
    for k, r in enumerate(data):
        n, s = int(r[ia]), int(r[iss])
        if cur and abs(cur["n"] - n) <= 0.02 * max(cur["n"], 1):
            cur["cnt"] += 1; cur["inst"] += n; cur["samp"] += s; cur["end"] = k
        else:
            cur = {"n": n, "cnt": 1, "inst": n, "samp": s, "start": k, "end": k}; segs.append(cur)
    for s in segs:
        if s["inst"] > tot * 0.004 or s["samp"] > tots * 0.006:
            print(f"lines {s['start']:4d}-{s['end']:4d} n={s['cnt']:3d} exec={s['n']:>11d} inst%={100*s['inst']/tot:5.1f} "
                  f"samp%={100*s['samp']/tots:5.1f} {data[s['start']][isrc].strip()[:46]}")
