"""Turn gpurun_out/*.ncu-rep and launch lists into the small text summaries kept under profiles/.
    python profiles/summarize.py gpurun_out/prof.ncu-rep profiles/r01/knn_tc_full.txt
    python profiles/summarize.py --launches gpurun_out/launches.csv profiles/r01/launches.txt
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_tensor_subpipe_hmma.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.avg.per_cycle_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"]


def full(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    lines = []
    for r in rows[2:]:
        name = r[h.index("Kernel Name")] if "Kernel Name" in h else "?"
        lines.append(f"kernel: {name}")
        for k in KEYS:
            if k in h:
                i = h.index(k)
                lines.append(f"  {k:90s} {r[i]:>18s} {units[i]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    if len(rows) > 2:
        h = rows[1]
        isrc, iss, ie = h.index("Source"), h.index("Warp Stall Sampling (All Samples)"), h.index("Instructions Executed")
        data = [(int(r[ie] or 0), int(r[iss] or 0), r[isrc]) for r in rows[2:] if len(r) > ie]
        ts, ti = sum(d[1] for d in data), sum(d[0] for d in data)
        ops = defaultdict(lambda: [0, 0])
        for e, s, t in data:
            tok = t.split()
            o = (tok[1] if tok and tok[0].startswith("@") else tok[0]).split(".")[0] if tok else "?"
            ops[o][0] += e
            ops[o][1] += s
        lines.append(f"SASS opcode mix (warp instructions executed {ti}, stall samples {ts}):")
        for o, (e, s) in sorted(ops.items(), key=lambda x: -x[1][0])[:18]:
            lines.append(f"  {o:14s} inst {100 * e / max(ti, 1):5.1f}%   samples {100 * s / max(ts, 1):5.1f}%")
        lines.append("top stalled instructions:")
        for e, s, t in sorted(data, key=lambda x: -x[1])[:12]:
            lines.append(f"  {100 * s / max(ts, 1):5.1f}%  exec {e:11d}  {t[:90]}")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:40]))


def launches(csvf, out):
    rows = list(csv.reader(open(csvf)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[hdr + 1:]:
        if len(r) < 10:
            continue
        name = r[4].split("(")[0][-70:]
        tot[name] += float(r[-1].replace(",", ""))
        cnt[name] += 1
    total = sum(tot.values())
    lines = ["ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised: compare shares)",
             f"{'kernel':72s} {'launches':>8s} {'total ms':>10s} {'avg ms':>10s} {'share':>7s}"]
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        lines.append(f"{k:72s} {cnt[k]:8d} {v / 1e6:10.3f} {v / 1e6 / cnt[k]:10.3f} {100 * v / total:6.1f}%")
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    if sys.argv[1] == "--launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[1], sys.argv[2])
