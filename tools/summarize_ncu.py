#!/usr/bin/env python
"""Turn an .ncu-rep into the two text summaries kept under profiles/: key metrics + hot SASS regions."""
import csv, subprocess, sys, io

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed.sum.per_cycle_elapsed", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]

def run(args):
    return subprocess.run(["ncu", "-i", *args], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout

def main(rep, out):
    raw = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units, vals = raw[0], raw[1], raw[2]
    lines = [f"# {rep}", "", "## key metrics (ncu --set full --clock-control none)", ""]
    idx = {h: i for i, h in enumerate(hdr)}
    lines.append("kernel: " + vals[idx.get("Kernel Name", 4)])
    for k in KEYS:
        if k in idx:
            lines.append(f"{k:90s} {vals[idx[k]]:>16s} {units[idx[k]]}")
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv"]))))[2:]
    tot = sum(int(r[5]) for r in src) or 1
    samp = sum(int(r[4]) for r in src) or 1
    lines += ["", f"## hot SASS regions (total warp instructions {tot}, {len(src)} SASS lines)", "",
              "share = % of executed warp instructions; thr = average active threads per instruction", ""]
    segs, cur = [], None
    for i, r in enumerate(src):
        ex = int(r[5])
        if cur is None or not (0.8 * cur["ex"] <= ex <= 1.25 * cur["ex"]):
            cur = {"s": i, "ex": ex, "n": 0, "sum": 0, "samp": 0, "thr": 0}
            segs.append(cur)
        cur["n"] += 1; cur["sum"] += ex; cur["samp"] += int(r[4]); cur["thr"] += int(r[6]); cur["e"] = i
    for s in sorted(sorted(segs, key=lambda s: -s["sum"])[:12], key=lambda s: s["s"]):
        lines.append(f"SASS {s['s']:5d}-{s['e']:5d} ({s['n']:3d} instr) executed {s['ex']:>11d}x  share {100*s['sum']/tot:5.1f}%  "
                     f"stall samples {100*s['samp']/samp:5.1f}%  thr {s['thr']/max(1,s['sum']):4.1f}   starts: {src[s['s']][1].strip()[:44]}")
    open(out, "w").write("\n".join(lines) + "\n")

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
