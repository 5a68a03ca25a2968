#!/usr/bin/env python
"""Aggregate an .ncu-rep's warp-stall samples and executed instructions by CUDA source line.

    python tools/ncu_by_line.py gpurun_out/x.ncu-rep [top_n]
"""
import csv, io, subprocess, sys


def main(rep, top=40):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    cur, line, agg, inst, tot, itot = None, None, {}, {}, 0, 0
    for r in csv.reader(io.StringIO(out)):
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) < 8 or r[0] in ("Line No", "Function Name"):
            continue
        if r[0]:
            try:
                line = (cur, int(r[0]), r[1].strip()[:72])
            except ValueError:
                continue
        if r[2]:
            try:
                s, ex = int(r[6]), int(r[7])
            except ValueError:
                continue
            agg[line] = agg.get(line, 0) + s
            inst[line] = inst.get(line, 0) + ex
            tot += s
            itot += ex
    print(f"samples {tot}  warp instructions {itot}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
        print(f"{100 * v / tot:5.1f}% samples {100 * inst[k] / itot:5.1f}% instr  {k[0]}:{k[1]}  {k[2]}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
