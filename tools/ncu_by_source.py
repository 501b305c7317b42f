#!/usr/bin/env python3
"""Executed warp instructions and stall samples of one kernel aggregated by SOURCE FILE and by function-sized line
ranges, from an ncu report taken with --set full --import-source on (kernels are built with -lineinfo).

  python tools/ncu_by_source.py gpurun_out/prof.ncu-rep [profiles/out.txt]
"""
import collections
import csv
import io
import os
import subprocess
import sys


def main(rep, dst=None):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    lines = []
    per_file = collections.Counter()
    samples_file = collections.Counter()
    per_line = collections.Counter()
    samples_line = collections.Counter()
    text = {}
    cur_file, hdr, ix, cur_line = None, None, None, None
    func = "?"
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = os.path.basename(r[1]); continue
        if r[0] == "Function Name":
            func = r[1]; continue
        if r[0] == "Line No":
            hdr = r; ix = {h: i for i, h in enumerate(hdr)}; continue
        if hdr is None or len(r) < len(hdr):
            continue
        if r[0] != "":
            cur_line = int(r[0]); text[(cur_file, cur_line)] = r[1]; continue
        def num(v):
            try:
                return int(float(v))
            except ValueError:
                return 0
        n = num(r[ix["Instructions Executed"]])
        s = num(r[ix["# Samples"]])
        per_file[cur_file] += n; samples_file[cur_file] += s
        per_line[(cur_file, cur_line)] += n; samples_line[(cur_file, cur_line)] += s
    tot, stot = sum(per_file.values()), sum(samples_file.values())
    lines.append(f"# {rep}: {func}")
    lines.append(f"# executed warp instructions {tot}, stall samples {stot}")
    lines.append("## by source file")
    for f, n in per_file.most_common():
        lines.append(f"{f:28s} {n:12d} {100 * n / tot:6.2f} %   samples {100 * samples_file[f] / max(stot, 1):6.2f} %")
    lines.append("## top source lines")
    for (f, l), n in per_line.most_common(45):
        lines.append(f"{f}:{l:<5d} {n:10d} {100 * n / tot:6.2f} %  smp {100 * samples_line[(f, l)] / max(stot, 1):5.2f} %  {text.get((f, l), '')[:110]}")
    txt = "\n".join(lines) + "\n"
    if dst:
        open(dst, "w").write(txt)
    print(txt)


if __name__ == "__main__":
    main(*sys.argv[1:3])
