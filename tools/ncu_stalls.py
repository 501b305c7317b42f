#!/usr/bin/env python3
"""Per-opcode warp-stall sampling of a kernel's loop body from an ncu report (--set full --import-source on).

  python tools/ncu_stalls.py gpurun_out/prof.ncu-rep profiles/rNN_stall_by_opcode.txt [min_executed]

FFMA is split by its number of register source operands (FFMA3 = three registers, FFMA2 = two registers
plus a constant / immediate): "selected" samples per instruction ~ issue slots the instruction occupies.
"""
import collections
import csv
import io
import re
import subprocess
import sys


def main(rep, dst, min_exec=30000):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    name = rows[0][1] if len(rows[0]) > 1 else "?"
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    data = []
    for r in rows[2:]:
        if len(r) < len(hdr) or not r[0].startswith("0x"):
            break
        data.append(r)
    tot = collections.Counter()
    cls = collections.defaultdict(lambda: [0, 0, collections.Counter()])
    for r in data:
        for s in stalls:
            tot[s] += int(r[ix[s]] or 0)
        if int(r[ix["Instructions Executed"]] or 0) < min_exec:
            continue  # outside the per-tile loop
        t = re.sub(r"^@!?U?P\w+\s+", "", r[1].strip())
        op = t.split()[0].split(".")[0]
        if op == "FFMA":
            ops = t.split(None, 1)[1].split(",")[1:]
            op = "FFMA%d" % sum(1 for o in ops if re.match(r"\s*[-|]?R\d+", o))
        c = cls[op]
        c[0] += 1
        c[1] += int(r[ix["# Samples"]] or 0)
        for s in stalls:
            c[2][s] += int(r[ix[s]] or 0)
    S = sum(tot.values())
    with open(dst, "w") as fh:
        fh.write(f"# {name}\n# warp-stall samples, whole kernel (%): " + ", ".join(f"{k[6:]} {100 * v / S:.1f}" for k, v in tot.most_common(10)) + "\n")
        fh.write(f"# loop body (instructions executed >= {min_exec} times): {sum(c[0] for c in cls.values())} SASS instructions, "
                 f"{sum(c[1] for c in cls.values())} samples\n")
        fh.write("# opcode  static_count  samples  samples/instr  selected/instr  top stall reasons\n")
        for op, c in sorted(cls.items(), key=lambda kv: -kv[1][1]):
            top = ", ".join(f"{k[6:]} {v}" for k, v in c[2].most_common(3))
            fh.write(f"{op:9s} {c[0]:5d} {c[1]:7d} {c[1] / c[0]:8.2f} {c[2]['stall_selected'] / c[0]:8.2f}   {top}\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 30000)
