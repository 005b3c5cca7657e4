#!/usr/bin/env python
"""Warp-stall samples of one kernel of an Nsight Compute report, aggregated per CUDA source line and per SASS opcode.
The report's source page is exported as SASS rows (`ncu -i rep --page source --csv`), the line of every SASS instruction
comes from `nvdisasm -g` on the cubin of the SAME build (compile with -lineinfo), matched by instruction offset.
    python tools/ncu_stall_lines.py <report.ncu-rep> <kernel regex> <object file .o> <mangled kernel name> <source file>"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile


def main():
    rep, kregex, obj, mangled, srcfile = sys.argv[1:6]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kregex],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr = rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = []
    for r in rows[2:]:  # one section per captured launch: the first one
        if len(r) < len(hdr):
            break
        data.append(r)
    with tempfile.TemporaryDirectory() as td:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=td, capture_output=True)
        cubin = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-g", os.path.join(td, cubin)], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(txt) if (".text." + mangled) in l and l.strip().startswith(".section"))
    cur, insts = None, {}
    for l in txt[start + 1:]:
        if l.strip().startswith(".section"):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            insts[int(m.group(1), 16)] = cur
    base = int(data[0][ix["Address"]], 16)
    lines, ops, tot = collections.Counter(), collections.Counter(), 0.0
    for r in data:
        n = float(r[ix["# Samples"]] or 0)
        tot += n
        lines[insts.get(int(r[ix["Address"]], 16) - base)] += n
        s = r[ix["Source"]].split()
        op = s[1] if s and s[0].startswith("@") else (s[0] if s else "?")
        ops[op.split(".")[0]] += n
    src = open(srcfile).read().splitlines()
    print("# %s  kernel %s: %d SASS rows, %d samples (all warps incl. the producer)" % (rep, mangled, len(data), tot))
    print("# share of samples per opcode: " + ", ".join("%s %.1f %%" % (o, 100 * n / tot) for o, n in ops.most_common(10)))
    for k, v in lines.most_common(30):
        text = ""
        if k and k[0] == os.path.basename(srcfile) and k[1] <= len(src):
            text = src[k[1] - 1].strip()[:100]
        print("%6.2f %%  %s:%s  %s" % (100 * v / tot, k[0] if k else "?", k[1] if k else "?", text))


if __name__ == "__main__":
    main()
