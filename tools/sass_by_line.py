#!/usr/bin/env python3
"""Join an `ncu --page source --csv --print-source sass` dump of ONE kernel launch with `nvdisasm -gi` line info of the same
cubin and print executed instructions / thread-instructions / stall samples per source line and per phase of traverse.cuh.

  cuobjdump -xelf all qaray_b200/libqaray_b200.so; nvdisasm -gi trace_kernels.sm_100a.cubin > trace.sass
  ncu -i X.ncu-rep --page source --csv --kernel-name regex:k_wf_trace --launch-skip 1 --launch-count 1 --print-source sass > src.csv
  python tools/sass_by_line.py trace.sass <mangled-name-substring> src.csv [file-to-attribute-to ...]
"""
import csv
import re
import sys
from collections import defaultdict

sass, fn, src = sys.argv[1], sys.argv[2], sys.argv[3]
prefer = sys.argv[4:] or ["traverse.cuh", "trace_kernels.cu", "trace.cuh"]
lines = open(sass).read().splitlines()
start = [i for i, l in enumerate(lines) if l.startswith(".text.") and fn in l][0]
instr = []  # (where, text)
cur = None
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith(".section"):
        if instr:
            break
    m = re.findall(r'(?:File|inlined at) "([^"]+)", line (\d+)', l)
    m = m[::-1]  # outermost frame first: the line of the preferred file that the instruction belongs to
    if l.strip().startswith("//## File"):
        where = None
        for want in prefer:
            for f, n in m:
                if f.endswith(want):
                    where = (want, int(n))
                    break
            if where:
                break
        cur = where or (m[0][0].split("/")[-1], int(m[0][1]))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        instr.append((cur, l.split("*/", 1)[1].strip()))
rows = list(csv.reader(open(src)))
blocks = [i for i, r in enumerate(rows) if r and r[0] == "Address"]  # one block per launch in the dump
import os
which = int(os.environ.get("LAUNCH", "0"))
hdr_i = blocks[which]
hdr = rows[hdr_i]
end = blocks[which + 1] - 1 if which + 1 < len(blocks) else len(rows)
data = [r for r in rows[hdr_i + 1:end] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
assert len(data) == len(instr), (len(data), len(instr))
per = defaultdict(lambda: [0, 0, 0, 0])
for (where, text), r in zip(instr, data):
    ie, te, smp = int(r[ci["Instructions Executed"]]), int(r[ci["Thread Instructions Executed"]]), int(r[ci["# Samples"]])
    lsb = int(r[ci["stall_long_sb"]])
    p = per[where]
    p[0] += ie; p[1] += te; p[2] += smp; p[3] += lsb
tot = [sum(p[k] for p in per.values()) for k in range(4)]
print("total: inst %d  thread-inst %d  (%.1f thr/inst)  samples %d  long_sb %d" % (tot[0], tot[1], tot[1] / max(1, tot[0]), tot[2], tot[3]))
print("%-28s %8s %6s %8s %8s %8s" % ("line", "inst%", "thr/i", "smp%", "lsb%", ""))
for where, p in sorted(per.items(), key=lambda kv: (str(kv[0][0]), kv[0][1])):
    if p[0] * 1000 < tot[0] and p[2] * 1000 < tot[2]:
        continue
    print("%-28s %8.2f %6.1f %8.2f %8.2f" % ("%s:%d" % where, 100.0 * p[0] / tot[0], p[1] / max(1, p[0]), 100.0 * p[2] / max(1, tot[2]), 100.0 * p[3] / max(1, tot[3])))
