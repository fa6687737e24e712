#!/usr/bin/env python3
"""Basic-block view of an `ncu --page source --csv --print-source sass` dump of one kernel launch: for every run of
instructions with the same execution count, its share of executed instructions, threads per instruction, stall samples and
the memory / shuffle / atomic instructions it contains (enough to recognise the phases of traverse.cuh).
  LAUNCH=k python tools/sass_blocks.py src.csv [min_share_pct]"""
import csv
import os
import sys

rows = list(csv.reader(open(sys.argv[1])))
minshare = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
blocks = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
which = int(os.environ.get("LAUNCH", "0"))
hdr = rows[blocks[which]]
end = blocks[which + 1] - 1 if which + 1 < len(blocks) else len(rows)
data = [r for r in rows[blocks[which] + 1:end] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
I = lambda r, k: int(r[ci[k]])
tot = sum(I(r, "Instructions Executed") for r in data)
tots = sum(I(r, "# Samples") for r in data)
print("kernel:", rows[blocks[which] - 1][1][:80], " instructions:", len(data), " executed:", tot, " thr/inst: %.1f" % (sum(I(r, "Thread Instructions Executed") for r in data) / tot))
out = []
i = 0
while i < len(data):
    j = i
    e = I(data[i], "Instructions Executed")
    while j + 1 < len(data) and I(data[j + 1], "Instructions Executed") == e and not data[j][1].strip().startswith(("BRA", "@", "EXIT", "BSYNC")):
        j += 1
    blk = data[i:j + 1]
    ie = sum(I(r, "Instructions Executed") for r in blk)
    te = sum(I(r, "Thread Instructions Executed") for r in blk)
    sm = sum(I(r, "# Samples") for r in blk)
    lsb = sum(I(r, "stall_long_sb") for r in blk)
    keys = [r[1].split()[0 if not r[1].strip().startswith("@") else 1] for r in blk]
    mem = [k for k in keys if k.startswith(("LDG", "STG", "LDS", "STS", "ATOM", "RED", "SHFL", "VOTE", "LDL", "STL", "MUFU", "WARPSYNC", "MATCH", "REDUX"))]
    out.append((i, len(blk), ie, te, sm, lsb, mem))
    i = j + 1
for (i, n, ie, te, sm, lsb, mem) in out:
    if 100.0 * ie / tot < minshare and 100.0 * sm / tots < minshare:
        continue
    from collections import Counter
    c = Counter(m.split(".")[0] + ("." + m.split(".")[-1] if m.startswith(("LDG", "STG")) else "") for m in mem)
    print("@%4d n=%3d  inst %5.2f%%  thr/i %4.1f  smp %5.2f%%  lsb %5.2f%%  %s" % (i, n, 100.0 * ie / tot, te / max(1, ie), 100.0 * sm / tots, 100.0 * lsb / max(1, sum(o[5] for o in out)),
                                                                        " ".join("%s*%d" % kv for kv in c.items())))
