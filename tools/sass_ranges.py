import csv, os, sys
rows = list(csv.reader(open(sys.argv[1])))
bounds = [int(x) for x in sys.argv[2].split(",")]
blocks = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
which = int(os.environ.get("LAUNCH", "0"))
hdr = rows[blocks[which]]
end = blocks[which + 1] - 1 if which + 1 < len(blocks) else len(rows)
data = [r for r in rows[blocks[which] + 1:end] if len(r) == len(hdr)]
ci = {h: i for i, h in enumerate(hdr)}
I = lambda r, k: int(r[ci[k]])
tot = sum(I(r, "Instructions Executed") for r in data); tots = sum(I(r, "# Samples") for r in data)
bounds = bounds + [len(data)]
for a, b in zip(bounds[:-1], bounds[1:]):
    blk = data[a:b]
    ie = sum(I(r, "Instructions Executed") for r in blk); te = sum(I(r, "Thread Instructions Executed") for r in blk); sm = sum(I(r, "# Samples") for r in blk)
    st = {k: sum(I(r, k) for r in blk) for k in ("stall_long_sb","stall_wait","stall_not_selected","stall_short_sb","stall_math","stall_branch_resolving","stall_selected","stall_no_inst","stall_mio","stall_lg","stall_barrier")}
    top = sorted(st.items(), key=lambda kv: -kv[1])[:5]
    print("[%4d,%4d) inst %5.1f%% thr/i %4.1f smp %5.1f%%  %s" % (a, b, 100.0*ie/tot, te/max(1,ie), 100.0*sm/tots, " ".join("%s %.1f" % (k[6:], 100.0*v/tots) for k, v in top)))
