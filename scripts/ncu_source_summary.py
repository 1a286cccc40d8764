"""Summarise an ncu source-page CSV (ncu -i X.ncu-rep --page source --csv): executed warp instructions by opcode,
and by source line when -lineinfo correlation (--import-source on + --print-source... ) is available in SASS view."""
import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
ci = {n: hdr.index(n) for n in ("Source", "Instructions Executed", "Thread Instructions Executed", "# Samples")}
ops = collections.Counter(); thr = collections.Counter(); smp = collections.Counter()
tot = 0; body = []
for r in rows[hi + 1:]:
    if len(r) <= ci["Instructions Executed"]: continue
    try: n = int(r[ci["Instructions Executed"]]); t = int(r[ci["Thread Instructions Executed"]]); s = int(r[ci["# Samples"]] or 0)
    except ValueError: continue
    src = r[ci["Source"]].strip()
    m = re.match(r"(@!?U?P\d+\s+)?([A-Z0-9_]+(\.[A-Z0-9_]+)*)", src)
    op = m.group(2) if m else src[:12]
    base = op.split(".")[0]
    ops[base] += n; thr[base] += t; smp[base] += s; tot += n
    body.append((n, t, s, r[0], src))
print("total warp instructions", tot)
for op, n in ops.most_common(28):
    print("%-10s %14d %5.1f%%  thr/warp %.1f  samples %d" % (op, n, 100.0 * n / tot, thr[op] / max(n, 1), smp[op]))
if len(sys.argv) > 2:
    print("--- hottest instructions")
    for n, t, s, a, src in sorted(body, reverse=True)[: int(sys.argv[2])]:
        print("%12d %5.1f %6d %s %s" % (n, t / max(n, 1), s, a, src[:90]))
