"""Join an ncu SASS source-page CSV with nvdisasm -gi line info: executed warp instructions per CUDA source line.
usage: ncu_by_line.py <ncu source csv> <cubin> <mangled kernel name> [source file] [top]"""
import csv, re, subprocess, sys, collections
src_csv, cubin, kname = sys.argv[1:4]
srcfile = sys.argv[4] if len(sys.argv) > 4 else None
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
dis = subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of = {}; cur = None; infn = False; pending = None
for ln in dis:
    if ln.startswith("//---") and ".text." in ln:
        infn = (".text." + kname + " ") in ln or ln.rstrip("- ").endswith(kname)
        continue
    if not infn: continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        if pending is None: pending = (m.group(1).split('/')[-1], int(m.group(2)))     # innermost frame comes first
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m:
        if pending is not None: cur = pending; pending = None
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; ia = 0; ii = hdr.index("Instructions Executed"); it = hdr.index("Thread Instructions Executed"); isamp = hdr.index("# Samples")
base = None; per = collections.Counter(); thr = collections.Counter(); smp = collections.Counter(); tot = 0
for r in rows[hi + 1:]:
    try: a = int(r[ia], 16); n = int(r[ii]); t = int(r[it]); s = int(r[isamp] or 0)
    except (ValueError, IndexError): continue
    if base is None: base = a
    l = line_of.get(a - base)
    per[l] += n; thr[l] += t; smp[l] += s; tot += n
import os
texts = {}
def src_line(key):
    if not key or not srcfile: return ""
    f, l = key
    if f not in texts:
        path = os.path.join(os.path.dirname(srcfile), f)
        texts[f] = open(path).read().splitlines() if os.path.exists(path) else []
    t = texts[f]
    return t[l - 1].strip()[:100] if 0 < l <= len(t) else ""
print("total warp instructions %d" % tot)
for l, n in per.most_common(top):
    code = src_line(l)
    tag = "%s:%d" % (l[0].replace("rtb_", "").replace(".cuh", "").replace(".cu", ""), l[1]) if l else "?"
    print("%5.1f%%  thr %4.1f  smp %5.1f%%  %-11s %s" % (100.0 * n / tot, thr[l] / max(n, 1), 100.0 * smp[l] / max(sum(smp.values()), 1), tag, code))
