"""Development aid: attribute an ncu SASS-level profile to source lines.

    python tools/ncu_by_line.py <report.ncu-rep> <kernel regex> <mangled-name substring> [launch index] [top N]

Joins `ncu --page source --csv` (per-SASS-instruction counts) with `nvdisasm -g` of the built library (file:line of every
instruction) by instruction order, and prints warp-instructions, lanes and stall samples per source line."""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kregex, mangled = sys.argv[1:4]
launch = int(sys.argv[4]) if len(sys.argv) > 4 else 0
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
by_samples = len(sys.argv) > 6 and sys.argv[6] == "samples"  # order by stall samples instead of issued instructions
lib = os.environ.get("AKB_LIB", os.path.join(ROOT, "akarirender-1_b200", "libakari_b200.so"))
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "kernels", lib], cwd=tmp, check=True, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
lines, cur, inside = [], None, False
for l in sass:
    if l.startswith(".text."):
        inside = mangled in l
        continue
    if not inside:
        continue
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        lines.append((int(m.group(1), 16), cur, m.group(2).strip()))
rows = list(csv.reader(subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kregex],
                                      capture_output=True, text=True).stdout.splitlines()))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
i0 = secs[launch]
i1 = secs[launch + 1] if launch + 1 < len(secs) else len(rows)
h = rows[i0 + 1]
ci = {n: h.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}
body = rows[i0 + 2:i1]
print("ncu instrs %d, nvdisasm instrs %d" % (len(body), len(lines)))
assert len(body) == len(lines), "binary differs from the profiled one"
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot = [0, 0, 0]
for r, (off, loc, txt) in zip(body, lines):
    a = agg[loc]
    e, t, s = int(r[ci["Instructions Executed"]]), int(r[ci["Thread Instructions Executed"]]), int(r[ci["# Samples"]])
    a[0] += e; a[1] += t; a[2] += s; a[3] += 1
    tot[0] += e; tot[1] += t; tot[2] += s
print("total warp-instr %d, lanes %.1f, samples %d" % (tot[0], tot[1] / max(1, tot[0]), tot[2]))
src_cache = {}
def src(loc):
    f, n = loc
    for d in ("akarirender-1_b200/csrc", "akarirender-1_b200/host"):
        p = os.path.join(ROOT, d, f)
        if os.path.exists(p):
            if p not in src_cache:
                src_cache[p] = open(p).read().splitlines()
            return src_cache[p][n - 1].strip()[:90] if n - 1 < len(src_cache[p]) else ""
    return ""
for loc, a in sorted(agg.items(), key=lambda kv: -kv[1][2 if by_samples else 0])[:top]:
    if loc is None:
        continue
    print("%-14s:%4d  sass %3d  warp-instr %5.1f%%  lanes %4.1f  samples %5.1f%%  | %s" % (
        loc[0], loc[1], a[3], 100 * a[0] / tot[0], a[1] / max(1, a[0]), 100 * a[2] / max(1, tot[2]), src(loc)))
