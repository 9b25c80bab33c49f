#!/usr/bin/env python
"""Attribute ncu warp-stall samples to CUDA source lines.

ncu's CSV source page carries SASS only, so this aligns it (by instruction order) with
`nvdisasm -g` of the same build, which carries `//## File ..., line N` markers.

usage: ncu_lines.py report.ncu-rep lib.so kernel_substring [launch_index] [top_n]
"""
import collections
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

rep, lib, kname = sys.argv[1], sys.argv[2], sys.argv[3]
launch = int(sys.argv[4]) if len(sys.argv) > 4 else 0
topn = int(sys.argv[5]) if len(sys.argv) > 5 else 30

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
lines_by_fn = {}
for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
    txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    fn, cur, seq = None, None, None
    for ln in txt.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            fn, seq, cur = m.group(1), [], None
            lines_by_fn[fn] = seq
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        if fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
            seq.append(cur)

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rawrows = list(csv.reader(io.StringIO(raw)))
names = [r[rawrows[0].index("Kernel Name")] for r in rawrows[2:]]
starts = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
# the CSV repeats every launch twice (sass + source views); keep unique blocks in launch order
blocks = []
for k, st in enumerate(starts):
    end = starts[k + 1] - 1 if k + 1 < len(starts) else len(rows)
    blocks.append((st, end))
blocks = blocks[::2] if len(blocks) == 2 * len(names) else blocks
sel = [i for i, n in enumerate(names) if kname in n]
bi = sel[launch]
st, end = blocks[bi]
h = rows[st]
ix = {n: i for i, n in enumerate(h)}
data = [r for r in rows[st + 1:end] if len(r) == len(h)]
fn = [f for f in lines_by_fn if kname.replace("<", "I").split("I")[0] in f and len(lines_by_fn[f]) == len(data)]
if not fn:
    cands = {f: len(v) for f, v in lines_by_fn.items() if kname.split("<")[0] in f}
    sys.exit(f"no disassembled function with {len(data)} instructions matches {kname}: {cands}")
seq = lines_by_fn[fn[0]]


def f(r, n):
    try:
        return float(r[ix[n]])
    except Exception:
        return 0.0


tot = sum(f(r, "# Samples") for r in data) or 1.0
agg = collections.defaultdict(lambda: [0.0, 0.0, collections.Counter()])
for r, loc in zip(data, seq):
    a = agg[loc]
    a[0] += f(r, "# Samples")
    a[1] += f(r, "Instructions Executed")
    for s in ("stall_long_sb", "stall_wait", "stall_short_sb", "stall_math", "stall_lg", "stall_branch_resolving", "stall_not_selected"):
        a[2][s[6:]] += f(r, s)
print(f"{names[bi][:70]}: {tot:.0f} samples, {len(data)} SASS instructions")
srcs = {}
for loc, (s, ex, st_) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    if loc is None:
        text = "?"
    else:
        path = [p for p in glob.glob(os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", "**", loc[0]), recursive=True)]
        if path and path[0] not in srcs:
            srcs[path[0]] = open(path[0]).read().splitlines()
        text = srcs[path[0]][loc[1] - 1].strip()[:90] if path else ""
    top = ", ".join(f"{k} {v / max(s, 1):.2f}" for k, v in st_.most_common(2))
    print(f"{s / tot:6.3f} exec={ex:10.0f} {str(loc):28s} [{top}] {text}")
