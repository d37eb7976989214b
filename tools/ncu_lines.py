"""Aggregate an ncu SASS page by CUDA source line (instructions executed, stall samples).
usage: ncu_lines.py <rep> <lib.so> <kernel-substring> [top]"""
import collections
import csv
import re
import subprocess
import sys
import tempfile

rep, so, kname = sys.argv[1:4]
import os
so = os.path.abspath(so)
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, check=True, capture_output=True)
import glob
cubin = glob.glob(tmp + "/*.cubin")
line_of = {}
for cb in cubin:
    out = subprocess.run(["nvdisasm", "-g", "-c", cb], capture_output=True, text=True).stdout
    cur_fn, cur_line, in_fn = None, None, False
    for ln in out.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            in_fn = kname in m.group(1)
            continue
        if not in_fn:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur_line = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            line_of[int(m.group(1), 16)] = cur_line
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
iw, ie = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Excessive")
base = None
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
tot_i = tot_s = 0
for r in rows[2:]:
    if len(r) <= isamp or not r[ia].startswith("0x"):
        continue
    a = int(r[ia], 16)
    if base is None:
        base = a
    key = line_of.get(a - base)
    agg[key][0] += int(r[ii]); agg[key][1] += int(r[isamp]); agg[key][2] += int(r[iw] or 0); agg[key][3] += int(r[ie] or 0)
    tot_i += int(r[ii]); tot_s += int(r[isamp])
print(f"total warp-instr {tot_i}, samples {tot_s}")
tot_w = sum(v[2] for v in agg.values())
print(f"shared wavefronts {tot_w}")
for k, (i, s, w, e) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{str(k):34s} instr {100*i/tot_i:5.1f}%  samples {100*s/max(tot_s,1):5.1f}%  smem-wf {100*w/max(tot_w,1):5.1f}% (excess {100*e/max(tot_w,1):4.1f}%)")
