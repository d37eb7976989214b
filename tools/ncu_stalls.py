"""Per source line: stall-reason breakdown of warp-state samples from an ncu SASS page.
usage: ncu_stalls.py <rep> <lib.so> <kernel-substring> [top]"""
import collections, csv, glob, re, subprocess, sys, tempfile
rep, so, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, check=True, capture_output=True)
line_of = {}
for cb in glob.glob(tmp + "/*.cubin"):
    out = subprocess.run(["nvdisasm", "-g", "-c", cb], capture_output=True, text=True).stdout
    cur, inside = None, False
    for ln in out.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m:
            inside = kname in m.group(1); continue
        if not inside: continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m: line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ia = hdr.index("Address")
cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: collections.Counter())
tot = collections.Counter()
base = None
for r in rows[2:]:
    if len(r) <= max(cols) or not r[ia].startswith("0x"): continue
    a = int(r[ia], 16)
    base = a if base is None else base
    key = line_of.get(a - base)
    for c in cols:
        v = int(r[c] or 0)
        if v:
            agg[key][hdr[c][6:]] += v; tot[hdr[c][6:]] += v
n = sum(tot.values())
print("all:", ", ".join(f"{k} {100*v/n:.1f}%" for k, v in tot.most_common(8)))
for key, c in sorted(agg.items(), key=lambda kv: -sum(kv[1].values()))[:top]:
    s = sum(c.values())
    print(f"{str(key):32s} {100*s/n:5.1f}%  " + ", ".join(f"{k} {100*v/s:.0f}%" for k, v in c.most_common(3)))
