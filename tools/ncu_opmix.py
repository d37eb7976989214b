"""Dynamic instruction mix of a kernel from an ncu report's SASS page: executed warp-instructions per opcode.
usage: ncu_opmix.py <rep> [top]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
cnt, smp = collections.Counter(), collections.Counter()
for r in rows[2:]:
    if len(r) <= iex:
        continue
    toks = r[isrc].split()
    if not toks:
        continue
    op = toks[1] if toks[0].startswith("@") and len(toks) > 1 else toks[0]
    parts = op.split(".")
    key = parts[0] + (".MOV" if len(parts) > 1 and parts[0] == "IMAD" and parts[1] == "MOV" else "")
    try:
        cnt[key] += int(r[iex]); smp[key] += int(r[ismp])
    except ValueError:
        pass
tot, ts = sum(cnt.values()), sum(smp.values())
print(f"total warp-instructions {tot}, samples {ts}")
for k, v in cnt.most_common(top):
    print(f"{k:14s} {v:12d} {100.0 * v / tot:5.1f}%   samples {100.0 * smp[k] / max(ts, 1):5.1f}%")
