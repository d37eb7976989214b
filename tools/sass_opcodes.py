"""Per-kernel SASS opcode evidence for libkws_b200.so (no GPU needed):
    python tools/sass_opcodes.py > profiles/sass_opcodes.txt
Counts, per kernel, the mnemonics that prove the Blackwell path: UTCHMMA (tcgen05.mma), LDTM/STTM (tensor
memory loads/stores), UBLKCP (cp.async.bulk = TMA), SYNCS (mbarrier), UTCBAR (tcgen05.commit), DMMA (fp64 tensor
core), FFMA2/FADD2/FMUL2 (packed f32x2), plus registers per thread."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "keyword-spotting_b200", sys.argv[1] if len(sys.argv) > 1 else "libkws_b200.so")
OPS = ["UTCHMMA", "LDTM", "STTM", "UTCBAR", "UBLKCP", "SYNCS", "DMMA", "FFMA2", "FADD2", "FMUL2", "FFMA", "HMMA", "LDGSTS"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True).stdout
regs = {}
for m in re.finditer(r"Function (\S+):\s*\n\s*REG:(\d+)", res):
    regs[m.group(1)] = int(m.group(2))
demangle = lambda s: subprocess.run(["c++filt", s], capture_output=True, text=True).stdout.strip()
cur, counts, total = None, {}, {}
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = dict.fromkeys(OPS, 0)
        total[cur] = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and cur:
        total[cur] += 1
        op = m.group(1)
        if op in counts[cur]:
            counts[cur][op] += 1
print("kernel | regs | SASS instr | " + " | ".join(OPS))
for k in sorted(counts, key=lambda k: -total[k]):
    name = re.sub(r"\(.*", "", demangle(k))
    print(f"{name} | {regs.get(k, '?')} | {total[k]} | " + " | ".join(str(counts[k][o]) for o in OPS))
