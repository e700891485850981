"""Per-kernel SASS opcode counts of libnmf_b200.so (cuobjdump -sass): the tcgen05 / TMA / TMEM mnemonics that prove which
hardware paths each kernel uses, and local-memory (spill) instructions.
    python scripts/sass_histogram.py > profiles/r2_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "nmf_gpu_b200", "libnmf_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UTCHMMA", "UTCHMMA.2CTA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS", "MUFU.RCP", "MUFU.LG2", "FFMA", "LDG", "STG",
        "LDS", "STS", "LDL", "STL", "ACQBULK", "ELECT", "REDUX", "SHFL"]
name, counts, total = None, None, 0
rows = []
def flush():
    if name:
        rows.append((name, total, dict(counts)))
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r"\(anonymous namespace\)::", "", name).replace("nmf::", "").split("(")[0][:70]
        counts, total = collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and name:
        op = m.group(1); total += 1
        for k in KEYS:
            if op == k or op.startswith(k + "."):
                counts[k] += 1
        if op.startswith("UTCHMMA") and ".2CTA" in op:
            counts["UTCHMMA.2CTA"] += 1
flush()
print("SASS opcode counts per kernel (cuobjdump -sass %s; sm_100a).  UTCHMMA = tcgen05.mma, UTMALDG = TMA tensor load,\n"
      "LDTM / STTM = tcgen05.ld / st (TMEM), UTCBAR = tcgen05.commit, SYNCS = mbarrier ops, LDL / STL = local memory (spills).\n" % os.path.basename(lib))
print("%-72s %6s  %s" % ("kernel", "instr", "  ".join(KEYS)))
for n, t, c in sorted(rows):
    print("%-72s %6d  %s" % (n, t, "  ".join("%*d" % (len(k), c.get(k, 0)) for k in KEYS)))
