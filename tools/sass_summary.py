"""SASS opcode counts per kernel of gnn_hex_b200/libhexgnn.so (cuobjdump -sass):  python tools/sass_summary.py > profiles/sass_rN.txt
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk, SYNCS = mbarrier, LDGMC = multimem.ld_reduce (NVSwitch)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gnn_hex_b200", "libhexgnn.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
OPS = ["UTCHMMA", "LDTM", "STTM", "UBLKCP", "UTMALDG", "SYNCS", "FADD2", "HMMA", "ATOMG", "ATOMS", "RED", "LDGMC"]
cur, counts = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if cur and m:
        op = m.group(1)
        counts[cur]["instructions"] += 1
        for o in OPS:
            if op == o or op.startswith(o + "."):
                counts[cur][o] += 1
print("SASS opcode counts per kernel of gnn_hex_b200/libhexgnn.so (cuobjdump -sass, sm_100a; tools/sass_summary.py).")
print("UTCHMMA = tcgen05.mma (kind::f16 and kind::tf32), LDTM / STTM = tcgen05.ld / st (tensor memory), UBLKCP = cp.async.bulk,")
print("SYNCS = mbarrier, FADD2 = packed f32x2 add, LDGMC = multimem.ld_reduce through the NVSwitch (multimem.st shows as STG.E.128.STRONG.SYS).  HMMA = legacy mma.sync.\n")
for fn, c in counts.items():
    name = demangle(fn)
    if not any(c[o] for o in OPS if o not in ("ATOMG", "ATOMS", "RED")) and "lw_" not in name and "hexara" not in name:
        continue
    print(name)
    print("    instructions %6d  " % c["instructions"] + "  ".join(f"{o} {c[o]}" for o in OPS if c[o]))
