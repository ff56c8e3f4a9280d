"""Opcode histogram per kernel of the built library (cuobjdump -sass), with the Blackwell-specific mnemonics called out.
usage: python scripts/sass_digest.py [libmaria_atm.so] > profiles/<tag>_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "maria_b200/libmaria_atm.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        hist[kern] = collections.Counter()
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
special = ("FFMA2", "FMUL2", "FADD2", "UBLKCP", "UTMA", "STAS", "SYNCS", "UTC", "LDS.128", "LDG.E.64", "STG.E.EF.128", "MUFU", "DFMA",
           "REDUX", "ATOMG", "LDGSTS")
print(f"SASS opcode digest of {lib} (cuobjdump -sass, sm_100a)")
for k, c in hist.items():
    tot = sum(c.values())
    if tot < 50:
        continue
    print(f"\n== {k}: {tot} instructions")
    sp = {s: sum(v for op, v in c.items() if op.startswith(s)) for s in special}
    print("   " + "  ".join(f"{s} x{v}" for s, v in sp.items() if v))
    print("   top: " + "  ".join(f"{op} {v}" for op, v in c.most_common(14)))
