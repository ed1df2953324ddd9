"""Static SASS mnemonic counts per kernel of libtdod.so (cuobjdump -sass): the mnemonics that prove TMA / mbarrier / cluster /
packed-fp32 / DPX use, and the instruction count of each kernel.    python tools/sass_summary.py > profiles/rNN_sass_summary.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "3dod_b200", "tdod", "libtdod.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WATCH = ["UBLKCP", "SYNCS", "UCGABAR", "FFMA2", "FADD2", "FMUL2", "VIADDMNMX", "MATCH", "REDUX", "CREDUX", "BAR", "LDS", "STS", "LDG", "STG", "ATOMS",
         "ATOMG", "RED", "SHFL", "MUFU", "DFMA", "DADD", "DMUL", "HMMA", "UTCMMA", "LDL", "STL"]
print("SASS mnemonic counts per kernel of 3dod_b200/tdod/libtdod.so (cuobjdump -sass, sm_100a; static instruction counts).")
print("UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier operations, UCGABAR = cluster barrier, FFMA2 / FADD2 = packed f32x2,")
print("VIADDMNMX = fused integer add + min (DPX), MATCH / REDUX / CREDUX = warp match / reduce, LDL / STL = local memory (spills).")
print("No HMMA / UTC*MMA anywhere: nothing on this path is a dense contraction (SURVEY 8(d)).\n")
cur, counts, total = None, None, 0
def short(mangled):
    try:
        d = subprocess.run(["cu++filt", mangled], capture_output=True, text=True).stdout.strip()
    except OSError:
        return mangled[:70]
    d = d.replace("void ", "").replace("<unnamed>::", "").replace("(anonymous namespace)::", "")
    return d[:d.rfind("(")] if "(" in d else d


def flush():
    if cur:
        print("%-44s %6d instr  %s" % (short(cur)[:44], total, "  ".join("%s %d" % (k, counts[k]) for k in WATCH if counts[k])))
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        cur, counts, total = m.group(1), collections.Counter(), 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        total += 1
        op = m.group(1)
        for k in WATCH:
            if op == k or op.startswith(k):
                counts[k] += 1
                break
flush()
