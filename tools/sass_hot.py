"""SASS evidence for profiles/: per hot kernel of libdirt_b200.so the instruction count and the instructions that
show what it is made of (64-bit key reductions, vector float reductions, match.any, the bulk async copy + mbarrier).
    python tools/sass_hot.py > profiles/r2_sass_hot.txt"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "dirt_b200", "libdirt_b200.so")], capture_output=True, text=True).stdout
KERNELS = ("k_setup", "k_grad_assemble", "k_tile_raster", "k_resolve", "k_hiz_build", "k_depth_split2", "k_small_raster", "k_project")
PATS = ["REDG.E.MIN.64", "REDG.E.ADD.F32", "ATOMG", "MATCH.ANY", "REDUX", "UBLKCP", "SYNCS", "SHFL", "LDG.E.128", "STG.E.128", "LDS.128",
        "VOTE", "POPC", "F2I", "MUFU"]
SHOW = ("REDG.E.MIN.64", "REDG.E.ADD.F32x4", "REDG.E.ADD.F32x2", "MATCH.ANY", "UBLKCP", "SYNCS.ARRIVE.TRANS64", "SYNCS.PHASECHK", "REDUX")
print("# SASS evidence (cuobjdump -sass dirt_b200/libdirt_b200.so, sm_100a cubins): per hot kernel, the instruction count, how often")
print("# the telling instructions occur, and one line of each.  Regenerate: python tools/sass_hot.py > profiles/r2_sass_hot.txt")
print("# arch lines of the dump: " + ", ".join(sorted(set(re.findall(r"arch = (sm_\w+)", txt)))))
print()
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n")[0].strip()
    if not any(k in name for k in KERNELS):
        continue
    lines = [l for l in f.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
    ops = collections.Counter()
    for l in lines:
        m = re.search(r"/\*[0-9a-f]{4}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            ops[m.group(2)] += 1
    demangled = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip() or name
    print("== %s  (%d instructions)" % (demangled[:150], len(lines)))
    print("   " + ", ".join("%s x%d" % (p, n) for p in PATS for n in [sum(v for k, v in ops.items() if k.startswith(p))] if n))
    shown = set()
    for l in lines:
        for p in SHOW:
            if p in l and p not in shown:
                shown.add(p)
                print("   " + l.strip()[:150])
    print()
