#!/usr/bin/env python
"""SASS evidence per kernel of libmfa_b200.so (no GPU needed): counts of the instructions that prove the sm_100a
features are used -- UTCHMMA (tcgen05.mma), LDTM / STTM (tcgen05.ld / st), UTMALDG (TMA tensor loads), UTCBAR
(tcgen05.commit), SYNCS (mbarrier), HMMA (mma.sync, decode), MUFU.EX2, LDSM (ldmatrix) -- and of LDL / STL (local-memory
traffic: spills or printf buffers).  Usage: python scripts/sass_summary.py > profiles/sass_r02.txt"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "mini_flash_attention_b200", "libmfa_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
pats = ["UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTCBAR", "SYNCS", "HMMA", "MUFU.EX2", "LDSM", "F2FP", "FMNMX3", "FFMA2", "LDL", "STL", "ELECT", "SETMAXREG"]
cur, tab = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(anonymous namespace\)::", "", cur)
        cur = re.sub(r"\(CUtensorMap_st.*", "", cur)
        tab[cur] = collections.Counter()
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        tab[cur]["total"] += 1
        for p in pats:
            if op.startswith(p):
                tab[cur][p] += 1
print(f"# SASS instruction counts per kernel, {os.path.relpath(lib, ROOT)} (cuobjdump -sass, sm_100a)")
print("# " + " ".join(f"{p:>9s}" for p in ["total"] + pats) + "  kernel")
for k, c in tab.items():
    print("  " + " ".join(f"{c[p]:9d}" for p in ["total"] + pats) + "  " + k)
tot = collections.Counter()
for c in tab.values():
    tot.update(c)
print("  " + " ".join(f"{tot[p]:9d}" for p in ["total"] + pats) + "  ALL KERNELS")
