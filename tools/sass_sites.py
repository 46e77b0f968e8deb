#!/usr/bin/env python
"""TMA / mbarrier / packed-SIMD instruction sites per kernel of libabub_cuda.so, from `cuobjdump -sass` of the objects in
autobub3hs_b200/csrc (sm_100a): the evidence that the TMA-fed kernels really issue UTMALDG (cp.async.bulk.tensor) and wait
on SYNCS (mbarrier), and where VIADDMNMX (fused add/max/relu on s16x2), VIMNMX (16-bit packed min/max), VABSDIFF4 (per-byte
|a-b|) and LDGSTS (cp.async) are used.   python tools/sass_sites.py > profiles/r2_sass_sites.txt"""
import collections
import glob
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = ["UTMALDG", "SYNCS", "VIADDMNMX", "VIMNMX", "VABSDIFF4", "LDGSTS", "ATOMS", "SHFL"]


def kernels(obj):
    txt = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
    cur, cnt = None, collections.OrderedDict()
    for l in txt.splitlines():
        m = re.search(r"Function : (\S+)", l)
        if m:
            cur = m.group(1)
            cnt[cur] = collections.Counter()
            continue
        m = re.match(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
        if m and cur:
            op = m.group(1)
            cnt[cur]["VIMNMX" if op == "VIMNMX3" else op] += 1
            cnt[cur]["_n"] += 1
    return cnt


def main():
    print(__doc__.strip().replace("\n", "\n# ").join(["# ", ""]))
    print("#\n# %-34s %-22s %7s " % ("kernel", "file", "instrs") + " ".join("%9s" % c for c in COLS))
    for obj in sorted(glob.glob(os.path.join(ROOT, "autobub3hs_b200", "csrc", "*.o"))):
        for fn, c in kernels(obj).items():
            dem = subprocess.run(["c++filt", fn], capture_output=True, text=True).stdout.strip()
            name = re.sub(r"\(.*", "", dem.replace("(anonymous namespace)::", "").replace("void ", ""))
            if name.startswith("cub::") or name.startswith("void cub::") or "cub::" in name[:20]:
                continue
            print("  %-34s %-22s %7d " % (name[:34], os.path.basename(obj).replace(".o", ".cu"), c["_n"]) + " ".join("%9d" % c[k] for k in COLS))


if __name__ == "__main__":
    main()
