#!/usr/bin/env python
"""Instruction histograms of the kernels in libllamakotlin_cuda.so (cuobjdump -sass), one block per kernel:
total instructions, the mnemonics that prove the data path (UBLKCP / UTMALDG = TMA, UTCHMMA / LDTM / STTM = tcgen05 +
tensor memory, IDP.4A = DP4A, SYNCS = mbarrier), local-memory traffic (LDL / STL = spills) and the top mnemonics.
No GPU needed.   python tools/sass_summary.py [name filter ...] > profiles/r02_sass_histograms.txt"""
import collections
import hashlib
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "llama.kotlin_b200", "libllamakotlin_cuda.so")
KEY = ["UBLKCP", "UTMALDG", "UTMAPF", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "IDP.4A", "SYNCS", "HMMA", "IMMA", "FFMA", "HFMA2",
       "LDS", "STS", "LDG", "STG", "LDL", "STL", "RED", "ATOM", "ATOMS", "BAR", "SHFL", "PRMT", "LOP3", "FMNMX3", "I2FP", "MUFU",
       "ACQBULK", "ELECT", "R2UR", "CALL", "MEMBAR", "ERRBAR", "CCTL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    sha = hashlib.sha256(open(LIB, "rb").read()).hexdigest()
    print(f"# SASS instruction histograms of {os.path.relpath(LIB, ROOT)} (sha256 {sha[:16]}...), cuobjdump -sass, sm_100a")
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
    want = sys.argv[1:]
    for name, cnt in kernels.items():
        dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip().replace("lk::", "")
        dem = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", dem)
        if want and not any(w in dem for w in want):
            continue
        fam = collections.Counter()
        for k, v in cnt.items():
            for key in KEY:
                if k == key or k.startswith(key + "."):
                    fam[key] += v
                    break
        print(f"\n## {dem}\n  instructions {sum(cnt.values())}")
        print("  key: " + ", ".join(f"{k} {fam[k]}" for k in KEY if fam.get(k)))
        print("  top: " + ", ".join(f"{k} {v}" for k, v in cnt.most_common(12)))


if __name__ == "__main__":
    main()
