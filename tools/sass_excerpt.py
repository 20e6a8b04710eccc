#!/usr/bin/env python
"""Per kernel, the SASS mnemonics DESIGN.md cites (count + first occurrence), from `cuobjdump -sass` of the built
library.  No GPU needed.

    python tools/sass_excerpt.py > profiles/sass_excerpt.txt
"""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "ekf-slam_b200", "libekf_slam_b200.so")
KERNELS = ("ekf_downdate_kernelILi64ELi2", "ekf_seq_kernelILi64", "ekf_seq_kernelILi256", "ekf_batch_kernelILi16ELb1ELi2",
           "shard_push_cols", "shard_begin", "mahalanobis_sweep")
KEYS = ["DMMA", "UBLKCP", "UBLKPF", "SYNCS", "LDG.E.NA.ENL2.256", "LDG.E.ENL2.256", "STG.E.ENL2.256", "LDG.E.256", "STG.E.256", "LDG.E.128",
        "STG.E.128", "LDS.128", "STS.128", "LDS.64", "REDG", "ATOMG", "MEMBAR", "CCTL", "REDUX", "SHFL", "MUFU.RCP64H",
        "MUFU.RSQ64H", "DFMA", "DMUL", "DADD"]

txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
print("# cuobjdump -sass ekf-slam_b200/libekf_slam_b200.so (sm_100a): per kernel, count and first occurrence of the SASS")
print("# mnemonics DESIGN.md cites.  Regenerate: python tools/sass_excerpt.py > profiles/sass_excerpt.txt\n")
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = f.split("\n", 1)[0].strip()
    if not any(k in name for k in KERNELS):
        continue
    lines = [l for l in f.split("\n") if re.search(r"/\*[0-9a-f]{4}\*/", l)]
    print(f"## {name}  ({len(lines)} instructions)")
    for k in KEYS:
        hits = [l for l in lines if re.search(r"(?<![A-Z.])" + re.escape(k) + r"(?![0-9])", l)]
        if hits:
            h = hits[0]
            addr = h.split("/*", 2)[1].split("*/")[0]
            ins = re.sub(r"\s+", " ", h.split("*/", 1)[1].split("/*")[0]).strip()
            print(f"  {k:16s} x{len(hits):5d}   first: {addr} {ins}")
    print()
