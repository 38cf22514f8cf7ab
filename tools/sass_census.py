#!/usr/bin/env python3
"""Opcode census of the shipped kernel library: per kernel, how many tcgen05 / TMA / TMEM /
legacy tensor-core / cp.async instructions the SASS holds (B200_PROFILING.md mnemonics).
usage: python tools/sass_census.py [lib.so] > profiles/rNN_sass_census.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "vllm_hymeta_b200", "csrc", "libhymeta_b200.so")
OPS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "LDTM", "STTM", "SYNCS", "HMMA", "LDSM", "LDGSTS", "MUFU.EX2", "R2P",
       "USETMAXREG", "STL", "LDL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    cur, counts, total = None, collections.OrderedDict(), collections.Counter()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = re.sub(r"\(.*", "", cur).replace("hymeta::", "").replace("void ", "")
            counts[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur:
            total[cur] += 1
            op = m.group(1)
            for o in OPS:
                if op == o or op.startswith(o + ".") or (o == "MUFU.EX2" and op.startswith("MUFU.EX2")):
                    counts[cur][o] += 1
    print(f"# SASS opcode census of {os.path.relpath(LIB, ROOT)} (cuobjdump -sass, sm_100a)")
    print(f"{'kernel':58s} {'instr':>6s} " + " ".join(f"{o:>8s}" for o in OPS))
    for k, c in counts.items():
        print(f"{k[:58]:58s} {total[k]:6d} " + " ".join(f"{c.get(o, 0):8d}" for o in OPS))


if __name__ == "__main__":
    main()
