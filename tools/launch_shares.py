#!/usr/bin/env python3
"""Per-kernel totals and shares from an `ncu --metrics gpu__time_duration.sum --csv` launch list
(tools/profile.sh).  usage: python tools/launch_shares.py gpurun_out/launches.csv > profiles/rNN_launch_shares.txt"""
import csv
import sys
from collections import defaultdict


def main():
    rows = [r for r in csv.reader(open(sys.argv[1], newline="")) if len(r) > 14 and r[0].isdigit()]
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows:
        if r[12] != "gpu__time_duration.sum":
            continue
        ns = float(r[14].replace(",", "")) * {"ns": 1, "us": 1e3, "usecond": 1e3, "msecond": 1e6, "ms": 1e6, "second": 1e9,
                                              "nsecond": 1}.get(r[13], 1)
        name = r[4].replace("hymeta::", "")
        tot[name] += ns
        cnt[name] += 1
    all_ns = sum(tot.values())
    print("# ncu --metrics gpu__time_duration.sum launch list (cold-cache, serialised: compare SHARES)")
    for name, ns in sorted(tot.items(), key=lambda kv: -kv[1]):
        print(f"{name[:70]:70s} n={cnt[name]:5d} total_ns={ns:14.0f} share={ns / all_ns:6.3f} avg_us={ns / cnt[name] / 1e3:10.2f}")


if __name__ == "__main__":
    main()
