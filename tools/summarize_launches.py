#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel share of
device time.  Usage: summarize_launches.py launches.csv ROUNDS [TOP]"""
import collections
import csv
import sys


def main():
    path, rounds = sys.argv[1], int(sys.argv[2])
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for r in rows:
        name = r["Kernel Name"].split("(")[0][:100]
        tot[name] += float(r["Metric Value"])
        cnt[name] += 1
    total = sum(tot.values())
    print("launches/round %d, device time/round %.1f us (serialised under ncu; compare shares)"
          % (len(rows) // rounds, total / rounds / 1e3))
    for name, v in sorted(tot.items(), key=lambda x: -x[1])[:top]:
        print("%9.1f us %5.1f%% x%-3d %s" % (v / rounds / 1e3, 100 * v / total, cnt[name] // rounds, name))


if __name__ == "__main__":
    main()
