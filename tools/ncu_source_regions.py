#!/usr/bin/env python3
"""Splits the per-instruction stall samples of an ncu source page (ncu -i X --page source --csv) into regions delimited by
barrier instructions (BAR / SYNCS / mbarrier waits), and prints per region: instructions, executed warp instructions, samples
and the dominant stall reasons. usage: python tools/ncu_source_regions.py src.csv [min_samples]"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    head = rows[1]
    idx = {h: i for i, h in enumerate(head)}
    stall_cols = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    regions, cur = [], dict(start=0, n=0, exec=0, samples=0, stalls={c: 0 for c in stall_cols}, first="")
    tot = 0
    for k, r in enumerate(rows[2:]):
        if len(r) < len(head):
            continue
        src = r[idx["Source"]].strip()
        if cur["n"] == 0:
            cur["first"] = src
            cur["start"] = k
        cur["n"] += 1
        cur["exec"] += int(r[idx["Instructions Executed"]] or 0)
        sm = int(r[idx["# Samples"]] or 0)
        cur["samples"] += sm
        tot += sm
        for c in stall_cols:
            cur["stalls"][c] += int(r[idx[c]] or 0)
        if src.startswith("BAR") or "SYNCS" in src or src.startswith("EXIT"):
            cur["last"] = src
            regions.append(cur)
            cur = dict(start=0, n=0, exec=0, samples=0, stalls={c: 0 for c in stall_cols}, first="")
    if cur["n"]:
        cur["last"] = "(end)"
        regions.append(cur)
    print(f"total samples {tot}")
    for g in regions:
        top = sorted(g["stalls"].items(), key=lambda kv: -kv[1])[:5]
        print(f"@{g['start']:5d} n={g['n']:4d} exec={g['exec']:9d} samples={g['samples']:7d} ({100.0 * g['samples'] / max(tot, 1):5.1f}%) "
              f"ends [{g['last'][:28]}]  " + " ".join(f"{k[6:]}={v}" for k, v in top if v))


if __name__ == "__main__":
    main()
