#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: per-instruction executions and stall samples.

    ncu -i prof.ncu-rep --page source --csv > src.csv ; python tools/ncu_src.py src.csv [top_n]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    rows = list(csv.reader(open(path)))
    hi = next(k for k, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    col = {h: k for k, h in enumerate(hdr)}
    data = rows[hi + 1:]
    def num(r, name):
        try:
            return float(r[col[name]])
        except (ValueError, KeyError, IndexError):
            return 0.0
    tot_s = sum(num(r, "# Samples") for r in data)
    tot_i = sum(num(r, "Instructions Executed") for r in data)
    print(f"instructions executed {tot_i:.0f}   samples {tot_s:.0f}")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    print("stall totals:", {s: int(sum(num(r, s) for r in data)) for s in stalls if sum(num(r, s) for r in data) > 0})
    print(f"{'idx':>5} {'exec':>12} {'thr':>5} {'samples':>8} {'%':>6} {'long':>6} {'short':>6} {'wait':>6} {'br':>5}  sass")
    order = sorted(range(len(data)), key=lambda k: -num(data[k], "# Samples"))[:top]
    for k in sorted(order):
        r = data[k]
        print(f"{k:5d} {num(r, 'Instructions Executed'):12.0f} {num(r, 'Avg. Threads Executed'):5.1f} {num(r, '# Samples'):8.0f} "
              f"{100 * num(r, '# Samples') / max(tot_s, 1):6.2f} {num(r, 'stall_long_sb'):6.0f} {num(r, 'stall_short_sb'):6.0f} "
              f"{num(r, 'stall_wait'):6.0f} {num(r, 'stall_branch_resolving'):5.0f}  {r[col['Source']][:90]}")


if __name__ == "__main__":
    main()
