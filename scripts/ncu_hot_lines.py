#!/usr/bin/env python
"""Print the hottest SASS lines (by warp-stall samples) of `ncu --page source --csv` output."""
import csv
import sys


def main():
    path = sys.argv[1]
    ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
    hdr = rows[hi]
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    i_src, i_s, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")

    def num(x):
        try:
            return int(float(x))
        except ValueError:
            return 0
    tot = sum(num(r[i_s]) for r in data) or 1
    print("total samples", tot, "instructions", len(data))
    names = ["stall_long_sb", "stall_wait", "stall_barrier", "stall_math", "stall_short_sb", "stall_branch_resolving",
             "stall_mio", "stall_lg", "stall_not_selected", "stall_selected", "stall_sleep", "stall_membar"]
    cols = [(h, hdr.index(h)) for h in names if h in hdr]
    top = sorted(((num(r[i_s]), k) for k, r in enumerate(data)), reverse=True)[:ntop]
    for n, k in top:
        r = data[k]
        st = " ".join(f"{h[6:]}={r[i]}" for h, i in cols if r[i] not in ("0", ""))
        print(f"{n:6d} {100 * n / tot:5.1f}% ex={r[i_ex]:>9} | {r[i_src][:64]:64s} | {st}")


if __name__ == "__main__":
    main()
