#!/usr/bin/env python
"""Top warp-stall locations of a kernel from an ``.ncu-rep`` captured with ``--set full --import-source on``.

    python tools/ncu_stalls.py gpurun_out/prof_x.ncu-rep [N]
"""
import csv
import io
import subprocess
import sys


def main():
    path = sys.argv[1]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    # one table per kernel: a short title row, a header row, then the SASS lines
    tables, i = [], 0
    while i < len(rows):
        if len(rows[i]) <= 3 and i + 1 < len(rows) and "Source" in rows[i + 1]:
            j = i + 2
            while j < len(rows) and not (len(rows[j]) <= 3 and j + 1 < len(rows) and "Source" in rows[j + 1]):
                j += 1
            tables.append((rows[i], rows[i + 1], [r for r in rows[i + 2:j] if len(r) == len(rows[i + 1])]))
            i = j
        else:
            i += 1
    for title, hdr, data in tables:
        report(title, hdr, data, n)


def report(title, hdr, data, n):
    print("=" * 100)
    print(title[1][:120] if len(title) > 1 else title)
    isrc, ins, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    stalls = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[ins] or 0) for r in data)
    print("total samples", tot)
    agg = {}
    for r in data:
        for i in stalls:
            agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i] or 0)
    print("by reason:", ", ".join(f"{k[6:]}={100 * v / max(tot, 1):.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for idx, r in sorted(enumerate(data), key=lambda ir: -int(ir[1][ins] or 0))[:n]:
        st = sorted(((hdr[i][6:], int(r[i] or 0)) for i in stalls if int(r[i] or 0) > 0), key=lambda kv: -kv[1])[:2]
        prev = data[idx - 1][isrc][:44] if idx > 0 else ""
        print(f"{int(r[ins]):6d} {100 * int(r[ins]) / tot:5.1f}% ex={r[iex]:>8} {r[isrc][:58]:58s} {st}  <- {prev}")


if __name__ == "__main__":
    main()
