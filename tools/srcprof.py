#!/usr/bin/env python
"""Summarise an `ncu --page source --csv` export: top SASS instructions by stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
ilsb = hdr.index("stall_long_sb")
data = [r for r in rows[2:] if len(r) > isamp and r[isamp].isdigit()]
tot = sum(int(r[isamp]) for r in data)
print("total samples", tot, "instructions", len(data))
for k, r in sorted(enumerate(data), key=lambda kr: -int(kr[1][isamp]))[:top]:
    print("%5d %6.2f%%  ex=%9s lsb=%5s  %s" % (k, 100.0 * int(r[isamp]) / tot, r[iex], r[ilsb], r[isrc][:110]))
