#!/usr/bin/env python
"""stdin: a bench.py JSON line -> the per-kernel table (us per frame, roofline fraction)."""
import json, sys
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        print("value", round(d["value"]), "Mpix/s; pipeline frac", round(d["roofline"]["pipeline"]["frac"], 4), "ms/step", round(d["ms_per_step"], 3))
        for k, v in d["roofline"]["kernels"].items():
            print("  %-24s %7.3f us/frame  frac %s  share %s" % (k, v.get("us_per_frame", 0), v.get("frac"), v["share"]))
