#!/usr/bin/env python
"""profiles/ncu_traffic_r01.json from the per-launch summary of an `ncu --set full` capture of one 32-frame step
(tools/ncusum.py output): DRAM bytes read + written per kernel, per frame, averaged over the kernel's launches."""
import csv, json, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
col = {n: hdr.index(n) for n in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum")}
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 32
acc = {}
seen_v = 0
for r in rows[1:]:
    name = r[col["Kernel Name"]]
    if "conf_run_kernel" in name: key = "conf_kernel<left+lrc>" if "<2, 1," in name or ", 1, 1>" in name else "conf_kernel<right>"
    elif "border_kernel" in name: key = "border_kernel"
    elif "weights_kernel" in name: key = "weights_kernel"
    elif "hsolve_kernel" in name: key = "hsolve_kernel"
    elif "vsolve_kernel" in name:
        seen_v += 1
        key = "vsolve_kernel<final>" if seen_v == 3 else "vsolve_kernel"      # third column sweep of the step carries the epilogue
    else: continue
    a = acc.setdefault(key, [0, 0.0, 0.0, 0.0])
    a[0] += 1; a[1] += float(r[col["gpu__time_duration.sum"]]); a[2] += float(r[col["dram__bytes_read.sum"]]) * 1e6; a[3] += float(r[col["dram__bytes_write.sum"]]) * 1e6
out = {"how": "ncu --set full --clock-control none -k regex:conf_run|weights_k|solve_k|border_k --launch-skip 10 --launch-count 10 python bench.py --frames %d --steps 1 "
              "--warmup 1 --no-cpu --no-e2e (the 10 launches of one step); dram__bytes_read.sum + dram__bytes_write.sum per launch divided by the frames of a launch; "
              "summary: profiles/ncu_full_r01_final_summary.csv" % frames, "kernels": {}}
for k, (n, us, rd, wr) in acc.items():
    out["kernels"][k] = {"dram_read_bytes_per_frame": round(rd / n / frames), "dram_write_bytes_per_frame": round(wr / n / frames),
                         "dram_bytes_per_frame": round((rd + wr) / n / frames), "ncu_us_per_launch": round(us / n, 1), "capture": "full_r01_final"}
json.dump(out, sys.stdout, indent=1)
