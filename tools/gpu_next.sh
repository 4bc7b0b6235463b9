#!/bin/bash
# the job the next gpurun call executes (edited between calls; gpurun snapshots the tree when the call starts)
set -x
python tools/single_kernels.py 1
DWLS_SMALL_GRIDS=0 python tools/single_kernels.py 1
python tools/single_kernels.py 2
python - <<'PY'
import json, subprocess, os, sys
# configs[2] FGS 4K with the default library and the 3-CTAs-per-SM row pass variant
for lib in ("", "exp/variants/libdwls_h16x3.so"):
    env = dict(os.environ)
    if lib: env["DWLS_LIB"] = lib
    out = subprocess.run([sys.executable, "tools/bench_configs.py"], capture_output=True, text=True, env=env).stdout
    try:
        d = json.loads(out[out.index("{"):])
        print(lib or "default", {k: (v.get("ms_per_call"), v.get("kernels_us_per_call")) for k, v in d.items() if "fgs" in k})
    except Exception as e:
        print("failed", lib, e, out[-500:])
PY
python tools/cpu_configs.py
