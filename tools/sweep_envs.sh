#!/bin/bash
# usage (under gpurun): tools/sweep_envs.sh <layout> N1 N2 ...  -- K_post / K_pre timings against the env count (fixed vs per-tile cost)
layout=$1; shift
for n in "$@"; do
  python bench.py --steps 800 --warmup 50 --no-e2e --no-cpu-baseline --no-other-configs --layout $layout --envs $n > /tmp/ab.json 2> /tmp/ab.err
  python - "$n" <<'PY'
import json, sys
try:
    d = json.loads(open("/tmp/ab.json").read().strip().splitlines()[-1]); r = d["roofline"]
    print("envs", sys.argv[1], "step_us", round(d["ms_per_step"] * 1e3, 2), "post_us", round(r["kernel_us"], 2), "pre_us", round(r["pre_kernel_us"], 2), "frac", round(r["frac"], 3))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("/tmp/ab.err").read()[-800:])
PY
done
