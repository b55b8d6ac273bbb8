#!/bin/bash
# usage (under gpurun): tools/gpu_check.sh <tag> [extra bench args]  -- GPU parity tests + kernel timings on both state layouts
tag=$1; shift
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for l in isaacgym soa; do
  python bench.py --steps 800 --warmup 50 --no-e2e --no-cpu-baseline --layout $l "$@" > gpurun_out/bench_${tag}_$l.json 2> gpurun_out/bench_${tag}_$l.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${tag}_$l.json").read().strip().splitlines()[-1]); r=d["roofline"]
    print("$l", "step_us", round(d["ms_per_step"]*1e3,2), "post_us", round(r["kernel_us"],2), "pre_us", round(r["pre_kernel_us"],2), "frac", round(r["frac"],3), "step_frac", round(r["step_frac"],3), "warm M/s", round(d["value_l2_warm"]/1e6))
except Exception as e:
    print("$l FAILED", e); print(open("gpurun_out/bench_${tag}_$l.err").read()[-1500:])
PY
done
