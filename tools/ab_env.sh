#!/bin/bash
# usage (under gpurun): tools/ab_env.sh <layout> "<VAR=val ...>" ...  -- kernel timings of the engine under different environment settings
layout=$1; shift
for envs in "" "$@"; do
  env $envs python bench.py --steps 800 --warmup 50 --no-e2e --no-cpu-baseline --no-other-configs --layout $layout > /tmp/ab.json 2> /tmp/ab.err
  python - "$envs" <<'PY'
import json, sys
try:
    d = json.loads(open("/tmp/ab.json").read().strip().splitlines()[-1]); r = d["roofline"]
    print(sys.argv[1] or "default", "step_us", round(d["ms_per_step"] * 1e3, 2), "post_us", round(r["kernel_us"], 2), "pre_us", round(r["pre_kernel_us"], 2), "frac", round(r["frac"], 3))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("/tmp/ab.err").read()[-800:])
PY
done
