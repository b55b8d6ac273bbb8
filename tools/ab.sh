#!/bin/bash
# usage (under gpurun): tools/ab.sh <layout> lib1.so lib2.so ...  -- kernel timings of A/B builds of the engine (HANDARM_B200_LIB)
layout=$1; shift
for lib in default "$@"; do
  if [ "$lib" = default ]; then unset HANDARM_B200_LIB; else export HANDARM_B200_LIB=$PWD/$lib; fi
  python bench.py --steps 800 --warmup 50 --no-e2e --no-cpu-baseline --no-other-configs --layout $layout > /tmp/ab.json 2> /tmp/ab.err
  python - "$lib" <<'PY'
import json, sys
try:
    d = json.loads(open("/tmp/ab.json").read().strip().splitlines()[-1]); r = d["roofline"]
    print(sys.argv[1], "step_us", round(d["ms_per_step"] * 1e3, 2), "post_us", round(r["kernel_us"], 2), "pre_us", round(r["pre_kernel_us"], 2), "frac", round(r["frac"], 3),
          "step_frac", round(r["step_frac"], 3))
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open("/tmp/ab.err").read()[-800:])
PY
done
