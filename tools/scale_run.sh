#!/bin/bash
# usage (under gpurun --gpus 8): tools/scale_run.sh <tag> [extra bench args]  -- the driver's scaling run: N = 1, 2, 4, 8 back to back
tag=$1; shift
for n in 1 2 4 8; do
  if [ $n = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs --layout soa "$@" > gpurun_out/scale_${tag}_n$n.json 2> gpurun_out/scale_${tag}_n$n.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --no-other-configs --layout soa "$@" > gpurun_out/scale_${tag}_n$n.json 2> gpurun_out/scale_${tag}_n$n.err
  fi
  python - $n $tag <<'PY'
import json, sys
n, tag = sys.argv[1], sys.argv[2]
try:
    d = json.loads(open(f"gpurun_out/scale_{tag}_n{n}.json").read().strip().splitlines()[-1]); r = d["roofline"]
    print("N", n, "value G/s", round(d["value"] / 1e9, 3), "us/step", round(d["ms_per_step"] * 1e3, 2), "post", round(r["kernel_us"], 2),
          "pre_in_step", round(r["pre_in_step_us"], 2), "e2e M/s", round(d.get("e2e", {}).get("value", 0) / 1e6, 1), "|", d["stats_exchange"][:70])
except Exception as e:
    print("N", n, "FAILED", e); print(open(f"gpurun_out/scale_{tag}_n{n}.err").read()[-1200:])
PY
done
