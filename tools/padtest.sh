for pad in 0 300 1100 2100 5000 33000; do
  HA_BENCH_PAD_KB=$pad python bench.py --steps 800 --warmup 50 --no-e2e --no-cpu-baseline --no-other-configs --layout soa 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('pad_kb', $pad, 'step_us', round(d['ms_per_step']*1e3,2), 'post_us', round(r['kernel_us'],2), 'pre_us', round(r['pre_kernel_us'],2))"
done
