#!/bin/bash
# usage (under gpurun): tools/ab_side.sh "<cfg keys>" lib1.so ...  -- step / kernel times of side configs for A/B builds of the engine
keys=$1; shift
for lib in default "$@"; do
  if [ "$lib" = default ]; then unset HANDARM_B200_LIB; else export HANDARM_B200_LIB=$PWD/$lib; fi
  for key in $keys; do
    python -c "
import bench, sys
d = bench.measure_other('$key', 'cuda:0', 20)
print('$lib', '$key', 'step_us', round(d['ms_per_step'] * 1e3, 2), 'post_us', round(d['post_kernel_us'], 2), 'pre_us', round(d['pre_kernel_us'], 2), 'step_frac', round(d['step_frac'], 3))
" 2>/dev/null | tail -1
  done
done
