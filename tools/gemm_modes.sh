#!/bin/bash
# A/B of the D=784 GEMM kernels: FLOWCUDA_DENSE_TC = 1cta (single-CTA kernel everywhere) | pair (2-CTA clusters everywhere) |
# quad (4-CTA clusters) | policy (unset: the per-Linear default of dt2_mode)
# usage: tools/gemm_modes.sh <tag>   (writes gpurun_out/gemm_<tag>_*.log)
tag=${1:-x}
for mode in ${MODES:-pair quad 1cta policy}; do
  export FLOWCUDA_DENSE_TC=$mode
  [ $mode = policy ] && unset FLOWCUDA_DENSE_TC
  echo "=== mode $mode"
  timeout 300 python -m pytest tests/test_gpu_tc.py tests/test_gpu_made.py tests/test_gpu_baseline_sizes.py -m gpu -q -x -k "dense or made or mnist or deep16" 2>&1 | tail -4
  for wl in maf_mnist maf_made deep16_nvp; do
    timeout 200 python bench.py --workload $wl --only-kernel --steps 20 2> gpurun_out/gemm_${tag}_${mode}_${wl}.err | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$mode', d['config']['workload'], 'ms', round(d['ms_per_step'], 4), 'frac', round(d['roofline']['frac'], 4))
"
  done
done
