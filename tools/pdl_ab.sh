#!/bin/bash
# A/B: programmatic dependent launch on/off, graph replay vs eager, per D=784 workload
export FLOWCUDA_DENSE_TC=${FLOWCUDA_DENSE_TC:-1cta}
for wl in ${WLS:-maf_mnist maf_made deep16_nvp}; do
  for pdl in 1 0; do
    for graph in 1 0; do
      FLOWCUDA_PDL=$pdl FLOWCUDA_BENCH_GRAPH=$graph timeout 200 python bench.py --workload $wl --only-kernel --steps 20 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print('$wl pdl=$pdl graph=$graph', 'ms', round(d['ms_per_step'], 4))
"
    done
  done
done
