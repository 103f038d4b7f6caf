#!/bin/bash
# Round-1 workload sweep on one B200: R-MAT and banded+random shapes, slab counts, kernels.
# Usage (under gpurun): bash tools/sweep_workloads.sh > gpurun_out/sweep.log
short() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('GNNZ/s %.1f  ms %.3f  frac %.3f  slabs %s  nnz %d  setup_s %s' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['config']['slabs'], d['config']['nnz'], d['config']['setup_s']))"; }
for cfg in "--workload rmat --rows 16777216" "--workload rmat --rows 16777216 --slabs 1" "--workload rmat --rows 16777216 --slabs 2" \
           "--workload rmat --rows 16777216 --kernel 5" \
           "--workload banded-power --rows 16777216" "--workload banded-power --rows 33554432" "--workload banded-power --rows 33554432 --slabs 1" \
           "--workload banded-power --rows 33554432 --slabs 2" "--workload banded-power --rows 33554432 --slabs 8"; do
  echo "== $cfg"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu $cfg 2>&1 | tail -1 | short
done
