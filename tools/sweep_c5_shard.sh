#!/bin/bash
# One rank's block (rank 3 of 8) of the C5 matrix (128M-row banded + random) alone on one B200: slab count and
# hypersparse-slab kernel.  bash tools/sweep_c5_shard.sh > gpurun_out/c5_shard.log
short() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('GNNZ/s %.1f  ms %.3f  slabs %s  launches %s' % (d['value'], d['ms_per_step'], d['config']['slabs'], d['gpu_launches']))"; }
for cfg in "" "--dcsr 0" "--slabs 18" "--slabs 12" "--slabs 6" "--slabs 4" "--slabs 1"; do
  echo "== shard 3/8 $cfg"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --workload banded-power --shard 3/8 --overwrite $cfg 2>&1 | tail -1 | short
done
