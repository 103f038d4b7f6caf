#!/bin/bash
# One rank's block (rank 3 of 8) of the C4 matrix (R-MAT scale 26, 2^31 entries) alone on one B200: slab count.
short() { python -c "import sys,json; d=json.loads(sys.stdin.read()); print('GNNZ/s %.1f  ms %.3f  slabs %s  launches %s' % (d['value'], d['ms_per_step'], d['config']['slabs'], d['gpu_launches']))"; }
for cfg in "" "--slabs 1" "--slabs 2" "--slabs 4" "--slabs 6" "--slabs 12"; do
  echo "== rmat shard 3/8 $cfg"
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --workload rmat --shard 3/8 --overwrite $cfg 2>&1 | tail -1 | short
done
