#!/bin/bash
# one rank's block of the 8-way split of C3 alone on one GPU: slab count x L2 prefetch of the x slice
for s in 1 2 3; do for pf in 0 1; do
  echo "== shard 0/8 slabs $s prefetch_x $pf"
  python bench.py --shard 0/8 --slabs $s --prefetch-x $pf --steps 30 --no-cpu --no-parity --no-others --per-slab 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('  GNNZ/s %.1f ms %.4f slabs %s kernel_ms %.4f per_slab %s' % (d['value'], d['ms_per_step'], d['config']['slabs'], d['roofline']['kernel_ms_per_step'], [p['ms'] for p in d['config'].get('per_slab',[])]))
    elif 'Error' in l or 'error' in l: print(l.strip())
"
done; done
