#!/bin/bash
# tuning aid: run tools/shape_bench.py (DNA 100 taxa, GTR+G4) with every library under tools/_variants/*/ (built with other -D settings)
cp iq-tree_b200/lib/liblktree_cuda.so /tmp/orig.so
for d in tools/_variants/*/; do
  n=$(basename $d)
  cp $d/liblktree_cuda.so iq-tree_b200/lib/liblktree_cuda.so
  for np in ${SIZES:-1000000 125000}; do
    echo -n "$n NP=$np: "
    timeout 90 python tools/shape_bench.py 100 $np 4 4 8 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('lnl_ms %.4f  lnl %.6f  partial_ms_per_launch %.4f' % (d['lnl_ms'], d['lnl'], d['kernels']['partial']['ms'] / d['kernels']['partial']['n']))
    elif 'rror' in l: print(l.strip())
"
  done
done
cp /tmp/orig.so iq-tree_b200/lib/liblktree_cuda.so
