#!/bin/bash
# tuning aid: protein 200 taxa x 200k patterns LG-like +G4 with every library under tools/_variants/*/
cp iq-tree_b200/lib/liblktree_cuda.so /tmp/orig.so
for d in tools/_variants/*/; do
  n=$(basename $d)
  cp $d/liblktree_cuda.so iq-tree_b200/lib/liblktree_cuda.so
  echo -n "$n: "
  timeout 120 python tools/shape_bench.py ${SHAPE:-200 200000 20 4} 4 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('lnl_ms %.4f  lnl %.6f  partial_ms_per_launch %.4f' % (d['lnl_ms'], d['lnl'], d['kernels']['partial']['ms'] / d['kernels']['partial']['n']))
    elif 'rror' in l: print(l.strip())
"
done
cp /tmp/orig.so iq-tree_b200/lib/liblktree_cuda.so
