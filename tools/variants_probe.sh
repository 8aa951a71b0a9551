#!/bin/bash
# tuning aid: for every library under tools/_variants/*/ run the coverage / identity probe (tools/trv_probe.py, NaN-poisoned engine) at the
# BASELINE shape and, with the slot limit forced down (spill and arena-read paths), at a smaller one; then time the shape.
cp iq-tree_b200/lib/liblktree_cuda.so /tmp/orig.so
for d in tools/_variants/*/; do
  n=$(basename $d)
  cp $d/liblktree_cuda.so iq-tree_b200/lib/liblktree_cuda.so
  echo "== $n"
  LKCUDA_POISON=1 timeout 300 python tools/trv_probe.py 100 1000000 ${RUNS:-2} 2>&1 | grep -v "^$" | tail -${RUNS:-2}
  for ms in 0 1 2; do
    echo -n "maxslots=$ms "; LKCUDA_TRV_MAXSLOTS=$ms LKCUDA_POISON=1 timeout 120 python tools/trv_probe.py 100 60011 1 2>&1 | tail -1
  done
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
