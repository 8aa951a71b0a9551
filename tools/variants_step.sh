#!/bin/bash
# tuning aid: tools/step_bench.py (the timed step of bench.py) with every library under tools/_variants/*/, twice each in alternation
cp iq-tree_b200/lib/liblktree_cuda.so /tmp/orig.so
for pass in 1 2; do
for d in tools/_variants/*/; do
  n=$(basename $d)
  cp $d/liblktree_cuda.so iq-tree_b200/lib/liblktree_cuda.so
  echo -n "$n: "; timeout 300 python tools/step_bench.py 20 3 2>&1 | tail -1
done
done
cp /tmp/orig.so iq-tree_b200/lib/liblktree_cuda.so
