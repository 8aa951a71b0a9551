#!/bin/bash
# tuning aid: for every library under tools/_variants/*/ print the probe digest (tools/trv_probe.py) and run the tests named in $TESTS
cp iq-tree_b200/lib/liblktree_cuda.so /tmp/orig.so
for d in tools/_variants/*/; do
  n=$(basename $d)
  cp $d/liblktree_cuda.so iq-tree_b200/lib/liblktree_cuda.so
  echo "== $n"
  LKCUDA_POISON=1 timeout 300 python tools/trv_probe.py 100 1000000 1 2>&1 | tail -1
  LKCUDA_POISON=1 timeout 300 python tools/trv_probe.py 100 60011 1 2>&1 | tail -1
  for i in 1 2; do timeout 600 python -m pytest ${TESTS:-tests/test_rate_em.py} -q -m gpu 2>&1 | tail -3; done
done
cp /tmp/orig.so iq-tree_b200/lib/liblktree_cuda.so
