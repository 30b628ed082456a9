#!/bin/bash
# A/B several builds of the library inside one GPU session: tools/abn.sh "<a.so> <b.so> ..." [scale_probe args]
libs=$1; shift
for i in 1 2 3; do
  for l in $libs; do
    echo "== $l"; GT_LIB=$l python tools/scale_probe.py "$@" 2>&1 | tail -n +1
  done
done
