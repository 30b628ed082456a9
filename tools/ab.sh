#!/bin/bash
# A/B two builds of the library inside one GPU session: tools/ab.sh <base.so> [scale_probe args]
base=$1; shift
for i in 1 2; do
  echo "== new";  python tools/scale_probe.py "$@" 2>&1 | tail -n +1
  echo "== base"; GT_LIB=$base python tools/scale_probe.py "$@" 2>&1 | tail -n +1
done
