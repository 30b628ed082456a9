#!/bin/bash
# A/B several builds inside one GPU session, three rounds.  Entries: lib.so[,ENV=VAL...]
#   tools/ab_probe.sh "ab/base.so,gpu-topology-on-k8s_b200/libgputopo.so" [scale_probe args]
entries=$1; shift
for i in 1 2 3; do
  for e in $entries; do
    lib=${e%%,*}; rest=${e#"$lib"}; envs=$(echo "$rest" | tr ',' ' ')
    echo "== $e"; env GT_NO_REBUILD=1 GT_LIB=$lib $envs python tools/scale_probe.py "$@" 2>&1 | tail -n +1
  done
done
