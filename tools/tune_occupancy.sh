#!/bin/bash
# rebuilds the library with different K1 register budgets and times the kernels (run on the GPU box)
for m in 4 5 6 8; do
  GT_NVCC_EXTRA="-DGT_SCORE_MIN_CTAS=$m" python gpu-topology-on-k8s_b200/build.py --force > /dev/null
  echo "== min_ctas=$m"
  python tools/scale_probe.py 4 100000 400000 2>&1 | tail -2
  python tools/scale_probe.py 3 50000 2>&1 | tail -1
done
python gpu-topology-on-k8s_b200/build.py --force > /dev/null
