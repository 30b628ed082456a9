#!/bin/bash
# rebuilds the library with different K1 register budgets and times the kernels (run on the GPU box)
for m in 4 5 6 8; do
  GT_NVCC_EXTRA="-DGT_SCORE_MIN_CTAS=$m" python gpu-topology-on-k8s_b200/build.py --force > /dev/null
  python bench.py --steps 500 --warmup 10 --no-cpu-baseline > /tmp/b.json 2>/tmp/b.err || tail -3 /tmp/b.err
  python -c "
import json;d=json.load(open('/tmp/b.json'));print('min_ctas=$m', d['ms_per_step']*1e3, d['kernels_us'])"
done
python gpu-topology-on-k8s_b200/build.py --force > /dev/null
