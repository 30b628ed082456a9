export GT_NO_REBUILD=1
GT_LIB=ab/cta64.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "synthetic or split or full_size or masks_change" 2>&1 | tail -1
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in cta128 cta64; do GT_LIB=ab/$v.so $B > gpurun_out/r2_cta2_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_cta2_*.json | tee gpurun_out/r2_score_cta64_ab.txt
