export GT_NO_REBUILD=1
GT_LIB=ab/k0t1.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "masks_change or synthetic or full_size or scoring_loop" 2>&1 | tail -1
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in k0t0 k0t1; do GT_LIB=ab/$v.so $B > gpurun_out/r2_k0t_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_k0t_*.json | tee gpurun_out/r2_k0_trigger_after_wait_ab.txt
