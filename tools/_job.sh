export GT_NO_REBUILD=1
GT_LIB=ab/smid4.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in smid8 smid4 smid6; do GT_LIB=ab/$v.so $B > gpurun_out/r2_smidb_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_smidb_*.json | tee gpurun_out/r2_small_ticket_mid_ab_v2.txt
