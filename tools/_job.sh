export GT_NO_REBUILD=1
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in ord3 ord4 ord5 ord6 ord7 ord8; do GT_LIB=ab/$v.so $B > gpurun_out/r2_ordb_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_ordb_*.json | tee gpurun_out/r2_bin_order_ab_v3.txt
