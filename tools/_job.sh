export GT_NO_REBUILD=1
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in tk9 bigtree early_all; do GT_LIB=ab/$v.so $B > gpurun_out/r2_tkc_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_tkc_*.json | tee gpurun_out/r2_ticket_timing_ab_v3.txt
