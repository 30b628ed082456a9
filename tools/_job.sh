export GT_NO_REBUILD=1
GT_LIB=ab/tick.so timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  for v in flip tick tick_t1; do GT_LIB=ab/$v.so $B > gpurun_out/r2_tick_${v}_$i.json 2>/dev/null; done
done
python tools/bench_brief.py gpurun_out/r2_tick_*.json | tee gpurun_out/r2_ticket_ab.txt
GT_LIB=ab/tick.so python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py /dev/stdin | tee -a gpurun_out/r2_ticket_ab.txt
