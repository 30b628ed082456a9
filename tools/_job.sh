export GT_NO_REBUILD=1
GT_LIB=ab/ahead.so timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  GT_LIB=ab/main3.so $B > gpurun_out/r2_ahead_main3_$i.json 2>/dev/null
  GT_LIB=ab/ahead.so $B > gpurun_out/r2_ahead_on_$i.json 2>/dev/null
  GT_NO_K0_AHEAD=1 GT_LIB=ab/ahead.so $B > gpurun_out/r2_ahead_off_$i.json 2>/dev/null
done
python tools/bench_brief.py gpurun_out/r2_ahead_*.json | tee gpurun_out/r2_k0_ahead_ab.txt
GT_LIB=ab/ahead.so python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py /dev/stdin | tee -a gpurun_out/r2_k0_ahead_ab.txt
