export GT_NO_REBUILD=1
for v in mixA mixB; do GT_LIB=ab/$v.so timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -1; done
GT_LIB=ab/trace2.so timeout 120 python tools/trace_k1.py > gpurun_out/r2_trace_warm.txt 2>&1
GT_LIB=ab/trace2.so GT_TRACE_COLD=8 timeout 120 python tools/trace_k1.py > gpurun_out/r2_trace_cold.txt 2>&1
for i in 1 2; do for v in main2 mixA mixB; do GT_LIB=ab/$v.so python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline > gpurun_out/r2_mix_${v}_$i.json 2>/dev/null; done; done
python tools/bench_brief.py gpurun_out/r2_mix_*.json | tee gpurun_out/r2_mix_ab.txt
tools/ab_probe.sh "ab/main2.so ab/empty.so" 4 100000 2>&1 | tee gpurun_out/r2_empty_k1.txt
