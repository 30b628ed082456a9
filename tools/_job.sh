export GT_NO_REBUILD=1
export GT_LIB=ab/flip.so
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
B="python bench.py --steps 2000 --warmup 50 --no-extras --no-cpu-baseline"
for i in 1 2; do
  GT_NO_EARLY_K0=1 GT_NO_CTL_FLIP=1 $B > gpurun_out/r2_hand_none_$i.json 2>/dev/null
  GT_NO_CTL_FLIP=1 $B > gpurun_out/r2_hand_early_$i.json 2>/dev/null
  GT_NO_EARLY_K0=1 $B > gpurun_out/r2_hand_flip_$i.json 2>/dev/null
  $B > gpurun_out/r2_hand_both_$i.json 2>/dev/null
done
python tools/bench_brief.py gpurun_out/r2_hand_*.json | tee gpurun_out/r2_handover_ab.txt
python bench.py --steps 20 --warmup 5 --no-extras --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py /dev/stdin | tee -a gpurun_out/r2_handover_ab.txt
