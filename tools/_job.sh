export GT_NO_REBUILD=1
python tools/e2e_cycles.py 2>&1 | tee gpurun_out/r2_e2e_cycles.txt
