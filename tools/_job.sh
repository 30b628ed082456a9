export GT_NO_REBUILD=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_driver_style_final.json 2> gpurun_out/r2_bench_driver_style_final.err; python tools/bench_brief.py gpurun_out/r2_bench_driver_style_final.json
python bench.py --steps 2000 --warmup 50 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_1gpu_final2.json 2>/dev/null; python tools/bench_brief.py gpurun_out/r2_bench_1gpu_final2.json
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_reference_arm_final.json 2>/dev/null; cut -c1-200 gpurun_out/r2_bench_reference_arm_final.json
python bench.py --steps 16 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_ncu_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 16 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_ncu_l.log 2>&1
python bench.py --steps 16 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_ncu_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"^k_score$|^k_bin$" -s 40 -c 4 -o gpurun_out/r2_k0k1_final python bench.py --steps 16 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2_ncu_f.log 2>&1
ls -la gpurun_out/r2_k0k1_final.ncu-rep
