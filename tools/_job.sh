export GT_NO_REBUILD=1
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -s 2>&1 | tail -25 > gpurun_out/r2_mp_world4_v5.log; tail -2 gpurun_out/r2_mp_world4_v5.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 4 --steps 2000 --warmup 50 --no-cpu-baseline > gpurun_out/r2_bench_4gpu_v5.json 2> gpurun_out/r2_bench_4gpu_v5.err
python tools/bench_brief.py gpurun_out/r2_bench_4gpu_v5.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus 2 --steps 2000 --warmup 50 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_2gpu_v5.json 2> gpurun_out/r2_bench_2gpu_v5.err
python tools/bench_brief.py gpurun_out/r2_bench_2gpu_v5.json
python bench.py --steps 2000 --warmup 50 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_1gpu_v5.json 2> gpurun_out/r2_bench_1gpu_v5.err
python tools/bench_brief.py gpurun_out/r2_bench_1gpu_v5.json
