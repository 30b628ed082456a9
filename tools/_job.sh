export GT_NO_REBUILD=1
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29617 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_2gpu_driver_style_final.json 2>/dev/null
python tools/bench_brief.py gpurun_out/r2_bench_2gpu_driver_style_final.json
