export GT_NO_REBUILD=1
GT_LIB=ab/fence_ar.so timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612"
for i in 1 2; do
GT_LIB=ab/fence_sc.so $T tools/mp_probe.py 2>/dev/null | grep "N  100000" | sed 's/^/fence.sc.sys:      /'
GT_LIB=ab/fence_ar.so $T tools/mp_probe.py 2>/dev/null | grep "N  100000" | sed 's/^/fence.acq_rel.sys: /'
done | tee gpurun_out/r2_mp_fence_ab.txt
