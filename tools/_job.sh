export GT_NO_REBUILD=1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612"
for i in 1 2; do
GT_LIB=ab/main5.so $T tools/mp_probe.py 2>/dev/null | grep "N  100000" | sed 's/^/previous protocol code:  /'
$T tools/mp_probe.py 2>/dev/null | grep "N  100000" | sed 's/^/general protocol, on K1: /'
done | tee gpurun_out/r2_mp_probe_v5.txt
