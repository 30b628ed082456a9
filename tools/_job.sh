export GT_NO_REBUILD=1
for i in 1 2; do
for v in main5 place place_nopeek place_bar; do GT_LIB=ab/$v.so python tools/scale_probe.py 4 100000 2>&1 | tail -1 | sed "s/^/$v /"; done
done | tee gpurun_out/r2_place_probe.txt
