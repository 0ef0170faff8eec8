mkdir -p gpurun_out/v
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for sm in 8 3 1; do
for wl in c2 c5; do
GM_GEO_SPLIT_MIN=$sm timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/s${sm}_$wl.json 2>gpurun_out/v/s${sm}_$wl.err
done; done
timeout 300 python bench.py --workload c3 --no-cpu-baseline --steps 20 > gpurun_out/v/main_c3.json 2>gpurun_out/v/main_c3.err
python tools/bench_kernels.py gpurun_out/v/s?_c2.json gpurun_out/v/s?_c5.json gpurun_out/v/main_c3.json
