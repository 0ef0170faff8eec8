mkdir -p gpurun_out/v
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c5 c2 c3; do
timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/main_$wl.json 2>gpurun_out/v/main_$wl.err
done
python tools/bench_kernels.py gpurun_out/v/main_c?.json
