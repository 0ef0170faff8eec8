mkdir -p gpurun_out/v
V=$PWD/deepmetv2_b200/csrc/variants
timeout 600 python -m pytest tests/test_gpu_geo.py tests/test_gpu_trainer.py -x -q 2>&1 | tail -3
for wl in c5 c2; do
timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/main_$wl.json 2>gpurun_out/v/main_$wl.err
for v in sc256_48 sc256_64; do
GRAPHMET_SO=$V/libgraphmet_$v.so timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/${v}_$wl.json 2>/dev/null
done; done
python tools/bench_kernels.py gpurun_out/v/*.json
