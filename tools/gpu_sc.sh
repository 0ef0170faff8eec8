mkdir -p gpurun_out/v
for v in s128_16 s64_32 s128_24 s128_40; do
for wl in c2 c5; do
GRAPHMET_SO=$PWD/deepmetv2_b200/csrc/variants/libgraphmet_$v.so timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/${v}_$wl.json 2>gpurun_out/v/${v}_$wl.err
done; done
python tools/bench_kernels.py gpurun_out/v/s*_c2.json gpurun_out/v/s*_c5.json | grep -v "^ " ; python tools/bench_kernels.py gpurun_out/v/s*_c2.json gpurun_out/v/s*_c5.json | grep -o "scatter_s_bwd=[0-9.]*"
