# backward node kernels compiled for more resident CTAs per SM: bench each variant library (GRAPHMET_SO)
mkdir -p gpurun_out/v
for v in main h4 h4e3 h4c3 h4e3c3; do
  so=deepmetv2_b200/csrc/variants/libgraphmet_$v.so; [ $v == main ] && so=deepmetv2_b200/csrc/libgraphmet.so
  for wl in c2 c5; do
    GRAPHMET_SO=$so timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/${v}_$wl.json 2>gpurun_out/v/${v}_$wl.err
  done
done
python tools/bench_kernels.py gpurun_out/v/main_c?.json gpurun_out/v/h4*_c?.json
