# ncu --set full capture of k_geo_edgeconv_tc at c3 bf16 -> gpurun_out/$1.ncu-rep (run under gpurun)
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_geo_edgeconv_tc -s 2 -c 1 -o gpurun_out/$1 python bench.py --workload c3 --precision ${2:-bf16} --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_geo_tc.log 2>&1; tail -1 gpurun_out/ncu_geo_tc.log | cut -c1-120
