# One measurement round on the GPU box: tests, smoke, bench lines, launch list, ncu captures -> gpurun_out/
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --workload c5 --no-cpu-baseline > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err
python bench.py --workload c3 --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python bench.py --workload c3 --precision bf16 --no-cpu-baseline > gpurun_out/bench_c3_bf16.json 2> gpurun_out/bench_c3_bf16.err
python bench.py --workload c3 --precision tf32 --no-cpu-baseline > gpurun_out/bench_c3_tf32.json 2> gpurun_out/bench_c3_tf32.err
python bench.py --workload c3 --precision bf16 --graph csr --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/bench_c3_bf16_csr.json 2> gpurun_out/bench_c3_bf16_csr.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c3_bf16.csv python bench.py --workload c3 --precision bf16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_geo_aggr -s 6 -c 1 -o gpurun_out/geo_aggr_c2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_geo_aggr -s 2 -c 1 -o gpurun_out/geo_aggr_c3 python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_c3f.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_geo_edgeconv_tc -s 2 -c 1 -o gpurun_out/geo_tc_c3_bf16 python bench.py --workload c3 --precision bf16 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_c3.log 2>&1
PYTHONPATH=. python tools/bench_knn.py > gpurun_out/bench_knn.json 2> gpurun_out/bench_knn.err
PYTHONPATH=. ncu --set full --clock-control none --import-source on -k regex:k_knn_search -s 3 -c 1 -o gpurun_out/knn_c2 python tools/bench_knn.py > gpurun_out/ncu_knn.log 2>&1
cat gpurun_out/bench_knn.json
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; python tools/bench_kernels.py gpurun_out/bench_c2.json gpurun_out/bench_c5.json gpurun_out/bench_c3.json gpurun_out/bench_c3_bf16.json gpurun_out/bench_c3_tf32.json gpurun_out/bench_c3_bf16_csr.json
