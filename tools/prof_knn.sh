# ncu --set full capture of k_knn_search (c2 shape, 32-d, k=20) -> gpurun_out/$1.ncu-rep (run under gpurun)
PYTHONPATH=. timeout 200 python tools/bench_knn.py 2>&1 | tail -3
PYTHONPATH=. timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_knn_search -s 3 -c 1 -o gpurun_out/$1 python tools/bench_knn.py > gpurun_out/ncu_knn.log 2>&1; tail -1 gpurun_out/ncu_knn.log | cut -c1-120
