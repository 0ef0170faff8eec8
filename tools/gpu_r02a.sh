# round 2, first GPU pass (2 GPUs): full GPU test suite incl. the 2-rank NCCL test, smoke, bench at N=1 and N=2
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_c2.err
GM_BENCH_WATCHDOG=500 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_c2_2gpu.json 2> gpurun_out/bench_c2_2gpu.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/bench_c2_2gpu.err
python tools/bench_kernels.py gpurun_out/bench_c2.json gpurun_out/bench_c2_2gpu.json
