# GPU tests + the 2-GPU bench lines (torchrun, NCCL) -> gpurun_out/
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c2 c5; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $wl --no-cpu-baseline > gpurun_out/bench_${wl}_2gpu.json 2> gpurun_out/bench_${wl}_2gpu.err
tail -1 gpurun_out/bench_${wl}_2gpu.json | cut -c1-260
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 1 --warmup 1 2>/dev/null | tail -1 | cut -c1-200
