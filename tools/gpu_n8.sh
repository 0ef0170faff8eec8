mkdir -p gpurun_out
for wl in c2 c5; do
GM_BENCH_WATCHDOG=150 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --workload $wl --no-cpu-baseline > gpurun_out/n8_$wl.json 2> gpurun_out/n8_$wl.err
echo "$wl rc=$?"; tail -1 gpurun_out/n8_$wl.json | cut -c1-160
done
