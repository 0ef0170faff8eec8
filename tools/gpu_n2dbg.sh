mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for wl in c2 c5; do
NCCL_DEBUG=WARN GM_BENCH_WATCHDOG=120 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload $wl --no-cpu-baseline > gpurun_out/n2b_$wl.json 2> gpurun_out/n2b_$wl.err
echo "$wl rc=$?"; tail -1 gpurun_out/n2b_$wl.json | cut -c1-160; grep -n "File \"/root/repo\|line [0-9]* in \|Timeout" gpurun_out/n2b_$wl.err | head -30
done
