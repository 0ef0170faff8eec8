# round 2, 2-GPU pass: whole GPU suite incl. the 2-rank NCCL test, N=2 bench (default line with side lines off at N>1)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_2gpu.log
tail -5 gpurun_out/pytest_gpu_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_c2_2gpu.json 2> gpurun_out/bench_c2_2gpu.err; echo "bench2 rc=$?"; tail -c 400 gpurun_out/bench_c2_2gpu.err
head -c 300 gpurun_out/bench_c2_2gpu.json; echo
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_c2_2gpu.json"))
print("N=2 c2", round(d["value"]), d["ms_per_step"], d["workload_stats"]["gradient_exchange"])
for k, v in (d.get("side") or {}).items():
    print(k, round(v["value"]), round(v["ms_per_step"], 3))
PY
