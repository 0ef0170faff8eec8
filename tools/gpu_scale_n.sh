# the driver's scaling run at ONE N, as the driver launches it: default flags (c2 + side line c5), K=20 W=5:  bash tools/gpu_scale_n.sh N
n=$1
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
echo "N=$n rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/scale_n$n.json"))
    print("N=$n c2", round(d["value"]), d["ms_per_step"], d["workload_stats"]["gradient_exchange"], "e2e", round(d["e2e"]["value"]))
    for k, v in (d.get("side") or {}).items():
        print("  ", k, round(v["value"]), round(v["ms_per_step"], 3))
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/scale_n$n.err").read()[-1500:])
PY
