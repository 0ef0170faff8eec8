# round 2, second full pass (1 GPU): whole GPU suite, smoke, default bench with side lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_c2.err
python tools/bench_kernels.py gpurun_out/bench_c2.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_c2.json"))
print("c2", round(d["value"]), d["ms_per_step"], "e2e", round(d["e2e"]["value"]))
for k, v in (d.get("side") or {}).items():
    r = v.get("roofline") or {}
    print(k, round(v["value"]), round(v["ms_per_step"], 3), r.get("kernel"), r.get("frac"), r.get("useful_mma_frac"))
PY
