# quick A/B: geo tests + c2 / c5 / c3 device-timed lines of the current build
timeout 600 python -m pytest tests/test_gpu_geo.py tests/test_gpu_trainer.py tests/test_gpu_fullsize.py -m gpu -x -q 2>&1 | tail -2
for w in c2 c5 c3; do
  python bench.py --workload $w --steps 12 --warmup 4 --no-side --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernels']
print('$w', round(d['value']), round(d['ms_per_step'],4), {n: round(v['avg_us'],1) for n, v in k.items()})"
done
