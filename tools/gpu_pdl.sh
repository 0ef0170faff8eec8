# programmatic dependent launch on (default) vs off (GM_PDL=0): tests once with it on, benches both ways
mkdir -p gpurun_out/v
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c2 c5 c3; do
  for pdl in 1 0; do
    GM_PDL=$pdl timeout 300 python bench.py --workload $wl --no-cpu-baseline --steps 20 > gpurun_out/v/pdl${pdl}_$wl.json 2>gpurun_out/v/pdl${pdl}_$wl.err
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/v/pdl?_c?.json')):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, j['value'], j['ms_per_step'], 'e2e', j['e2e']['value'], j['e2e'].get('pipelined_value'))
    except Exception as e:
        print(f, 'ERR', e)
PY
