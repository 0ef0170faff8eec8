# scatter-kernel variants (GEO_SC_CH / GEO_SC_STAGES): tools/gpu_variants_sc.sh name1 name2 ...
for v in base "$@"; do
  for w in c2 c5; do
    if [ "$v" == "base" ]; then so=$PWD/deepmetv2_b200/csrc/libgraphmet.so; else so=$PWD/deepmetv2_b200/csrc/variants/libgraphmet_$v.so; fi
    GRAPHMET_SO=$so python bench.py --workload $w --steps 12 --warmup 4 --no-side --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernels']['scatter_s_bwd']
print('$v $w', round(d['value']), round(d['ms_per_step'],4), 'scatter', round(k['avg_us'],1))"
  done
done
