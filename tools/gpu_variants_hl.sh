# headline (timers-off) step time of library variants: tools/gpu_variants_hl.sh name1 name2 ...   ("base" = the in-tree build)
for v in "$@"; do
  for w in c2 c5; do
    if [ "$v" == "base" ]; then so=$PWD/deepmetv2_b200/csrc/libgraphmet.so; else so=$PWD/deepmetv2_b200/csrc/variants/libgraphmet_$v.so; fi
    GRAPHMET_SO=$so python bench.py --workload $w --steps 20 --warmup 4 --no-side --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernels']
print('$v $w', round(d['value']), round(d['ms_per_step'],4), 'instr', round(d['instrumented_ms_per_step'],4), 'aggr', round(k['edgeconv_fwd']['avg_us'],1), 'scatter', round(k['scatter_s_bwd']['avg_us'],1))"
  done
done
