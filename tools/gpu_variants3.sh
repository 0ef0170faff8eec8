for v in mb2 mb4; do
  for w in c2 c5 c3; do
    GRAPHMET_SO=$PWD/deepmetv2_b200/csrc/variants/libgraphmet_$v.so python bench.py --workload $w --steps 12 --warmup 4 --no-side --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernels']['edgeconv_fwd']
print('$v $w', round(d['value']), round(d['ms_per_step'],4), 'edgeconv_fwd', round(k['avg_us'],1))"
  done
done
