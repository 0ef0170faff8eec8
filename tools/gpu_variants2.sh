# A/B of k_geo_aggr ring depth / chunk size variants on c2, c5, c3 (device-timed value + the edgeconv_fwd timer)
mkdir -p gpurun_out
for v in base aga agb agc agd age; do
  so=deepmetv2_b200/csrc/variants/libgraphmet_$v.so
  [ "$v" == "base" ] && so=deepmetv2_b200/csrc/libgraphmet.so
  for w in c2 c5 c3; do
    GRAPHMET_SO=$PWD/$so python bench.py --workload $w --steps 12 --warmup 4 --no-side --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernels']['edgeconv_fwd']
print('$v $w', round(d['value']), round(d['ms_per_step'],4), 'edgeconv_fwd', round(k['avg_us'],1))"
  done
done
