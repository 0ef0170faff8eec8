"""Print the per-kernel-family timings of bench.py JSON lines: python tools/bench_kernels.py file.json [...]"""
import json, sys
for f in sys.argv[1:]:
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f"{f}: {d['value']:.0f} {d['unit']}  {d['ms_per_step']:.3f} ms/step  e2e {d['e2e']['value']:.0f}")
    ks = sorted(d.get("kernels", {}).items(), key=lambda kv: -kv[1]["avg_us"] * kv[1]["regions"])
    print("   " + "  ".join(f"{k}={v['avg_us']:.1f}" for k, v in ks))
    for name, sd in (d.get("side") or {}).items():
        r = sd.get("roofline") or {}
        print(f"   side {name}: {sd['value']:.0f} {sd['unit']}  {sd['ms_per_step']:.3f} ms/step  roofline {r.get('kernel')} frac {r.get('frac')}")
