"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one training/inference step
(the kernels between the last two occurrences of a marker kernel) with per-kernel device time and share."""
import collections, csv, sys

def main(path, marker="k_reduce_partials"):   # the last kernel of a single-GPU training step (AdamW rides in it)
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    names = [r["Kernel Name"] for r in rows]
    idx = [i for i, n in enumerate(names) if marker in n]
    if len(idx) >= 2:
        s, e = idx[-2] + 1, idx[-1] + 1
    else:
        s, e = 0, len(rows)
    tot = 0.0
    agg = collections.OrderedDict()
    for r in rows[s:e]:
        d = float(r["Metric Value"].replace(",", "")) / 1000.0
        tot += d
        short = r["Kernel Name"].split("(")[0].replace("gm::<unnamed>::", "").replace("void ", "")
        a = agg.setdefault(short, [0, 0.0, r["Grid Size"], r["Block Size"]])
        a[0] += 1; a[1] += d
    print(f"# one step = {e - s} launches, {tot:.1f} us of kernel time (ncu: serialised, cold caches)")
    print(f"{'kernel':44s} {'n':>3s} {'us/launch':>10s} {'us total':>9s} {'share':>6s}  grid x block")
    for k, (n, d, grid, blk) in agg.items():
        print(f"{k[:44]:44s} {n:3d} {d / n:10.1f} {d:9.1f} {100 * d / tot:5.1f}%  {grid} x {blk}")

if __name__ == "__main__":
    main(sys.argv[1], *(sys.argv[2:3]))
