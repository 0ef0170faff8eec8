"""Hot spots of an `ncu --page source --csv` export: per kernel, instruction totals grouped by execution
frequency and the top stall-sample instructions."""
import collections, csv, sys

def sections(path):
    rows = list(csv.reader(open(path)))
    cur = None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            yield cur
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and len(r) == len(cur["hdr"]):
            cur["rows"].append(r)

def main(path, top=25):
    for sec in list(sections(path)):
        hdr, data = sec["hdr"], sec["rows"]
        isrc, ins, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        tot_i = sum(int(r[ins]) for r in data); tot_s = sum(int(r[ismp]) for r in data)
        print(f"== {sec['name'][:90]}\n   instructions {tot_i}  samples {tot_s}  sass lines {len(data)}")
        bands = collections.Counter(); nb = collections.Counter()
        for r in data:
            n = int(r[ins]); k = 0 if n == 0 else 10 ** (len(str(n)) - 1) * int(str(n)[0])
            bands[k] += n; nb[k] += 1
        for k in sorted(bands):
            if bands[k] > 0.02 * tot_i:
                print(f"   exec~{k:>9d}: {nb[k]:4d} sass lines, {bands[k]:10d} inst ({100 * bands[k] / tot_i:4.1f}%)")
        order = sorted(range(len(data)), key=lambda i: -int(data[i][ismp]))[:top]
        for i in sorted(order):
            r = data[i]
            print(f"   [{i:4d}] {r[isrc].strip()[:80]:80s} inst={int(r[ins]):9d} samp={int(r[ismp]):6d} ({100 * int(r[ismp]) / max(tot_s, 1):4.1f}%)")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
