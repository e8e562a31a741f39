"""Dev tool: per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list (profiles/*launches*.csv)."""
import collections, csv, re, sys

def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    tot, cnt = collections.Counter(), collections.Counter()
    for x in csv.DictReader(lines):
        if x.get("Metric Name") != "gpu__time_duration.sum":
            continue
        k = re.sub(r"\(.*", "", x["Kernel Name"])
        v = float(x["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(x["Metric Unit"], 1.0)
        tot[k] += v
        cnt[k] += 1
    s = sum(tot.values())
    for k, v in tot.most_common():
        print(f"{k[:64]:64s} {cnt[k]:4d} {v:9.1f} us {100 * v / s:5.1f}%")
    print(f"{'total':64s} {sum(cnt.values()):4d} {s:9.1f} us")

if __name__ == "__main__":
    main(sys.argv[1])
