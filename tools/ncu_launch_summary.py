"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (shares of one step).
usage: python tools/ncu_launch_summary.py launches.csv <launches_per_step> [out.json]"""
import collections, csv, json, re, sys
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
n = int(sys.argv[2])
step = rows[-n:]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in step:
    name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("pfml::<unnamed>::", "").replace("void ", "")
    grid = r["Grid Size"].strip()
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1e3 if r["Metric Unit"] == "ns" else (v * 1e3 if r["Metric Unit"] == "ms" else v)
    key = name if "gemm" not in name else name + " grid=" + grid
    agg[key][0] += 1; agg[key][1] += v
tot = sum(v[1] for v in agg.values())
out = []
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append({"kernel": k, "launches": v[0], "us": round(v[1], 1), "share": round(v[1] / tot, 4)})
    print(f"{v[1]:10.1f} us {100*v[1]/tot:5.1f}% n={v[0]:4d} avg {v[1]/v[0]:8.1f}  {k[:110]}")
print("total us", round(tot, 1))
if len(sys.argv) > 3:
    json.dump({"total_us": tot, "launches": n, "kernels": out}, open(sys.argv[3], "w"), indent=1)
