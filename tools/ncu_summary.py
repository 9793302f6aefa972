"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and launch count per kernel."""
import csv, re, sys, collections
path = sys.argv[1]
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.DictReader(lines)
tot = collections.defaultdict(lambda: [0.0, 0])
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "ns")
    ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void ", "", name)
    tot[name][0] += ns
    tot[name][1] += 1
total = sum(v[0] for v in tot.values())
n = sum(v[1] for v in tot.values())
print(f"# {path}: {n} launches, {total/1e6:.1f} ms of kernel time (cold-cache, serialised under ncu: compare SHARES)")
print(f"{'share':>7} {'ms':>9} {'launches':>9} {'avg us':>8}  kernel")
for name, (ns, c) in sorted(tot.items(), key=lambda kv: -kv[1][0])[: int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print(f"{100*ns/total:6.2f}% {ns/1e6:9.2f} {c:9d} {ns/c/1e3:8.1f}  {name[:110]}")
