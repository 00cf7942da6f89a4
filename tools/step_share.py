"""Per-kernel share of one sampling step from an ncu launch list (gpu__time_duration.sum per launch, tools/step_launches.py):
python tools/step_share.py launches.csv > profiles/<tag>_step_share.txt"""
import csv, re, sys

rows = []
with open(sys.argv[1]) as fh:
    lines = [l for l in fh if l.startswith('"')]
rd = csv.reader(lines)
hdr = next(rd)
col = {h: i for i, h in enumerate(hdr)}
for r in rd:
    if r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    v = float(r[col["Metric Value"]].replace(",", ""))
    u = r[col["Metric Unit"]]
    ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(u, 1)
    name = r[col["Kernel Name"]]
    name = re.sub(r"^void ", "", name).split("(")[0]
    rows.append((name, ns))
agg = {}
for n, ns in rows:
    a = agg.setdefault(n, [0, 0.0]); a[0] += 1; a[1] += ns
tot = sum(a[1] for a in agg.values())
print(f"# {len(rows)} launches, {tot / 1e6:.3f} ms summed kernel time (ncu per-launch times are serialised and cold-cache: compare SHARES with the bench)")
print(f"{'share':>7} {'ms':>9} {'n':>5} {'us/launch':>10}  kernel")
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{100 * a[1] / tot:6.2f}% {a[1] / 1e6:9.3f} {a[0]:5d} {a[1] / a[0] / 1e3:10.2f}  {n}")
