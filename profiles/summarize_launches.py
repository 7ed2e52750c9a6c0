"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel for the LAST full step in it.
usage: python profiles/summarize_launches.py gpurun_out/launches.csv"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
names = [r["Kernel Name"] for r in rows]
# a step starts at the forward pack-plan kernel (round 2: varlen packing; round 1 lists: the forward squeeze kernel)
sq = [i for i, n in enumerate(names) if "pack_plan_kernel" in n]
# the weight packs of a step are enqueued (auxiliary stream) just before its pack-plan kernel: they belong to the step
sq = [i - sum(1 for j in (i - 1, i - 2) if j >= 0 and "fold_pack" in names[j]) for i in sq]
per_step = 1
if not sq:
    sq = [i for i, n in enumerate(names) if "squeeze_kernel" in n and "unsqueeze" not in n]
    per_step = 2
# argv[2] = k picks the k-th step from the end (0 = last, which may be cut off by ncu's -c limit)
k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
start = sq[-per_step * (1 + k)] if len(sq) >= per_step * (1 + k) else 0
end = sq[-per_step * k] if k > 0 else len(rows)
sel = rows[start:end]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in sel:
    n = re.sub(r"[(<].*", "", r["Kernel Name"]).replace("void ", "")
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1e3 if r["Metric Unit"] == "ns" else (v * 1e3 if r["Metric Unit"] == "ms" else v)
    agg[n][0] += 1
    agg[n][1] += v
tot = sum(v[1] for v in agg.values())
print(f"launches {len(sel)}  total {tot:.1f} us (cold-cache, serialised: compare shares)")
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"{t:10.1f} us {100 * t / tot:5.1f}% {c:5d}  {n}")
