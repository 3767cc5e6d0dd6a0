#!/usr/bin/env python3
"""ncu launch-list CSV (gpu__time_duration.sum) -> markdown table of per-kernel totals and shares."""
import collections, csv, sys
src, dst, title = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(open(src)))
hdr = next(r for r in rows if "Kernel Name" in r)
ik, iu, iv = hdr.index("Kernel Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if len(r) == len(hdr) and r[0].isdigit():
        agg[r[ik]][0] += 1; agg[r[ik]][1] += float(r[iv].replace(",", "")) * scale.get(r[iu], 1e-6)
tot = sum(v[1] for v in agg.values())
with open(dst, "w") as f:
    f.write(f"# {title}\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares)\n\n"
            "| kernel | launches | total ms | avg ms | share |\n|---|---|---|---|---|\n")
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| `{k[:100]}` | {n} | {ms:.3f} | {ms / n:.4f} | {100 * ms / tot:.2f}% |\n")
print(open(dst).read())
