#!/usr/bin/env python3
"""Per-kernel share of ONE bench step from an ncu launch list (--metrics gpu__time_duration.sum --csv).

    python tools/launch_share.py profiles/r01/launches_bench_steps2.csv > profiles/r01/launch_share.md

One full-size step of the capture is taken: from a commit-pass leaf launch (`k_sha256_fixed<0, *>`) to the step's
`k_root_compare`."""
import csv, re, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 14 and r[0].isdigit()]
names = [re.sub(r"\(.*", "", r[4]).replace("void ", "").replace("<unnamed>::", "").strip() for r in rows]
ns = [float(r[14].replace(",", "")) for r in rows]
unit = rows[0][13]
scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "ms": 1.0, "msecond": 1.0}[unit]
# a full-size step: the last commit-pass leaf launch that is as long as the longest one (the end-to-end part
# of the bench runs the same kernels on small session chunks), up to the next root compare
leaf = [i for i, n in enumerate(names) if n.startswith("k_sha256_fixed<0")]
longest = max(ns[i] for i in leaf)
start = [i for i in leaf if ns[i] >= 0.9 * longest and any(n.startswith("k_root_compare") for n in names[i:])][-1]
end = next(i for i in range(start, len(names)) if names[i].startswith("k_root_compare"))
agg, order = {}, []
for n, t in zip(names[start:end + 1], ns[start:end + 1]):
    if n not in agg:
        agg[n] = [0, 0.0]; order.append(n)
    agg[n][0] += 1; agg[n][1] += t * scale
total = sum(v[1] for v in agg.values())
print(f"# ncu launch list of `python bench.py --steps 2 --warmup 3 --no-cpu-baseline` ({sys.argv[1]})\n")
print("One step (commit + verify), per-kernel device time under ncu (serialised, cold cache):\n")
print("| kernel | launches | ms | share |\n|---|---|---|---|")
for n in order:
    print(f"| `{n}` | {agg[n][0]} | {agg[n][1]:.3f} | {100 * agg[n][1] / total:.1f} % |")
print(f"| total | {sum(v[0] for v in agg.values())} | {total:.3f} | |")
