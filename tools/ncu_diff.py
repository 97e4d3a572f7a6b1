#!/usr/bin/env python3
"""Side-by-side of two launches of the SAME kernel in one `ncu --page raw --csv --print-units base` export:
every metric whose value differs by more than a threshold, largest relative difference first.
    ncu_diff.py raw.csv <grid A> <grid B> [min relative difference, default 0.05] [out.txt]
Rows are picked by their "Grid Size" column (first launch with that grid after the warm-up ones: the LAST match).
Developer tool (round 2: why do persistent grids lose to one tile per CTA?)."""
import csv
import io
import sys

src, ga, gb = sys.argv[1], sys.argv[2], sys.argv[3]
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.05
text = "".join(ln for ln in open(src, errors="replace") if ln.startswith('"'))
rows = list(csv.reader(io.StringIO(text)))
head, units, data = rows[0], rows[1], rows[2:]
gi = head.index("Grid Size")


def pick(g):
    m = [r for r in data if r[gi].replace(" ", "") == f"({g},1,1)"]
    if not m:
        sys.exit(f"no launch with grid {g}; grids seen: {sorted(set(r[gi] for r in data))}")
    return m[-1]


a, b = pick(ga), pick(gb)
out = []
for i, h in enumerate(head):
    if i <= gi + 2:
        continue
    try:
        x, y = float(a[i].replace(",", "")), float(b[i].replace(",", ""))
    except ValueError:
        continue
    if x == y:
        continue
    rel = abs(x - y) / max(abs(x), abs(y))
    if rel >= thr:
        out.append((rel, h, units[i], x, y))
out.sort(reverse=True)
lines = [f"# {src}: grid {ga} (A) vs grid {gb} (B), metrics differing by >= {thr:.0%}, {len(out)} of {len(head)}",
         f"# {'metric':<88} {'unit':<14} {'A':>18} {'B':>18} {'B/A':>8}"]
for rel, h, u, x, y in out:
    lines.append(f"{h:<90} {u:<14} {x:>18.6g} {y:>18.6g} {(y / x if x else float('inf')):>8.3f}")
text = "\n".join(lines) + "\n"
if len(sys.argv) > 5:
    open(sys.argv[5], "w").write(text)
print(text)
