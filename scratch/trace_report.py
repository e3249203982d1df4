"""Per-kernel timeline of one decode step from a B200_TRACE file: python scratch/trace_report.py <file>
Columns (us): gap = this kernel's CTA-0 entry minus the previous kernel's last exit (negative = resident early),
wait = entry -> dependency wait passed, pro = prologue, body = prologue done -> last CTA exit, span = prev exit -> this exit."""
import collections
import sys

rows = [l.split() for l in open(sys.argv[1])]
rows = [(r[1], *[int(x) for x in r[2:6]]) for r in rows]
agg = collections.defaultdict(lambda: collections.defaultdict(list))
prev_end = None
for name, t0, t1, t2, t3 in rows[5:-5]:          # skip the first / last layer's edges
    if prev_end:
        a = agg[name]
        a["gap"].append((t0 - prev_end) / 1e3); a["wait"].append((t1 - t0) / 1e3); a["pro"].append((t2 - t1) / 1e3)
        a["body"].append((t3 - t2) / 1e3); a["span"].append((t3 - prev_end) / 1e3); a["after_prev"].append((t1 - prev_end) / 1e3)
    prev_end = t3
print(f"{'kernel':9s} {'n':>3s} {'gap':>7s} {'wait':>7s} {'wait-prev_end':>13s} {'pro':>7s} {'body':>7s} {'span':>7s}")
tot = 0
for name, a in agg.items():
    m = lambda k: sum(a[k]) / len(a[k])
    print(f"{name:9s} {len(a['gap']):3d} {m('gap'):7.2f} {m('wait'):7.2f} {m('after_prev'):13.2f} {m('pro'):7.2f} {m('body'):7.2f} {m('span'):7.2f}")
    tot += sum(a["span"])
print(f"sum of spans {tot:.1f} us over {len(rows) - 10} launches; step (first entry -> last exit) {(rows[-1][4] - rows[0][1]) / 1e3:.1f} us")
