"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
    python profiles/summarize_launches.py in.csv out.txt "header line" """
import collections, csv, sys
src, dst, title = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    v, u = float(r[vi].replace(",", "")), r[ui]
    us = v / 1e3 if u.startswith("n") else (v if u.startswith("u") else v * 1e3)
    a = agg[r[ki][:100]]
    a[0] += 1
    a[1] += us
tot = sum(a[1] for a in agg.values())
out = [title, "# total %.1f ms over %d launches; per-launch times are cold-cache and serialised: compare SHARES"
       % (tot / 1e3, sum(a[0] for a in agg.values())), "share%  launches  avg_us  kernel"]
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    out.append("%6.2f %7d %10.1f  %s" % (100 * a[1] / tot, a[0], a[1] / a[0], k))
own = sum(a[1] for k, a in agg.items() if "tp::" in k)
out.append("# share of hand-written kernels (tp::*): %.1f %%" % (100 * own / tot))
open(dst, "w").write("\n".join(out) + "\n")
print("\n".join(out))
