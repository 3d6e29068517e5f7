#!/usr/bin/env python3
"""Per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launches_summary.txt"""
import collections, csv, re, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
h = rows[hdr]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3}
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    a = agg.setdefault(re.sub(r"\(.*", "", r[ki]), [0, 0.0])
    a[0] += 1
    a[1] += float(r[vi].replace(",", "")) * scale[r[ui]]
ours = {n: a for n, a in agg.items() if "zvb::" in n or "gemv::" in n}
tot = sum(a[1] for a in agg.values())
step = sum(a[1] for n, a in ours.items() if re.search(r"scan_|merge_|prep_queries|ivf_plan|lm_", n))
print("# %s: %d launches, %.3f ms total (ncu-serialised, cold cache: shares, not absolutes)" % (sys.argv[1], sum(a[0] for a in agg.values()), tot))
print("# kernels inside a search step (prep_queries + scan_* + merge_*): %.3f ms" % step)
for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    inside = bool(re.search(r"(zvb::|gemv::).*(scan_|merge_|prep_queries|ivf_plan|lm_)", n))
    print("%-72s %5d launches %10.3f ms (%8.1f us each)  %5.1f%% of all%s" % (n[:72], c, t, 1e3 * t / c, 100 * t / tot,
          ("  %5.1f%% of step" % (100 * t / step)) if inside and step else "  (setup: synthetic data / ingest)"))
