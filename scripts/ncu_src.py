#!/usr/bin/env python
"""Per-SASS-instruction hot spots of one kernel from an .ncu-rep (source page).
Usage: python scripts/ncu_src.py <rep> <kernel-name> [top_n]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
# several launches are concatenated: keep the first
hdr_idx = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hdr = rows[hdr_idx[0]]
end = hdr_idx[1] - 1 if len(hdr_idx) > 1 else len(rows)
ia, ie, isamp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
data = []
for r in rows[hdr_idx[0] + 1:end]:
    try:
        data.append((r[ia].strip(), int(r[ie]), int(r[isamp])))
    except Exception:
        pass
tot, ts = sum(d[1] for d in data), sum(d[2] for d in data)
print("instructions", len(data), "executed", tot, "samples", ts)
top = sorted(enumerate(data), key=lambda x: -x[1][2])[:topn]
for i, (s, e, sm) in sorted(top):
    print("%5d %9d %6d  %s" % (i, e, sm, s[:100]))
# cumulative by 100-instruction buckets
print("buckets (start, executed, samples):")
for a in range(0, len(data), 100):
    print(a, sum(d[1] for d in data[a:a + 100]), sum(d[2] for d in data[a:a + 100]))
