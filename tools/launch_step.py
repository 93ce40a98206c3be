"""Print the launches of the last full step in an ncu launch-list CSV (gpu__time_duration.sum per launch)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi = H.index("Kernel Name"), H.index("Metric Value")
seq = [(r[ki][:72], float(r[vi].replace(",", ""))) for r in rows[hdr + 1:] if len(r) > vi]
heads = [i for i, (k, _) in enumerate(seq) if "tc::head_kernel<" in k]  # the filter opens a step
step = seq[heads[-2]:heads[-1]]
for k, v in step:
    print(f"{v / 1000:9.1f} us  {k}")
print(f"{sum(v for _, v in step) / 1000:9.1f} us  total, {len(step)} launches")
