"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel totals over a launch range."""
import csv, sys, collections, re
path = sys.argv[1]
args = [a for a in sys.argv[2:] if not a.startswith("--")]
lo = int(args[0]) if len(args) > 0 else 0
hi = int(args[1]) if len(args) > 1 else 10**9
rows = [r for r in csv.reader(open(path, errors="ignore")) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
seq = [(re.sub(r"\(.*", "", r[ki]).replace("zkb::", "").replace("<unnamed>::", ""), float(r[vi].replace(",", ""))) for r in rows[1:]]
if "--list" in sys.argv:
    for i, (k, v) in enumerate(seq):
        print(i, f"{v/1e3:9.1f} us", k)
    sys.exit()
seq = seq[lo:hi]
tot = sum(v for _, v in seq)
agg = collections.OrderedDict()
for k, v in seq:
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += v
print(f"launches {len(seq)}  total {tot/1e6:.3f} ms")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v/1e3:10.1f} us {100*v/tot:5.1f}%  x{n:<3d} {k}")
