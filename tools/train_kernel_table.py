"""Aggregate an ncu --csv metrics log of one training step (tools/profile_train_step.py) by kernel:
launches, total us, share, DRAM bytes and achieved DRAM GB/s, tensor-pipe %.  usage: python tools/train_kernel_table.py <csv>"""
import collections
import csv
import re
import sys

rows = [l for l in open(sys.argv[1]) if l.startswith('"')]
k = collections.OrderedDict()
for x in csv.DictReader(rows):
    d = k.setdefault(x["ID"], {"name": x["Kernel Name"]})
    v = float(x["Metric Value"].replace(",", ""))
    u, m = x["Metric Unit"], x["Metric Name"]
    if m == "gpu__time_duration.sum":
        v = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}.get(u, 1.0) * v
    if m.startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    d[m] = v


def short(n):
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\(.*", "", n)
    return n.replace("s2p::", "").replace("at::native::", "at::")[:58]


agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for d in k.values():
    a = agg[short(d["name"])]
    t = d["gpu__time_duration.sum"]
    a[0] += 1
    a[1] += t
    a[2] += d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0)
    a[3] += d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", 0) * t
tot = sum(a[1] for a in agg.values())
print(f"# one G + D training step under ncu (--clock-control none; cold-cache, serialised launches: compare SHARES); "
      f"{sum(a[0] for a in agg.values())} launches, {tot / 1e3:.1f} ms of kernel time")
print(f"{'kernel':58s} {'n':>4s} {'us':>9s} {'share':>6s} {'dram MB':>9s} {'GB/s':>6s} {'tensor%':>7s}")
for n, a in sorted(agg.items(), key=lambda z: -z[1][1])[:40]:
    print(f"{n:58s} {a[0]:4d} {a[1]:9.0f} {a[1] / tot * 100:5.1f}% {a[2] / 1e6:9.0f} {a[2] / a[1] / 1e3:6.0f} {a[3] / a[1]:7.1f}")
