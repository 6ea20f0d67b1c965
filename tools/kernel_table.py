"""Summarise an ncu per-kernel metrics CSV (tools/gpu_round.sh: *_kernel_metrics.csv) as a table."""
import csv
import sys
from collections import OrderedDict

path = sys.argv[1]
lines = [l for l in open(path) if not l.startswith("==")]
d = OrderedDict()
for x in csv.DictReader(lines):
    e = d.setdefault(x["ID"], {"name": x["Kernel Name"].split("(")[0].replace("void ", "").replace("s2p::", "")[:34], "grid": x["Grid Size"]})
    e[x["Metric Name"]] = float(x["Metric Value"].replace(",", ""))
print(f"{'id':>3} {'kernel':34} {'us':>8} {'tensor%':>7} {'L2->SM B/clk/SM':>15} {'dramR MB':>8} {'dramW MB':>8} {'dram%':>6} {'GHz':>5}")
tot = 0.0
for k, v in d.items():
    cyc = v.get("sm__cycles_elapsed.max", 1.0)
    us = v["gpu__time_duration.sum"] / 1000.0
    tot += us
    print(f"{k:>3} {v['name']:34} {us:8.1f} {v.get('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 0):7.1f} "
          f"{v.get('l1tex__m_xbar2l1tex_read_bytes.sum', 0) / cyc / 148:15.1f} {v.get('dram__bytes_read.sum', 0) / 1e6:8.0f} "
          f"{v.get('dram__bytes_write.sum', 0) / 1e6:8.0f} {v.get('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 0):6.1f} {cyc / us / 1000:5.2f}")
print(f"total {tot:.1f} us over {len(d)} launches")
