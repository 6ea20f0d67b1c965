"""Copy the judged evidence of one GPU round from gpurun_out/ (scratch) into profiles/ (tracked).
usage: python tools/make_profile_summary.py <tag-in-gpurun_out> <name-in-profiles>"""
import csv
import os
import shutil
import subprocess
import sys

tag, name = sys.argv[1], sys.argv[2]
src = lambda s: os.path.join("gpurun_out", f"{tag}_{s}")
dst = lambda s: os.path.join("profiles", f"{name}_{s}")
os.makedirs("profiles", exist_ok=True)
for f in ("bench.json", "launches.csv", "kernel_metrics.csv", "smi.txt", "plain.log"):
    if os.path.exists(src(f)):
        shutil.copy(src(f), dst(f))
if os.path.exists(src("pytest.log")):
    with open(src("pytest.log")) as f, open(dst("pytest_tail.txt"), "w") as o:
        o.writelines(f.readlines()[-6:])
if os.path.exists(src("kernel_metrics.csv")):
    out = subprocess.run([sys.executable, "tools/kernel_table.py", src("kernel_metrics.csv")], capture_output=True, text=True).stdout
    with open(dst("kernel_table.txt"), "w") as o:
        o.write("# one forward of 128 frames (tools/profile_step.py) under ncu, --clock-control none; per-launch times are\n"
                "# cold-cache and serialised: compare SHARES.  tensor% = sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,\n"
                "# L2->SM = l1tex__m_xbar2l1tex_read_bytes.sum / sm__cycles_elapsed.max / 148, dram = dram__bytes_{read,write}.sum\n")
        o.write(out)
rep = src("prof.ncu-rep")
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    keep = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "dram__bytes_read.sum",
            "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic", "launch__cluster_size", "lts__t_sector_hit_rate.pct",
            "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__inst_executed.sum", "sm__cycles_active.avg"]
    with open(dst("ncu_full_summary.csv"), "w", newline="") as o:
        w = csv.writer(o)
        cols = [i for i, h in enumerate(hdr) if h in keep]
        w.writerow([hdr[i] for i in cols])
        w.writerow([units[i] for i in cols])
        for r in data:
            w.writerow([r[i] for i in cols])
    print(f"wrote {dst('ncu_full_summary.csv')} ({len(data)} launches from ncu --set full)")
print("done")
