"""Summarise an `ncu --page raw --csv` dump: one line per profiled launch."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
cols = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg", "sm__cycles_elapsed.avg",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__registers_per_thread",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct"]
units = rows[1]
def val(r, c):
    v = float(r[idx[c]]); u = units[idx[c]]
    if c.startswith("dram__bytes"):
        v *= {"Mbyte": 1.0, "Kbyte": 1e-3, "byte": 1e-6, "Gbyte": 1e3}.get(u, 1.0)
    if c == "gpu__time_duration.sum":
        v *= {"us": 1.0, "ns": 1e-3, "ms": 1e3}.get(u, 1.0)
    return v
print(f"{'kernel':34s} {'us':>8s} {'Minst':>7s} {'issue%':>6s} {'warps%':>6s} {'act/el':>6s} {'rdMB':>8s} {'wrMB':>8s} {'grid':>6s} {'regs':>4s} {'dram%':>6s} {'fma%':>5s} {'L2hit':>5s}")
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")[:34]
    v = [val(r, c) for c in cols]
    print(f"{name:34s} {v[0]:8.1f} {v[1]/1e6:7.2f} {v[2]:6.1f} {v[3]:6.1f} {v[4]/v[5]:6.2f} {v[6]:8.2f} {v[7]:8.2f} {int(v[8]):6d} {int(v[9]):4d} {v[10]:6.1f} {v[11]:5.1f} {v[12]:5.1f}")
