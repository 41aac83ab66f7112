"""Summarise an ncu report (read on the CPU box): one row per captured launch with the metrics the
roofline needs.   python tools/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_xxx.md"""
import csv
import subprocess
import sys

COLS = [
    ("time_us", "gpu__time_duration.sum"),
    ("dram_rd_MB", "dram__bytes_read.sum"),
    ("dram_wr_MB", "dram__bytes_write.sum"),
    ("dram_%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("tensor_%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
    ("sm_GHz", "smsp__cycles_elapsed.avg.per_second"),
    ("regs", "launch__registers_per_thread"),
    ("smem_dyn_KB", "launch__shared_mem_per_block_dynamic"),
    ("L2_hit_%", "lts__t_sector_hit_rate.pct"),
    ("grid", "launch__grid_size"),
]


def main():
    rep = sys.argv[1]
    labels = sys.argv[2].split(",") if len(sys.argv) > 2 else []
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print(f"ncu --set full --clock-control none, report `{rep}` (per launch; cold-cache, serialised replays)\n")
    print("| # | kernel | " + " | ".join(c for c, _ in COLS) + " |")
    print("|---|---|" + "---|" * len(COLS))
    for i, r in enumerate(data):
        name = r[hdr.index("Kernel Name")].replace("void cid::", "").split("(")[0]
        if i < len(labels):
            name = f"{labels[i]}: {name}"
        vals = []
        for c, m in COLS:
            if m in hdr:
                v, u = r[hdr.index(m)], units[hdr.index(m)]
                try:
                    f = float(v)
                    if u == "byte":
                        f /= 1e6 if "MB" in c else 1e3
                    elif u == "Kbyte" and "MB" in c:
                        f /= 1e3
                    elif u == "Gbyte" and "MB" in c:
                        f *= 1e3
                    elif u == "ns" and c == "time_us":
                        f /= 1e3
                    elif u == "ms" and c == "time_us":
                        f *= 1e3
                    elif u == "Mhz" and c == "sm_GHz":
                        f /= 1e3
                    vals.append(f"{f:.2f}" if f < 1000 else f"{f:.0f}")
                except ValueError:
                    vals.append(v)
            else:
                vals.append("-")
        print(f"| {i} | {name} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()
