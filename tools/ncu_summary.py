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


LAYERS = ["down1.0", "down1.2", "down2.0", "down2.2", "bottleneck.0", "bottleneck.2", "up2", "upconv2.0", "upconv2.2", "up1",
          "upconv1.0", "upconv1.2"]


def traffic_json(hdr, units, data, path, batch, size):
    """profiles/rNN_ncu_traffic.json for bench.py's roofline.traffic: dram bytes per launch of the 12 layer kernels of one
    forward (the report's first 12 rows), with the hash of the library that was profiled."""
    import hashlib
    import json
    from pathlib import Path
    so = Path(__file__).resolve().parent.parent / "celebrity-image-denoiser_b200" / "csrc" / "libcid_b200.so"

    def val(r, m):
        v, u = float(r[hdr.index(m)]), units[hdr.index(m)]
        return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3}.get(u, 1)
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    from cid_b200.build import source_sha
    out = {"batch": batch, "height": size, "width": size, "source_sha": source_sha(),
           "how": "ncu --set full --clock-control none, one forward; dram__bytes_read.sum / dram__bytes_write.sum per launch",
           "layers": {}}
    assert len(data) >= 12, "expected the 12 layer kernels of one forward"
    for name, r in zip(LAYERS, data[:12]):
        out["layers"][name] = {"kernel": r[hdr.index("Kernel Name")].split("(")[0],
                               "dram_read_bytes": val(r, "dram__bytes_read.sum"), "dram_write_bytes": val(r, "dram__bytes_write.sum"),
                               "time_us": val(r, "gpu__time_duration.sum")}
    Path(path).write_text(json.dumps(out, indent=1) + "\n")


def main():
    rep = sys.argv[1]
    labels = sys.argv[2].split(",") if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else []
    tj = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--traffic-json=")]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    if tj:
        dims = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--shape=")]
        b, sz = (int(v) for v in dims[0].split("x")) if dims else (256, 128)
        traffic_json(hdr, units, data, tj[0], b, sz)
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
