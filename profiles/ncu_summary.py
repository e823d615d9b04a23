"""Condense `ncu -i X.ncu-rep --page raw --csv` into the handful of per-launch metrics DESIGN.md / bench.py cite.

  ncu -i gpurun_out/r01_full.ncu-rep --page raw --csv > /tmp/raw.csv ; python profiles/ncu_summary.py /tmp/raw.csv > profiles/rNN_ncu_full.md
"""
import csv
import sys

WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("lts__t_bytes.sum", "l2_bytes"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_%peak"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%peak"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pipe_%"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_%"),
        ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("smsp__inst_executed.sum", "warp_inst"), ("l1tex__t_sector_hit_rate.pct", "l1_hit_%"),
        ("lts__t_sector_hit_rate.pct", "l2_hit_%")]


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [(lab, hdr.index(m)) for m, lab in WANT if m in hdr]
    kn = hdr.index("Kernel Name")
    print("| kernel | " + " | ".join(lab for lab, _ in cols) + " |")
    print("|---|" + "---|" * len(cols))
    for r in data:
        cells = []
        for lab, i in cols:
            v = r[i]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}" if abs(f) < 1e6 else f"{f:.4e}"
            except ValueError:
                pass
            cells.append(f"{v} {units[i]}".strip())
        print(f"| `{r[kn].split('(')[0]}` | " + " | ".join(cells) + " |")


def traffic(path):
    """{kernel: dram bytes (read + write) per launch, averaged over the captured launches} as JSON (bench.py's roofline.traffic)."""
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    acc = {}
    for r in data:
        name = r[kn].split("(")[0].replace("void ", "").split("<")[0]
        b = float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]]
        acc.setdefault(name, []).append(b)
    print(json.dumps({k: sum(v) / len(v) for k, v in acc.items()}, indent=1))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[2] == "--traffic":
        traffic(sys.argv[1])
    else:
        main(sys.argv[1])
