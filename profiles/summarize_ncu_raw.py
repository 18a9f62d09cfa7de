"""Per-kernel table from `ncu -i X.ncu-rep --page raw --csv`: the counters that show
what bounds each kernel (duration, DRAM bytes and throughput, L1TEX / L2 / issue /
fp64 utilisation).

    python profiles/summarize_ncu_raw.py gpurun_out/r2_stages_raw.csv [hbm_gbs]
"""
import csv
import re
import sys

COLS = [
    ("ms", "gpu__time_duration.sum"),
    ("DRAM rd MB", "dram__bytes_read.sum"),
    ("DRAM wr MB", "dram__bytes_write.sum"),
    ("DRAM %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("L1TEX %", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
    ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
    ("fp64 %", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"),
    ("warps %", "sm__warps_active.avg.pct_of_peak_sustained_active"),
    ("regs", "launch__registers_per_thread"),
    ("grid", "launch__grid_size"),
]
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}


def main(path, hbm=None):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    print("| kernel | " + " | ".join(c for c, _ in COLS) + (" | GB/s | of HBM peak |" if hbm else " |"))
    print("|---|" + "---:|" * (len(COLS) + (2 if hbm else 0)))
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("amclb200::", "").replace("(anonymous namespace)::", "")
        vals, by = [], {}
        for label, key in COLS:
            if key not in ix or r[ix[key]] == "":
                vals.append("-")
                continue
            v = float(r[ix[key]].replace(",", "")) * SCALE.get(units[ix[key]], 1.0)
            by[label] = v
            vals.append("%.3f" % v if label == "ms" else ("%.1f" % v if v < 1000 else "%.0f" % v))
        extra = ""
        if hbm and "ms" in by:
            gbs = (by.get("DRAM rd MB", 0) + by.get("DRAM wr MB", 0)) / by["ms"]
            extra = " | %.0f | %.2f |" % (gbs, gbs / hbm)
        print("| `%s` | " % name.strip() + " | ".join(vals) + (extra if hbm else " |"))


if __name__ == "__main__":
    main(sys.argv[1], float(sys.argv[2]) if len(sys.argv) > 2 else None)
