"""Reads `ncu -i X.ncu-rep --page raw --csv` of ONE likelihood_field_kernel launch and
records the counters bench.py quotes (roofline.ncu_recorded, roofline.traffic) in
profiles/lf_counters.json, keyed by particles per launch.

    python profiles/lf_counters_from_ncu.py gpurun_out/r2_lf_c_raw.csv 16777216 <commit>
"""
import csv
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "lf_counters.json")


def main():
    raw, particles, commit = sys.argv[1], sys.argv[2], sys.argv[3]
    rows = list(csv.reader(open(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = dict(zip(hdr, vals))
    u = dict(zip(hdr, units))

    def num(k):
        return float(d[k].replace(",", ""))

    def nbytes(k):
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u[k]]
        return num(k) * scale

    rec = {
        "commit": commit,
        "kernel": d.get("Kernel Name", "likelihood_field_kernel"),
        "kernel_ms_under_ncu": num("gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}[u["gpu__time_duration.sum"]],
        "l2_sectors_read": int(num("lts__t_sectors_srcunit_tex_op_read.sum")),
        # map cells arrive as global loads (round-2 kernel up to 3b66385) or texture fetches (after)
        "l1tex_sectors": int(num("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum")
                             + num("l1tex__t_sectors_pipe_tex_mem_texture.sum")),
        "l1tex_sectors_texture": int(num("l1tex__t_sectors_pipe_tex_mem_texture.sum")),
        "l1tex_tex_pipe_pct": num("l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "dram_bytes": int(nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")),
        "l1tex_lsu_pipe_pct": num("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
        "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "fp64_pipe_pct": num("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        "l2_hit_pct": num("lts__t_sector_hit_rate.pct"),
        "inst_executed": int(num("smsp__inst_executed.sum")),
        "shared_wavefronts": int(num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")),
        "registers": int(num("launch__registers_per_thread")),
        "grid": int(num("launch__grid_size")), "block": int(num("launch__block_size")),
    }
    allrec = json.load(open(OUT)) if os.path.exists(OUT) else {
        "note": "ncu --set full --clock-control none of ONE likelihood_field_kernel<LF> launch of bench.py "
                "(1059 used beams, 4000x4000 map), keyed by particles per launch; bench.py quotes these as "
                "'ncu, recorded at <commit>' -- they are not measured in the timed run"}
    allrec[str(particles)] = rec
    json.dump(allrec, open(OUT, "w"), indent=1)
    print(json.dumps(rec, indent=1))


if __name__ == "__main__":
    main()
