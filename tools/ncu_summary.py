"""Summarise an .ncu-rep (ncu --set full) into the files profiles/ keeps:

    python tools/ncu_summary.py gpurun_out/r02n/ncu_features_tc2.ncu-rep profiles/r02n_ncu_features_tc2 \
        --alg-bytes 3118579200 --what "4 874 x 8 s lmfe launch"

writes <out>_summary.csv (the metrics DESIGN.md quotes, one row per captured launch) and, when
--alg-bytes is given, <out>_ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per
ALGORITHMIC byte of the launch (bench.py reads the newest such file for roofline.traffic).
"""
import argparse
import csv
import io
import json
import subprocess

KEEP = [
    "gpu__time_duration.sum", "sm__cycles_elapsed.avg", "launch__registers_per_thread", "launch__block_size",
    "launch__grid_size", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum", "lts__t_sector_hit_rate.pct",
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("out")
    ap.add_argument("--alg-bytes", type=float, default=None)
    ap.add_argument("--what", default="")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    cols = [i for i, h in enumerate(hdr) if h in KEEP or h == "Kernel Name" or
            (h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio"))]
    with open(a.out + "_summary.csv", "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow([hdr[i] for i in cols])
        w.writerow([units[i] for i in cols])
        for r in rows[2:]:
            w.writerow([r[i] for i in cols])
    if a.alg_bytes:
        recs = []
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))

            def gb(k):
                v = float(d[k])
                return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[u[k]]
            tot = gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum")
            recs.append({"kernel": d["Kernel Name"], "dram_bytes": tot, "algorithmic_bytes": a.alg_bytes,
                         "dram_bytes_per_alg_byte": tot / a.alg_bytes, "what": a.what, "report": a.rep})
        with open(a.out + "_ncu_traffic.json", "w") as fh:
            json.dump(recs, fh, indent=1)
    print("wrote", a.out + "_summary.csv")


if __name__ == "__main__":
    main()
