"""Summarise an .ncu-rep (raw page) into the metrics we track.  usage: ncu_summary.py rep [out.csv]"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
pats = ('Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct', 'sm__throughput.avg.pct', 'sm__warps_active.avg.pct',
        'launch__registers_per_thread', 'launch__occupancy_limit', 'launch__grid_size', 'launch__block_size',
        'sm__inst_executed_pipe_fma.avg.pct', 'sm__inst_executed_pipe_alu.avg.pct', 'sm__inst_executed_pipe_fp64.avg.pct',
        'sm__inst_executed_pipe_lsu.avg.pct', 'sm__inst_executed_pipe_xu.avg.pct', 'sm__pipe_tensor',
        'smsp__issue_active.avg.pct', 'smsp__inst_executed.sum', 'lts__t_sector_hit_rate.pct',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared', 'smsp__average_warp', 'smsp__warp_issue_stalled',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'lts__t_bytes.sum', 'sm__cycles_elapsed.max')
keep = [i for i, h in enumerate(hdr) if any(p in h for p in pats)]
out = sys.argv[2] if len(sys.argv) > 2 else None
if out:
    with open(out, "w") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in keep]); w.writerow([units[i] for i in keep])
        for r in rows[2:]:
            w.writerow([r[i] for i in keep])
for r in rows[2:]:
    print("----")
    for i in keep:
        v = r[i]
        if 'warp_issue_stalled' in hdr[i] and 'ratio' not in hdr[i]:
            continue
        print("  %-80s %s %s" % (hdr[i][:80], v[:60], units[i]))
