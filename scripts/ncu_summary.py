"""Condenses `ncu -i X.ncu-rep --page raw --csv` into the handful of columns profiles/ keeps.
   ncu -i gpurun_out/X.ncu-rep --page raw --csv | python scripts/ncu_summary.py > profiles/rNN_X.csv"""
import csv, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
rows = [r for r in csv.reader(l for l in sys.stdin if not l.startswith("=="))]
hdr, units = rows[0], rows[1]
col = {}
for w in WANT:
    hits = [i for i, h in enumerate(hdr) if h == w or h.endswith("." + w)]
    if hits: col[w] = hits[0]
name = hdr.index("Kernel Name")
out = csv.writer(sys.stdout)
out.writerow(["kernel"] + ["%s [%s]" % (w, units[col[w]]) for w in WANT if w in col])
for r in rows[2:]:
    out.writerow([r[name][:90]] + [r[col[w]] for w in WANT if w in col])
