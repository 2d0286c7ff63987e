"""Condense `ncu -i X.ncu-rep --page raw --csv` into the per-kernel table committed under profiles/:
python tools/ncu_summary.py raw.csv > profiles/ncu_full_rNN.csv   (one column per captured kernel, first launch each)."""
import csv, sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__grid_size", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
kn = hdr.index("Kernel Name")
seen, cols = set(), []
for r in rows[2:]:
    name = r[kn].replace("hexgnn::", "")
    if name in seen:
        continue
    seen.add(name)
    cols.append((name, r))
w = csv.writer(sys.stdout)
w.writerow(["metric", "unit"] + [c[0] for c in cols])
for m in WANT:
    if m in hdr:
        i = hdr.index(m)
        w.writerow([m, units[i]] + [c[1][i] for c in cols])
