"""Per-launch table (duration, DRAM bytes, tensor-pipe activity, registers) from `ncu -i X.ncu-rep --page raw --csv`:
ncu -i gpurun_out/rNN.ncu-rep --page raw --csv | python tools/ncu_table.py > profiles/rNN_summary.txt"""
import csv, sys

rows = list(csv.reader(l for l in sys.stdin if l.startswith('"')))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [("Kernel Name", "kernel", 34), ("Grid Size", "grid", 14), ("gpu__time_duration.sum", "time", 10),
        ("dram__bytes_read.sum", "dram rd", 10), ("dram__bytes_write.sum", "dram wr", 10),
        ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", "tensor%", 9),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm thr%", 9),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram thr%", 10),
        ("sm__cycles_elapsed.max", "cycles", 11), ("launch__registers_per_thread", "regs", 6),
        ("launch__shared_mem_per_block_dynamic", "dyn smem", 10)]
cols = [c for c in cols if c[0] in hdr]
idx = [hdr.index(c[0]) for c in cols]
print("  ".join(f"{c[1] + (' [' + units[i] + ']' if units[i] and len(units[i]) < 8 else ''):>{c[2]}}" for c, i in zip(cols, idx)))
for r in data:
    out = []
    for c, i in zip(cols, idx):
        v = r[i]
        if c[1] == "kernel":
            v = v.replace("void ", "").replace("unnamed>::", "").split("(")[0]
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        out.append(f"{v[:c[2]]:>{c[2]}}")
    print("  ".join(out))
