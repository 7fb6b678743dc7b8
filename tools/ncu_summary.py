"""Table of the per-launch counters the DESIGN quotes from an `ncu --set full` report:
    ncu -i report.ncu-rep --page raw --csv > raw.csv;  python tools/ncu_summary.py raw.csv [layer names...]
tensor% = sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32 (pct of peak, elapsed), tc smem% = l1tex__data_pipe_tc_wavefronts_mem_shared
(tensor-core operand reads from shared memory, pct of peak, elapsed).  (measurement aid)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
TO_MB = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6}
idx = {h: i for i, h in enumerate(hdr)}
def col(suffix):
    return idx[suffix] if suffix in idx else idx[[h for h in hdr if h.endswith(suffix)][0]]
cols = [("us", "gpu__time_duration.sum"), ("tensor%", "sm__ops_path_tensor_op_utchmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"),
        ("tc smem%", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
        ("L2 thr%", "lts__throughput.avg.pct_of_peak_sustained_elapsed"), ("dram rd MB", "dram__bytes_read.sum"),
        ("dram wr MB", "dram__bytes_write.sum"), ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        ("regs", "launch__registers_per_thread")]
names = sys.argv[2:]
print(f"{'layer':22s} {'kernel':40s} " + " ".join(f"{n:>10s}" for n, _ in cols))
for i, r in enumerate(rows[2:]):
    k = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
    vals = []
    for _, c in cols:
        try:
            v = float(r[col(c)].replace(',', ''))
            if units[col(c)] in TO_MB:
                v *= TO_MB[units[col(c)]]
            vals.append(f"{v:10.1f}")
        except Exception:
            vals.append(f"{'-':>10s}")
    print(f"{(names[i] if i < len(names) else ''):22s} {k:40s} " + " ".join(vals))
