"""Condenses `ncu -i x.ncu-rep --page raw --csv` (exported on the GPU box; the .ncu-rep of a --set full capture is
too large to bring back) into the per-launch numbers the design notes cite.
Usage: python tools/ncu_raw_summary.py profiles/x_raw.csv.gz [label ...] > profiles/x.txt
Labels name the launches in order (e.g. the op names given to tools/run_ops.py)."""
import csv
import gzip
import sys

METRICS = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_%"),
    ("sm__inst_executed_pipe_tensor.sum", "tensor_inst"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_%"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("sm__cycles_elapsed.avg", "sm_cycles"),
]


def main():
    path, labels = sys.argv[1], sys.argv[2:]
    rows = list(csv.reader(gzip.open(path, "rt") if path.endswith(".gz") else open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    hdr, units, data = rows[hi], rows[hi + 1], rows[hi + 2:]
    col = {h: i for i, h in enumerate(hdr)}
    have = [(m, s) for m, s in METRICS if m in col]
    print(f"# {path}: {len(data)} launches; columns: " + ", ".join(f"{s} [{units[col[m]]}]" for m, s in have))
    for k, r in enumerate(data):
        if len(r) < len(hdr):
            continue
        name = r[col["Kernel Name"]].split("(")[0]
        label = labels[k] if k < len(labels) else ""
        vals = []
        for m, s in have:
            v = r[col[m]].replace(",", "")
            try:
                f = float(v)
                vals.append(f"{s}={f:.4g}")
            except ValueError:
                vals.append(f"{s}={v}")
        print(f"{k:3d} {name:24s} {label:18s} " + " ".join(vals))


if __name__ == "__main__":
    main()
