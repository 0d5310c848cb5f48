"""Condenses an ncu report (.ncu-rep, read here on the CPU box) into the few per-launch numbers the
design notes cite: duration, DRAM bytes, tensor-pipe %, DRAM %, L2 hit rate, registers, grid.
Usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep [label ...] > profiles/x.txt"""
import csv
import io
import subprocess
import sys

METRICS = [
    ("gpu__time_duration.sum", "dur"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_%"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%"),
    ("lts__t_sector_hit_rate.pct", "l2_hit_%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_%"),
    ("smsp__inst_executed.sum", "warp_inst"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__cycles_elapsed.avg", "sm_cycles"),
]


def main():
    rep = sys.argv[1]
    labels = sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: {len(data)} launches (ncu --set full --clock-control none; per-launch, cold-ish cache, serialised)")
    for k, r in enumerate(data):
        name = r[col["Kernel Name"]].split("(")[0]
        lab = labels[k] if k < len(labels) else ""
        parts = [f"{k}: {name} {lab}"]
        for m, short in METRICS:
            if m in col:
                parts.append(f"{short}={r[col[m]]}{units[col[m]]}")
        print("  ".join(parts))


if __name__ == "__main__":
    main()
