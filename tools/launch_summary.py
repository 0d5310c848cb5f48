"""Condenses an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` launch
list of `bench.py` into per-kernel totals (share of the step, DRAM traffic) and, with --json, writes the
per-launch DRAM traffic of the dominant kernel for bench.py's roofline.traffic field.
Usage: python tools/launch_summary.py gpurun_out/launches.csv [--json profiles/roofline_traffic.json]"""
import collections
import csv
import gzip
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    path = sys.argv[1]
    rows = list(csv.reader(gzip.open(path, "rt") if path.endswith(".gz") else open(path)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr, data = rows[hi], rows[hi + 1:]
    ci = {h: i for i, h in enumerate(hdr)}
    agg = collections.defaultdict(lambda: collections.defaultdict(float))
    cnt = collections.Counter()
    for r in data:
        if len(r) < len(hdr):
            continue
        name = r[ci["Kernel Name"]].split("(")[0].split("<")[0].replace("void ", "")
        m, u = r[ci["Metric Name"]], r[ci["Metric Unit"]]
        v = float(r[ci["Metric Value"]].replace(",", ""))
        if m == "gpu__time_duration.sum":
            v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}[u]
            agg[name]["ns"] += v
            cnt[name] += 1
        else:
            agg[name][m] += v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    tot = sum(a["ns"] for a in agg.values())
    print(f"# {path}: {sum(cnt.values())} launches, {tot / 1e6:.1f} ms of kernel time (cold-cache, serialised under ncu)")
    top = None
    for n, a in sorted(agg.items(), key=lambda x: -x[1]["ns"]):
        if top is None:
            top = n
        rd, wr = a["dram__bytes_read.sum"], a["dram__bytes_write.sum"]
        print(f"{n:60s} launches={cnt[n]:5d} time={a['ns'] / 1e6:9.3f} ms share={a['ns'] / tot * 100:5.1f}%  "
              f"dram_rd={rd / 1e9:8.2f} GB dram_wr={wr / 1e9:8.2f} GB  per launch {(rd + wr) / cnt[n] / 1e6:8.1f} MB")
    if "--json" in sys.argv:
        out = sys.argv[sys.argv.index("--json") + 1]
        a = agg[top]
        import bench   # the build the pass was taken on: bench.py quotes the figure only for the same sources

        json.dump({"kernel": top.split("::")[-1], "launches": cnt[top], "build": bench.source_hash(),
                   "traffic_bytes_per_launch": (a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]) / cnt[top],
                   "share_of_kernel_time_under_ncu": a["ns"] / tot, "source": path}, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
