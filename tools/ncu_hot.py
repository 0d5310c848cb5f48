"""Lists the SASS instructions of one launch in an ncu report that collect the most warp-stall samples.
Usage: python tools/ncu_hot.py report.ncu-rep [launch_index] [min_share]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.004
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    # the export holds one block per launch: a "Kernel Name" row, a header row, then instructions
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    starts.append(len(rows))
    blk = rows[starts[k]:starts[k + 1]]
    hdr, data = blk[1], [r for r in blk[2:] if len(r) == len(blk[1])]
    ci = {h: i for i, h in enumerate(hdr)}
    ie, ss = ci["Instructions Executed"], ci["Warp Stall Sampling (All Samples)"]
    tot = sum(int(r[ie]) for r in data)
    tots = sum(int(r[ss]) for r in data)
    print(f"# launch {k}: {blk[0][1]}  warp instructions {tot}  stall samples {tots}")
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for i, r in enumerate(data):
        if int(r[ss]) > tots * share:
            st = {h[6:]: int(r[ci[h]]) for h in stalls if int(r[ci[h]]) > 0}
            top = sorted(st.items(), key=lambda x: -x[1])[:3]
            print(f"{i:6d} {r[1].strip()[:72]:72s} exec={r[ie]:>9s} samples={r[ss]:>6s} {top}")


if __name__ == "__main__":
    main()
