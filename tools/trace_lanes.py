"""Timeline of one enhance() pass as the graph runs it (uv_trace_ops: both utterance lanes on their own streams,
a CUDA event pair around every launch).  Prints what each lane does while the other lane's GRU recurrence runs
and how busy the device is, and optionally dumps the raw (lane, name, start, end) rows.
Usage: python tools/trace_lanes.py [--workload cfg2] [--B 64] [--dump out.tsv]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from open_universe_b200 import _abi, config, networks, program, testing  # noqa: E402,F401


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--B", type=int, default=None)
    ap.add_argument("--dump", default=None)
    args = ap.parse_args()
    wl = dict(bench.WORKLOADS[args.workload])
    if args.B:
        wl["B"] = args.B
    cfg = bench.load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    model = model.to("cuda").eval()
    B, T, n_steps = wl["B"], wl["T"], wl["n_steps"]
    plan = model._get_plan(B, T, n_steps)
    pg = plan.pg
    mix = testing.synthetic_noisy_speech(B, T).cuda()
    noise = torch.randn(n_steps, B, pg.Lp, device="cuda")
    out = torch.empty(B, T, device="cuda")
    coefs = program.step_coefs(model.model_config(), n_steps, 1.3)
    for _ in range(3):
        t0, t1 = plan.trace_ops(mix.data_ptr(), noise.data_ptr(), out.data_ptr(), None, coefs, False)
    rows = [(op.lane, name, a, b, op.type) for op, name, a, b in zip(pg.ops, pg.names, t0, t1) if op.step != -2]
    total = max(r[3] for r in rows)
    print(f"{len(pg.lanes)} lane(s), {len(rows)} launches, pass = {total:.2f} ms (eager, two streams, events around every launch)")
    for lane in range(len(pg.lanes)):
        lr = [r for r in rows if r[0] == lane]
        busy = sum(r[3] - r[2] for r in lr)
        print(f"  lane {lane}: first start {min(r[2] for r in lr):.2f} ms, last end {max(r[3] for r in lr):.2f} ms, "
              f"sum of op intervals {busy:.2f} ms")
    # for every GRU recurrence: how much of its interval is covered by kernels of the other lane
    grus = [r for r in rows if r[4] == _abi.OP_GRU]
    cov_tot = gru_tot = 0.0
    for g in grus:
        other = [r for r in rows if r[0] != g[0]]
        cov = sum(max(0.0, min(g[3], r[3]) - max(g[2], r[2])) for r in other if r[4] != _abi.OP_GRU)
        gcov = sum(max(0.0, min(g[3], r[3]) - max(g[2], r[2])) for r in other if r[4] == _abi.OP_GRU)
        cov_tot += cov
        gru_tot += g[3] - g[2]
        print(f"  {g[1]:14s} lane {g[0]} {g[2]:7.2f} -> {g[3]:7.2f} ms ({g[3]-g[2]:.2f}); other lane: non-GRU kernels {cov:.2f} ms, "
              f"GRU {gcov:.2f} ms")
    if grus:
        print(f"GRU time {gru_tot:.2f} ms, of which overlapped by the other lane's non-GRU kernels {cov_tot:.2f} ms")
    if args.dump:
        with open(args.dump, "w") as f:
            for r in rows:
                f.write(f"{r[0]}\t{r[1]}\t{r[2]:.4f}\t{r[3]:.4f}\n")


if __name__ == "__main__":
    main()
