"""Runs a few named ops of a compiled enhance() program in isolation (uv_run_ops), for ncu captures
and quick timings.  Usage: python tools/run_ops.py --ops s0.enc0.conv2,s0.enc1.conv2 [--B 32] [--reps 3]
The whole program is run once first so the selected ops see realistic operand data."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from open_universe_b200 import config, networks, program, testing  # noqa: E402,F401


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--ops", required=True)
    ap.add_argument("--B", type=int, default=None)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--no-warm", action="store_true")
    args = ap.parse_args()
    wl = dict(bench.WORKLOADS[args.workload])
    if args.B:
        wl["B"] = args.B
    cfg = bench.load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    model = model.to("cuda").eval()
    model.use_cuda_graph = False
    B, T, n_steps = wl["B"], wl["T"], wl["n_steps"]
    plan = model._get_plan(B, T, n_steps)
    pg = plan.pg
    mix = testing.synthetic_noisy_speech(B, T).cuda()
    noise = torch.randn(n_steps, B, pg.Lp, device="cuda")
    out = torch.empty(B, T, device="cuda")
    coefs = program.step_coefs(model.model_config(), n_steps, 1.3)
    st = torch.cuda.current_stream().cuda_stream
    io = (mix.data_ptr(), noise.data_ptr(), out.data_ptr(), None, coefs, False, st)
    idx = {}
    for i, n in enumerate(pg.names):
        idx.setdefault(n, i)  # score-net ops repeat per step: take step 0
    sel = [idx[n] for n in args.ops.split(",")]
    if not args.no_warm:
        last = max(sel) + 1
        first = 0
        plan.run_ops(first, last, *io)
    for _ in range(args.reps):
        for i in sel:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            plan.run_ops(i, i + 1, *io)
            ev1.record()
            torch.cuda.synchronize()
            print(f"{pg.names[i]:20s} {ev0.elapsed_time(ev1)*1e3:9.1f} us (incl. launch + sync overhead)")


if __name__ == "__main__":
    main()
