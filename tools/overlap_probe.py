"""Probe: do consecutive enhance() passes overlap usefully when they alternate between two plans on two streams
(the tail of pass k, where one utterance lane runs alone, under the head of pass k + 1)?  Prints ms per pass for one
plan on one stream and for two plans on two streams.  Usage: python tools/overlap_probe.py [--workload cfg2] [--steps 12]"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from open_universe_b200 import _native, config, networks, program, testing  # noqa: E402,F401


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--plans", type=int, default=2)
    args = ap.parse_args()
    wl = bench.WORKLOADS[args.workload]
    cfg = bench.load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    model = model.to("cuda").eval()
    B, T, n_steps = wl["B"], wl["T"], wl["n_steps"]
    plans = [model._get_plan(B, T, n_steps)]
    for _ in range(args.plans - 1):
        plans.append(_native.NativePlan(model._native_model, plans[0].pg, use_graph=1))
    Lp = plans[0].pg.Lp
    mix = testing.synthetic_noisy_speech(B, T).cuda()
    noise = torch.randn(n_steps, B, Lp, device="cuda")
    outs = [torch.empty(B, T, device="cuda") for _ in plans]
    coefs = program.step_coefs(model.model_config(), n_steps, 1.3)
    streams = [torch.cuda.Stream() for _ in plans]
    main_s = torch.cuda.current_stream()

    def run(k, n_plans):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record(main_s)
        for s in streams[:n_plans]:
            s.wait_event(e0)
        for i in range(k):
            j = i % n_plans
            plans[j].enhance(mix.data_ptr(), noise.data_ptr(), outs[j].data_ptr(), None, coefs, False, streams[j].cuda_stream)
        for s in streams[:n_plans]:
            ev = torch.cuda.Event()
            ev.record(s)
            main_s.wait_event(ev)
        e1.record(main_s)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k

    for n_plans in range(1, args.plans + 1):
        run(4, n_plans)
        ms = [run(args.steps, n_plans) for _ in range(3)]
        print(f"{n_plans} plan(s) on {n_plans} stream(s): {min(ms):.2f} ms per pass (runs {['%.2f' % m for m in ms]}) -> {B * T / 16000 / min(ms) * 1e3:.0f} audio-s/s")
    ref = outs[0].clone()
    print("outputs of the plans agree:", all(torch.equal(ref, o) for o in outs))


if __name__ == "__main__":
    main()
