"""Per-op device times of one enhance() pass (uv_profile_ops, CUDA events), grouped by layer.
Usage: python tools/profile_ops.py [--workload cfg2] [--top 40]"""
import argparse
import os
import re
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from open_universe_b200 import _abi, config, networks, program, testing  # noqa: E402,F401


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2")
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--B", type=int, default=None)
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
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(2):
        ms = plan.profile_ops(mix.data_ptr(), noise.data_ptr(), out.data_ptr(), None, coefs, False, st)
    # aggregate by layer name (steps folded together)
    agg = {}
    for op, name, t in zip(pg.ops, pg.names, ms):
        key = re.sub(r"^s\d+\.", "s.", name)
        a = agg.setdefault(key, [0.0, 0, op])
        a[0] += t
        a[1] += 1
    total = sum(ms)
    print(f"total eager {total:.2f} ms over {len(ms)} ops")
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    for key, (t, n, op) in rows[: args.top]:
        extra = ""
        if op.type == _abi.OP_GEMM:
            K = op.a.nch * 8 * op.n_src * (op.smax - op.smin + 1)
            flop = 2.0 * op.m_rows * op.N_total * K
            nbytes_out = sum(op.out[i].t.nch * 8 * (2 if op.out[i].t.dtype == 1 else 4) for i in range(op.n_out)) * op.m_rows * op.up
            extra = f"M={op.m_rows} N={op.N_total} K={K} nsrc={op.n_src} up={op.up} nout={op.n_out} {flop*n/t/1e9:8.1f} TF/s(exec) out {nbytes_out/1e6:.0f} MB"
        print(f"{key:22s} {_abi.OP_NAMES[op.type]:8s} x{n:2d} {t:8.3f} ms  {t/n*1e3:8.1f} us/launch  {extra}")


if __name__ == "__main__":
    main()
