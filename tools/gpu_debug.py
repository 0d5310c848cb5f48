"""GPU bring-up tool: runs a compiled program op by op on the GPU (uv_run_ops) and on the CPU
emulator, and reports the relative error of every op's outputs -- localises a kernel bug to
one launch.  Usage: python tools/gpu_debug.py [case] [--tol 0.02]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from open_universe_b200 import _abi, _native, config, networks, program, testing, weights  # noqa: E402,F401
from oracle.emulator import Emulator, op_outputs  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="universepp_16k")
    ap.add_argument("--B", type=int, default=2)
    ap.add_argument("--T", type=int, default=3000)
    ap.add_argument("--tol", type=float, default=0.02)
    ap.add_argument("--lockstep", action="store_true", help="copy the emulator's arena to the GPU before every op")
    args = ap.parse_args()
    with open(os.path.join(ROOT, "tests", "golden", f"config_{args.model}.json")) as f:
        cfg = config.to_attr(json.load(f))["model"]
    model = config.instantiate(cfg)
    testing.seed_weights(model, seed=7)
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    mcfg = model.model_config()
    pm = weights.pack_model(sd, mcfg)
    n_steps = int(cfg.diffusion.n_steps)
    pg = program.build_program(pm, mcfg, args.B, args.T, n_steps)
    print(f"{len(pg.ops)} ops, arena {pg.arena/1e6:.1f} MB, blob {pm.blob.size/1e6:.1f} MB")
    g = torch.Generator().manual_seed(1235)
    mix = 0.1 * torch.randn(args.B, args.T, generator=g)
    noise = torch.stack([torch.randn(args.B, pg.Lp, generator=g) for _ in range(n_steps)])
    coefs = program.step_coefs(mcfg, n_steps, float(cfg.diffusion.epsilon))

    nm = _native.NativeModel(pm.blob.bytes(), 0)
    plan = _native.NativePlan(nm, pg, use_graph=False)
    dmix, dnoise = mix.cuda(), noise.cuda()
    dout = torch.zeros(args.B, args.T, device="cuda")
    ddbg = torch.zeros(n_steps, args.B, pg.Lp, device="cuda")
    em = Emulator(pm, pg)
    io = dict(mix=mix, noise=noise, coef=coefs, keep_rms=False, dbg={})
    bad = 0
    worst = {}
    for i, (op, name) in enumerate(zip(pg.ops, pg.names)):
        t0 = time.time()
        em.run_op(i, io)
        plan.run_ops(i, i + 1, dmix.data_ptr(), dnoise.data_ptr(), dout.data_ptr(), ddbg.data_ptr(), coefs, False, None)
        errs = []
        for label, off, nbytes, dt in op_outputs(op):
            gbytes = plan.arena_read(off, nbytes)
            gt = torch.frombuffer(bytearray(gbytes), dtype=dt).float()
            et = em.arena[off : off + nbytes].view(dt).float()
            den = et.norm().item()
            err = (gt - et).norm().item() / max(den, 1e-12)
            nan = not torch.isfinite(gt).all().item()
            errs.append((label, err, nan, den))
        if op.type == _abi.OP_POST:
            err = (dout.cpu() - io["out"]).norm().item() / io["out"].norm().item()
            errs.append(("out", err, False, io["out"].norm().item()))
        mx = max([e[1] for e in errs] + [0.0])
        kind = _abi.OP_NAMES[op.type]
        worst[kind] = max(worst.get(kind, 0.0), mx)
        flag = mx > args.tol or any(e[2] for e in errs)
        if flag or i < 12 or i % 40 == 0:
            print(f"op {i:4d} {kind:8s} {name:24s} " + " ".join(f"{l}:{e:.2e}{'(NaN)' if n else ''}[{d:.2g}]" for l, e, n, d in errs)
                  + ("   <-- MISMATCH" if flag else ""))
        if flag:
            bad += 1
            if bad > 25:
                print("too many mismatches, stopping")
                break
    print("worst per op kind:", {k: f"{v:.2e}" for k, v in worst.items()})
    if "out" in io:
        print("final out rel err vs emulator:", ((dout.cpu() - io["out"]).norm() / io["out"].norm()).item())
    print("MISMATCHING OPS:", bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
