"""The fp32 parity tier (north_star: rel-L2 <= 1e-4 "in tf32 / fp32", SI-SDR >= 40 dB; VERDICT row N2).

model.precision = "fp32" selects a program whose activations are f32 planes and whose dense layers / recurrences run
as fp32 FMA kernels (csrc/hp_ops.cuh) from fp32 weight images; everything goes through the same C ABI (uv_plan_create /
uv_enhance / uv_run_ops).  Checked against the golden outputs of the real reference, against the CPU oracle at a
benchmark length and with two utterance lanes, and launch by launch against the CPU emulator of the op program."""
import pytest
import torch

from conftest import GOLDEN_CASES, build_model, golden_inputs, load_fixture
from open_universe_b200 import _abi, _native, program, weights
from oracle import universe_oracle as uo
from oracle.emulator import Emulator, op_outputs

pytestmark = pytest.mark.gpu

TOL = 1e-4          # north_star: per-step score outputs and final waveform, fp32 tier
SISDR_MIN = 40.0


def _gpu_model(name, seed=7):
    model, cfg = build_model(name, seed)
    model = model.to("cuda")
    model.eval()
    model.precision = "fp32"
    return model, cfg


def test_fp32_every_op_matches_emulator():
    """Every launch of a 3 x 0.3 s fp32 program against the CPU emulator (same layouts, fp32 arithmetic)."""
    B, T, n_steps = 3, 5001, 8
    model, cfg = build_model("universepp_16k")
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    mcfg = model.model_config()
    pm = weights.pack_model(sd, mcfg, hp=True)
    pg = program.build_program(pm, mcfg, B, T, n_steps, precision="fp32")
    nm = _native.NativeModel(pm.blob.bytes(), 0)
    plan = _native.NativePlan(nm, pg, use_graph=False)
    coefs = program.step_coefs(mcfg, n_steps, float(cfg.model.diffusion.epsilon))
    g = torch.Generator().manual_seed(1235)
    mix = 0.1 * torch.randn(B, T, generator=g)
    noise = torch.stack([torch.randn(B, pg.Lp, generator=g) for _ in range(n_steps)])
    dmix, dnoise = mix.cuda(), noise.cuda()
    dout, ddbg = torch.zeros(B, T, device="cuda"), torch.zeros(n_steps, B, pg.Lp, device="cuda")
    em = Emulator(pm, pg)
    eio = dict(mix=mix, noise=noise, coef=coefs, keep_rms=False, dbg={})
    worst, bad = {}, []
    for i, (op, name) in enumerate(zip(pg.ops, pg.names)):
        em.run_op(i, eio)
        plan.run_ops(i, i + 1, dmix.data_ptr(), dnoise.data_ptr(), dout.data_ptr(), ddbg.data_ptr(), coefs, False, None)
        for label, off, nbytes, dt in op_outputs(op):
            gt = torch.frombuffer(bytearray(plan.arena_read(off, nbytes)), dtype=dt).float()
            et = em.arena[off : off + nbytes].view(dt).float()
            err = (gt - et).norm().item() / max(et.norm().item(), 1e-12)
            kind = _abi.OP_NAMES[op.type]
            worst[kind] = max(worst.get(kind, 0.0), err)
            # dense layers: split-bf16 products on the tensor cores (16 mantissa bits per operand), fp32 everywhere else
            if not (err < (5e-5 if op.type == _abi.OP_GEMM else 2e-5)) or not torch.isfinite(gt).all():
                bad.append((i, name, label, err))
    e_out = uo.rel_l2(dout.cpu(), eio["out"])
    print("fp32 tier: worst rel-L2 per op kind vs emulator:", {k: f"{v:.1e}" for k, v in worst.items()}, f"final out {e_out:.2e}")
    assert not bad, bad[:10]
    assert e_out < TOL


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_fp32_enhance_matches_reference_golden(case):
    fix = load_fixture(case)
    model, cfg = _gpu_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    out, steps = model.enhance_debug(mix.cuda(), noise.cuda(), **fix["kwargs"])
    out, steps = out.cpu(), steps.cpu()
    assert torch.isfinite(out).all()
    errs = [uo.rel_l2(steps[i], fix["steps"][i]) for i in range(steps.shape[0])]
    e_out, sisdr = uo.rel_l2(out, fix["out"]), uo.si_sdr(out, fix["out"]).min().item()
    print(f"fp32 tier vs reference golden {case}: worst per-step rel-L2 {max(errs):.2e}, final {e_out:.2e}, SI-SDR {sisdr:.1f} dB")
    assert max(errs) < TOL
    assert e_out < TOL
    assert sisdr > SISDR_MIN


def test_fp32_enhance_matches_reference_golden_4s():
    """A benchmark-length clip (T = 64 000: 401 GRU steps) recorded from the real reference (oracle/gen_golden.py)."""
    fix = load_fixture("pp16k_4s")
    model, cfg = _gpu_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    out, steps = model.enhance_debug(mix.cuda(), noise.cuda(), **fix["kwargs"])
    out, steps = out.cpu(), steps.cpu()
    r = fix["keep_row"]   # the fixture keeps one row of three steps
    errs = [uo.rel_l2(steps[n, r : r + 1], fix["steps"][k]) for k, n in enumerate(fix["keep_steps"])]
    e_out, sisdr = uo.rel_l2(out, fix["out"]), uo.si_sdr(out, fix["out"]).min().item()
    print(f"fp32 tier vs reference golden 4 s clip: worst per-step rel-L2 {max(errs):.2e}, final {e_out:.2e}, SI-SDR {sisdr:.1f} dB")
    assert max(errs) < TOL and e_out < TOL and sisdr > SISDR_MIN


def test_fp32_two_lanes_graph_and_lengths_match_oracle():
    """17 utterances (two lanes on two captured streams), replayed as a CUDA graph, against the CPU oracle; then the
    same plan with per-row lengths: every row equals a single-utterance run of that length."""
    model, cfg = _gpu_model("universepp_16k")
    B, T = 17, 2400
    g = torch.Generator().manual_seed(3)
    mix = 0.05 * torch.randn(B, T, generator=g)
    Lp = T + (model.tot_ds - T % model.tot_ds)
    noise = torch.randn(8, B, Lp, generator=g)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref, ref_steps = uo.enhance(sd, cfg.model, mix, [n[:, None, :] for n in noise], return_steps=True)
    for rep in range(3):   # eager first call, captured graph from the second on
        out, steps = model.enhance_debug(mix.cuda(), noise.cuda())
        e_steps = max(uo.rel_l2(steps[i].cpu(), ref_steps[i][:, 0]) for i in range(8))
        e_out = uo.rel_l2(out.cpu(), ref)
        assert e_steps < TOL and e_out < TOL, (rep, e_steps, e_out)
    plan = next(iter(model._plans.values()))
    assert len(plan.pg.lanes) == 2 and all(op.type != _abi.OP_CONVBLOCK for op in plan.pg.ops)
    print(f"fp32 tier, 17 x {T} (two lanes, graph): worst per-step rel-L2 {e_steps:.2e}, final {e_out:.2e}")
    lengths = [T - 97 * b for b in range(B)]
    out_r, _ = model.enhance_debug(mix.cuda(), noise.cuda(), lengths=lengths)
    for b in (0, 5, 16):
        Tb = lengths[b]
        Lpb = Tb + (model.tot_ds - Tb % model.tot_ds)
        one, _ = model.enhance_debug(mix[b : b + 1, :Tb].cuda(), noise[:, b : b + 1, :Lpb].contiguous().cuda())
        err = uo.rel_l2(out_r[b, :Tb].cpu(), one[0].cpu())
        assert err < 1e-5, (b, err)


def test_precision_switch_keeps_both_tiers():
    """One model object serves both tiers: switching `precision` repacks the weights once and keys the plans."""
    model, cfg = _gpu_model("universepp_16k")
    mix = 0.05 * torch.randn(2, 1600, generator=torch.Generator().manual_seed(1)).cuda()
    noise = torch.randn(8, 2, 1760, generator=torch.Generator().manual_seed(2)).cuda()
    hi, _ = model.enhance_debug(mix, noise)
    model.precision = "bf16"
    lo, _ = model.enhance_debug(mix, noise)
    assert len(model._plans) == 2
    err = uo.rel_l2(lo.cpu(), hi.cpu())
    print(f"bf16 tier against fp32 tier on the same draws: rel-L2 {err:.2e}")
    assert 1e-6 < err < 1e-2
    model.precision = "tf64"
    with pytest.raises(ValueError):
        model.enhance_debug(mix, noise)


def test_fp32_config_length_24k_matches_cpu_oracle():
    """BASELINE configs[3] length at reduced batch: UNIVERSE++ 24 kHz, 1 x 10 s = 1001 GRU steps of the H = 384
    recurrence, 48-channel levels (16-channel K groups of the split GEMM)."""
    model, cfg = _gpu_model("universepp_24k")
    B, T = 1, 240000
    g = torch.Generator().manual_seed(T + B)
    mix = 0.05 * torch.randn(B, T, generator=g) + 0.003
    Lp = T + (model.tot_ds - T % model.tot_ds)
    noise = torch.randn(8, B, Lp, generator=g)
    out, steps = model.enhance_debug(mix.cuda(), noise.cuda())
    out, steps = out.cpu(), steps.cpu()
    model._invalidate()
    torch.cuda.empty_cache()
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref, ref_steps = uo.enhance(sd, cfg.model, mix, [n[:, None, :] for n in noise], return_steps=True)
    e_steps = max(uo.rel_l2(steps[i], ref_steps[i][:, 0]) for i in range(8))
    e_out, sisdr = uo.rel_l2(out, ref), uo.si_sdr(out, ref).min().item()
    print(f"fp32 tier, 24 kHz 1 x 10 s (1001 GRU steps): worst per-step rel-L2 {e_steps:.2e}, final {e_out:.2e}, SI-SDR {sisdr:.1f} dB")
    assert e_steps < TOL and e_out < TOL and sisdr > SISDR_MIN
