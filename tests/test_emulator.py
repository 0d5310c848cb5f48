"""Compiler bookkeeping on the CPU: the op program + packed weights, executed by the CPU
emulator with the kernels' layouts and bf16 rounding points, against the reference's golden
outputs.  Tolerance: rel-L2 <= 1e-2 (bf16 operands, north_star), SI-SDR >= 40 dB."""
import pytest
import torch

from conftest import GOLDEN_CASES, build_model, golden_inputs, load_fixture
from open_universe_b200 import program, weights
from oracle import universe_oracle as uo
from oracle.emulator import Emulator


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_program_emulation_matches_reference(case):
    fix = load_fixture(case)
    model, cfg = build_model(fix["model"], fix["seed_weights"])
    mcfg = model.model_config()
    kw = fix["kwargs"]
    n_steps = kw.get("n_steps", int(cfg.model.diffusion.n_steps))
    pm = weights.pack_model(model.state_dict(), mcfg)
    pg = program.build_program(pm, mcfg, fix["B"], fix["T"], n_steps)
    mix, noise = golden_inputs(fix, model.tot_ds)
    coefs = program.step_coefs(mcfg, n_steps, kw.get("epsilon", float(cfg.model.diffusion.epsilon)))
    io = Emulator(pm, pg).run(mix, noise, coefs, keep_rms=kw.get("keep_rms", False), dbg=True)
    assert uo.rel_l2(io["out"], fix["out"]) < 1e-2
    assert uo.si_sdr(io["out"], fix["out"]).min() > 40.0
    for i in range(n_steps):
        assert uo.rel_l2(io["dbg"][i], fix["steps"][i]) < 1e-2, i


def test_step_coefs_reduce_to_reference_update():
    """x + s^2 eta score + beta z  ==  a x + b net + c z with the EDM wrapper folded in."""
    _, cfg = build_model("universepp_16k")
    sigma, eta, beta = uo.sampler_constants(cfg.model)
    coefs = program.step_coefs(cfg.model, 8, 1.3)
    x, net, z = 0.37, -0.21, 0.9
    sd = 10 ** (-26 / 20)
    for n in range(8):
        s = float(sigma[n])
        est = sd**2 / (s**2 + sd**2) * x + s * sd / (s**2 + sd**2) ** 0.5 * net
        score = (est - x) / s**2
        ref = x + s**2 * eta * score + beta * float(sigma[n + 1]) * z if n < 7 else x + s**2 * score
        a, b, c = coefs[n]
        assert abs(a * x + b * net + c * z - ref) < 1e-6


def test_arena_tensors_do_not_overlap():
    model, cfg = build_model("universepp_16k")
    pm = weights.pack_model(model.state_dict(), model.model_config())
    pg = program.build_program(pm, model.model_config(), 3, 2000, 8)
    spans = []
    for name, t in pg.tensors.items():
        if isinstance(t, int):
            continue
        es = 2 if t.dtype == 1 else 4
        spans.append((t.off, t.off + t.split * t.nch * t.P * 8 * es, name))
    spans.sort()
    for (a0, a1, an), (b0, b1, bn) in zip(spans, spans[1:]):
        assert a1 <= b0, (an, bn)
    assert spans[-1][1] <= pg.arena


def test_two_lane_program_equals_single_lane_programs():
    """The two-lane program (halves of the batch with their own tensors, io offsets i1 / i2) computes, for each
    half, bit for bit what a single-lane program of that half computes (the CPU emulator's fp32 matmuls block
    differently for another batch size, so the comparison is per lane, not against one program of the whole batch)."""
    model, cfg = build_model("universepp_16k")
    mcfg = model.model_config()
    pm = weights.pack_model(model.state_dict(), mcfg)
    B, T, n_steps = 3, 700, 3
    g = torch.Generator().manual_seed(5)
    mix = 0.1 * torch.randn(B, T, generator=g)
    Lp = T + (model.tot_ds - T % model.tot_ds)
    noise = torch.stack([torch.randn(B, Lp, generator=g) for _ in range(n_steps)])
    coefs = program.step_coefs(mcfg, n_steps, 1.3)
    pg2 = program.build_program(pm, mcfg, B, T, n_steps, lanes=2)
    assert len(pg2.lanes) == 2 and [l.B for l in pg2.lanes] == [2, 1] and pg2.lanes[1].b_off == 2
    assert sum(op.lane for op in pg2.ops) == pg2.lane_ops[1]
    assert sum(1 for op in pg2.ops if op.sync & 1) == 1 and sum(1 for op in pg2.ops if op.sync & 2) == 1
    assert sum(1 for op in pg2.ops if op.step == -2) == 1          # one shared FiLM table
    io2 = Emulator(pm, pg2).run(mix, noise, coefs, dbg=True)
    for b0, b1 in ((0, 2), (2, 3)):
        pg1 = program.build_program(pm, mcfg, b1 - b0, T, n_steps, lanes=1)
        io1 = Emulator(pm, pg1).run(mix[b0:b1], noise[:, b0:b1], coefs, dbg=True)
        assert torch.equal(io1["out"], io2["out"][b0:b1])
        for i in range(n_steps):
            assert torch.equal(io1["dbg"][i], io2["dbg"][i][b0:b1])
    assert program.default_lanes(64) == 2 and program.default_lanes(8) == 1 and program.default_lanes(64, aux=True) == 1
