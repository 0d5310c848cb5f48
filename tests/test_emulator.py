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


def _spans(pg):
    out = []
    for name, t in pg.tensors.items():
        if isinstance(t, int):
            continue
        out.append((t.off, t.off + program._tens_bytes(t), name, (t.Lh, t.P, t.dtype)))
    return sorted(out)


def test_arena_tensors_do_not_overlap_without_recycling():
    model, cfg = build_model("universepp_16k")
    pm = weights.pack_model(model.state_dict(), model.model_config())
    pg = program.build_program(pm, model.model_config(), 3, 2000, 8, recycle=False)
    spans = _spans(pg)
    for (a0, a1, an, _), (b0, b1, bn, _) in zip(spans, spans[1:]):
        assert a1 <= b0, (an, bn)
    assert spans[-1][1] <= pg.arena and spans[0][0] >= pg.sig


@pytest.mark.parametrize("name,B", [("universepp_16k", 3), ("universepp_16k", 17), ("universe_16k", 2), ("universepp_24k", 2)])
def test_recycled_arena_shares_slots_only_between_dead_tensors_of_one_geometry(name, B):
    """finalize_arena(): two tensors may overlap in memory only if their live ranges are disjoint (an op's outputs
    never share planes with its inputs) and they have the same row geometry (the never-written padding rows of a plane
    stay zero); raw buffers are never shared; the program is smaller than with one slot per tensor."""
    model, cfg = build_model(name)
    mcfg = model.model_config()
    pm = weights.pack_model(model.state_dict(), mcfg)
    pg = program.build_program(pm, mcfg, B, 2000, 8, recycle=True)
    plain = program.build_program(pm, mcfg, B, 2000, 8, recycle=False)
    assert pg.arena < 0.6 * plain.arena
    for lane in pg.lanes:
        spans = _spans(lane)
        assert min(s[0] for s in spans) >= max(lane.x, lane.sig) and max(s[1] for s in spans) <= lane.arena
        for i, (a0, a1, an, ag) in enumerate(spans):
            for b0, b1, bn, bg in spans[i + 1:]:
                if b0 >= a1:
                    break
                (alo, ahi), (blo, bhi) = lane.live[an], lane.live[bn]
                assert ahi < blo or bhi < alo, (an, bn, lane.live[an], lane.live[bn])
                assert ag == bg, (an, bn)
    if len(pg.lanes) == 2:   # lanes run concurrently: disjoint arenas
        assert max(s[1] for s in _spans(pg.lanes[0])) <= min(s[0] for s in _spans(pg.lanes[1]))


@pytest.mark.parametrize("case", ["pp16k_aux", "pp16k_warm"])
def test_aux_program_emulation_matches_reference(case):
    """use_aux_signal / warm_start as the native path runs them: front end + conditioner, ONE AUX record
    (aux_to_wav + state initialisation, csrc/aux_ops.cuh), the resumed sampler steps, POST -- against the
    reference's golden outputs (universe.py:313-331, universe_gan.py:147-151)."""
    from open_universe_b200 import testing

    fix = load_fixture(case)
    model, cfg = build_model(fix["model"], fix["seed_weights"])
    testing.perturb_snake_alpha(model)
    mcfg = model.model_config()
    kw = fix["kwargs"]
    n_steps = kw.get("n_steps", int(cfg.model.diffusion.n_steps))
    pm = weights.pack_model({k: v.detach().float() for k, v in model.state_dict().items()}, mcfg)
    assert pm.layers["s.aux"]["mode"] == 0 and pm.layers["s.aux"]["KU"] == 15 and pm.layers["s.aux"]["KD"] == 28
    pg = program.build_program(pm, mcfg, fix["B"], fix["T"], n_steps, aux=True)
    assert pg.names[pg.idx_aux] == "aux" and len(pg.ops) == pg.idx_aux + n_steps + 1
    mix, noise = golden_inputs(fix, model.tot_ds)
    nat = torch.zeros_like(noise)
    n0 = kw.get("warm_start")
    if n0 is not None:   # the reference consumes its draws from the start of the list
        nat[n0:] = noise[: noise.shape[0] - n0]
    coefs = program.step_coefs(mcfg, n_steps, kw.get("epsilon", float(cfg.model.diffusion.epsilon)))
    emu = Emulator(pm, pg)
    io = emu.run(mix, nat, coefs, keep_rms=kw.get("keep_rms", False), first=0, last=pg.idx_init_x)
    if n0 is None:
        emu.run_op(pg.idx_aux, io)
    else:
        emu.run_op(pg.idx_aux + 1 + n0, io)
        for i in range(pg.step_first_op[n0], pg.idx_post):
            emu.run_op(i, io)
    emu.run_op(pg.idx_post, io)
    assert uo.rel_l2(io["out"], fix["out"]) < 1e-2
    assert uo.si_sdr(io["out"], fix["out"]).min() > 40.0


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_fp32_program_emulation_matches_reference(case):
    """precision="fp32" (the rel-L2 <= 1e-4 tier of north_star): the same compiler emits f32 planes and names the fp32
    weight images; executed op by op on the CPU it reproduces the reference's golden outputs to fp32 round-off."""
    fix = load_fixture(case)
    model, cfg = build_model(fix["model"], fix["seed_weights"])
    mcfg = model.model_config()
    kw = fix["kwargs"]
    n_steps = kw.get("n_steps", int(cfg.model.diffusion.n_steps))
    pm = weights.pack_model(model.state_dict(), mcfg, hp=True)
    with pytest.raises(ValueError):
        program.build_program(weights.PackedModel(), mcfg, fix["B"], fix["T"], n_steps, precision="fp32")
    pg = program.build_program(pm, mcfg, fix["B"], fix["T"], n_steps, precision="fp32")
    from open_universe_b200 import _abi
    assert all(op.type != _abi.OP_CONVBLOCK for op in pg.ops)
    assert all(t.dtype == _abi.DT_F32 for t in pg.tensors.values() if not isinstance(t, int))
    mix, noise = golden_inputs(fix, model.tot_ds)
    coefs = program.step_coefs(mcfg, n_steps, kw.get("epsilon", float(cfg.model.diffusion.epsilon)))
    io = Emulator(pm, pg).run(mix, noise, coefs, keep_rms=kw.get("keep_rms", False), dbg=True)
    assert uo.rel_l2(io["out"], fix["out"]) < 1e-5
    for i in range(n_steps):
        assert uo.rel_l2(io["dbg"][i], fix["steps"][i]) < 1e-5, i
