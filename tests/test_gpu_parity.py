"""GPU parity tests proper: the native sm_100a path, called through the reference-facing API
(model.enhance -> C ABI), against (a) the golden outputs of the real reference and (b) the CPU
oracle on the same seeded inputs.  Gates (north_star): per-step network output and final
waveform rel-L2 <= 1e-2 (bf16 operands, fp32 accumulate), output-vs-reference SI-SDR >= 40 dB."""
import os
import subprocess

import pytest
import torch

from conftest import GOLDEN_CASES, ROOT, build_model, golden_inputs, load_fixture  # noqa: F401
from oracle import universe_oracle as uo

pytestmark = pytest.mark.gpu

TOL = 1e-2
SISDR_MIN = 40.0


def _gpu_model(name, seed=7):
    model, cfg = build_model(name, seed)
    model = model.to("cuda")
    model.eval()
    return model, cfg


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_enhance_matches_reference_golden(case):
    fix = load_fixture(case)
    model, cfg = _gpu_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    out, steps = model.enhance_debug(mix.cuda(), noise.cuda(), **fix["kwargs"])
    out, steps = out.cpu(), steps.cpu()
    assert torch.isfinite(out).all()
    for i in range(steps.shape[0]):
        assert uo.rel_l2(steps[i], fix["steps"][i]) < TOL, f"step {i}"
    assert uo.rel_l2(out, fix["out"]) < TOL
    assert uo.si_sdr(out, fix["out"]).min() > SISDR_MIN


@pytest.mark.parametrize("name,B,T", [("universepp_16k", 3, 5001), ("universepp_16k", 1, 1600), ("universepp_16k", 2, 37),
                                      ("universe_16k", 1, 4799), ("universepp_24k", 1, 2400)])
def test_enhance_matches_cpu_oracle_ragged_shapes(name, B, T):
    """Lengths that are not / exactly a multiple of the total down-sampling (a full extra block is
    padded then, reference universe.py:219-223), very short clips, odd batch sizes."""
    model, cfg = _gpu_model(name)
    g = torch.Generator().manual_seed(T)
    mix = 0.05 * torch.randn(B, T, generator=g)
    Lp = T + (model.tot_ds - T % model.tot_ds)
    noise = torch.randn(8, B, Lp, generator=g)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref, ref_steps = uo.enhance(sd, cfg.model, mix, [n[:, None, :] for n in noise], return_steps=True)
    out, steps = model.enhance_debug(mix.cuda(), noise.cuda())
    for i in range(8):
        assert uo.rel_l2(steps[i].cpu(), ref_steps[i][:, 0]) < TOL, f"step {i}"
    assert uo.rel_l2(out.cpu(), ref) < TOL
    assert uo.si_sdr(out.cpu(), ref).min() > SISDR_MIN


def test_enhance_public_api_rng_ranks_and_ensemble():
    model, cfg = _gpu_model("universepp_16k")
    T = 2000
    mix = 0.05 * torch.randn(2, T, generator=torch.Generator().manual_seed(5)).cuda()
    Lp = T + (model.tot_ds - T % model.tot_ds)
    # the rng is consumed exactly like the reference: n_steps draws of (B, 1, Lp)
    rng = torch.Generator(device="cuda").manual_seed(11)
    out = model.enhance(mix, rng=rng)
    rng2 = torch.Generator(device="cuda").manual_seed(11)
    noise = torch.stack([torch.randn((2, 1, Lp), device="cuda", generator=rng2)[:, 0] for _ in range(8)])
    out2, _ = model.enhance_debug(mix, noise)
    assert torch.equal(out, out2)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref = uo.enhance(sd, cfg.model, mix.cpu(), [n[:, None, :].cpu() for n in noise])
    assert uo.rel_l2(out.cpu(), ref) < TOL
    # rank handling (universe.py:251-257, 370-373)
    assert model.enhance(mix[0], rng=rng).shape == (T,)
    assert model.enhance(mix[:, None, :], rng=rng).shape == (2, 1, T)
    # ensemble (universe.py:261-264, 359-364)
    e = model.enhance(mix, rng=rng, ensemble=3, ensemble_stat="median")
    assert e.shape == (2, T) and torch.isfinite(e).all()
    with pytest.raises(NotImplementedError):
        model.enhance(mix, ensemble=2, ensemble_stat="mode")


def test_signal_median_ensemble_matches_reference_golden():
    """ensemble=3, ensemble_stat="signal_median" (universe.py:261-264,359-368; utils/stats.py:22-66)."""
    fix = load_fixture("pp16k_sigmed")
    model, _ = _gpu_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    kw = fix["kwargs"]
    out = model._enhance_impl(mix.cuda(), kw["n_steps"], None, None, None, None, False, False, kw["ensemble"],
                              kw["ensemble_stat"], None, noise=noise.cuda(), return_steps=False).cpu()
    assert out.shape == fix["out"].shape
    assert uo.rel_l2(out, fix["out"]) < TOL and uo.si_sdr(out, fix["out"]).min() > SISDR_MIN


@pytest.mark.parametrize("case", ["pp16k_aux", "pp16k_warm"])
def test_aux_signal_and_warm_start_match_reference_golden(case):
    """use_aux_signal / warm_start (universe.py:313-331): conditioner incl. the decoder tail and the resumed sampler
    run natively, aux_to_wav + the state initialisation as one AUX launch (csrc/aux_ops.cuh)."""
    from open_universe_b200 import testing

    fix = load_fixture(case)
    model, _ = build_model(fix["model"], fix["seed_weights"])
    testing.perturb_snake_alpha(model)
    model = model.to("cuda").eval()
    mix, noise = golden_inputs(fix, model.tot_ds)
    kw = fix["kwargs"]
    nat = torch.zeros_like(noise)
    if kw.get("warm_start") is not None:   # the reference consumes its draws from the start of the list
        n0 = kw["warm_start"]
        nat[n0:] = noise[: noise.shape[0] - n0]
    out = model._enhance_impl(mix.cuda(), None, None, None, None, None, bool(kw.get("use_aux_signal")), bool(kw.get("keep_rms")),
                              None, "median", kw.get("warm_start"), noise=nat.cuda(), return_steps=False).cpu()
    assert uo.rel_l2(out, fix["out"]) < TOL and uo.si_sdr(out, fix["out"]).min() > SISDR_MIN
    # public call with a device generator
    o2 = model.enhance(mix.cuda(), rng=torch.Generator(device="cuda").manual_seed(1), **kw)
    assert o2.shape == mix.shape and torch.isfinite(o2).all()


def test_known_score_hook_matches_reference_golden():
    """target= / fake_score_snr= (universe.py:269-298): analytic score, closed-form sampler trajectory."""
    fix = load_fixture("pp16k_target")
    model, _ = _gpu_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    kw = fix["kwargs"]
    out = model._enhance_impl(mix.cuda(), None, None, fix["target"].cuda(), kw["fake_score_snr"], None, False, kw["keep_rms"],
                              None, "median", None, noise=noise.cuda(), return_steps=False).cpu()
    assert uo.rel_l2(out, fix["out"]) < 1e-4
    # public call, device generator, default 5 dB score noise: runs, finite, peak-limited
    o2 = model.enhance(mix.cuda(), target=fix["target"].cuda(), rng=torch.Generator(device="cuda").manual_seed(3))
    assert o2.shape == mix.shape and torch.isfinite(o2).all() and o2.abs().max() <= 1.0 + 1e-6


def test_deterministic_and_batch_independent():
    model, _ = _gpu_model("universepp_16k")
    B, T = 4, 3200
    g = torch.Generator().manual_seed(3)
    mix = 0.05 * torch.randn(B, T, generator=g).cuda()
    Lp = T + 160
    noise = torch.randn(8, B, Lp, generator=g).cuda()
    a, _ = model.enhance_debug(mix, noise)
    b, _ = model.enhance_debug(mix, noise)
    assert torch.equal(a, b)
    perm = torch.tensor([2, 0, 3, 1]).cuda()
    c, _ = model.enhance_debug(mix[perm], noise[:, perm])
    assert torch.equal(c, a[perm])
    # a single utterance alone gives the same result as inside the batch
    d, _ = model.enhance_debug(mix[1:2], noise[:, 1:2])
    assert uo.rel_l2(d.cpu(), a[1:2].cpu()) < 1e-6


def test_graph_replay_equals_eager_launches():
    model, _ = _gpu_model("universe_16k")
    g = torch.Generator().manual_seed(9)
    mix = 0.05 * torch.randn(2, 2500, generator=g).cuda()
    noise = torch.randn(8, 2, 2560, generator=g).cuda()
    a0, sa0 = model.enhance_debug(mix, noise)   # first use of a plan: eager launches (the graph is captured lazily)
    a, sa = model.enhance_debug(mix, noise)     # second use: captured + replayed
    a2, sa2 = model.enhance_debug(mix, noise)   # replay
    assert torch.equal(a0, a) and torch.equal(sa0, sa) and torch.equal(a2, a) and torch.equal(sa2, sa)
    model._invalidate()
    model.use_cuda_graph = False
    b, sb = model.enhance_debug(mix, noise)
    assert torch.equal(a, b) and torch.equal(sa, sb)


def test_plan_cache_is_bounded_and_evicted_shapes_reproduce():
    """Many distinct lengths (every file of a real corpus has its own): the cache keeps at most plan_cache_max plans /
    plan_cache_bytes of arenas, closes the least recently used ones, and a shape that was evicted gives the same
    bits when it comes back."""
    model, _ = _gpu_model("universepp_16k")
    model.plan_cache_max = 4
    g = torch.Generator().manual_seed(21)
    first = {}
    lengths = [1600 + 37 * k for k in range(12)]
    for T in lengths + lengths[:3]:
        mix = 0.05 * torch.randn(1, T, generator=torch.Generator().manual_seed(T)).cuda()
        noise = torch.randn(8, 1, T + (160 - T % 160), generator=torch.Generator().manual_seed(T + 1)).cuda()
        out, _ = model.enhance_debug(mix, noise)
        assert len(model._plans) <= 4
        if T in first:
            assert torch.equal(out, first[T])
        first[T] = out
    model.plan_cache_bytes = 1   # a budget below one arena: only the plan in use survives
    model.enhance_debug(0.05 * torch.randn(1, 2500, generator=g).cuda(), torch.randn(8, 1, 2560, generator=g).cuda())
    assert len(model._plans) == 1


def test_two_lanes_equal_one_lane_and_oracle(monkeypatch):
    """B >= 16 is compiled as two utterance lanes on two captured streams (odd B: 9 + 8 here).  Same bits as the
    single-lane program of the whole batch (rows are computed independently of their tile / lane), and both
    within the gate of the CPU oracle."""
    model, cfg = _gpu_model("universepp_16k")
    B, T = 17, 1700
    g = torch.Generator().manual_seed(11)
    mix = 0.05 * torch.randn(B, T, generator=g)
    Lp = T + (model.tot_ds - T % model.tot_ds)
    noise = torch.randn(8, B, Lp, generator=g)
    monkeypatch.delenv("UV_LANES", raising=False)
    out2, st2 = model.enhance_debug(mix.cuda(), noise.cuda(), keep_rms=True)
    pg = model._plans[(B, T, 8, False)].pg
    assert [l.B for l in pg.lanes] == [9, 8]
    monkeypatch.setenv("UV_LANES", "1")
    model._invalidate()
    out1, st1 = model.enhance_debug(mix.cuda(), noise.cuda(), keep_rms=True)
    assert len(model._plans[(B, T, 8, False)].pg.lanes) == 1
    assert torch.equal(out1, out2) and torch.equal(st1, st2)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    ref, ref_steps = uo.enhance(sd, cfg.model, mix, [n[:, None, :] for n in noise], keep_rms=True, return_steps=True)
    for i in range(8):
        assert uo.rel_l2(st2[i].cpu(), ref_steps[i][:, 0]) < TOL, f"step {i}"
    assert uo.rel_l2(out2.cpu(), ref) < TOL and uo.si_sdr(out2.cpu(), ref).min() > SISDR_MIN


def test_gemm_kernel_unit_cases():
    exe = os.path.join(ROOT, "open_universe_b200", "csrc", "test_gemm")
    if not os.path.exists(exe):
        pytest.skip("test_gemm binary not built")
    cases = subprocess.run([exe, "list"], capture_output=True, text=True).stdout.split()
    assert cases
    for c in cases:
        r = subprocess.run([exe, c, "0"], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "PASS" in r.stdout, r.stdout + r.stderr
