"""GPU tests of the keyword tail of enhance() and of the throughput-mode noise, through the C ABI:
  * the kernels' own sampler draws (uv_set_noise_seed + noise == NULL) against the CPU restatement oracle/philox.py,
    in INIT_X, CONV_OUT and the AUX record; statistics of what the GPU drew; a seeded enhance() against the CPU
    emulator fed with the same draws,
  * the AUX record (aux_to_wav + state initialisation, csrc/aux_ops.cuh) against the emulator / the oracle's
    aux_to_wav (reference universe_gan.py:147-151),
  * uv_signal_median against the reference algorithm (utils/stats.py:22-66)."""
import numpy as np
import pytest
import torch

from conftest import build_model, load_fixture, golden_inputs
from gpu_util import read_f32, read_planes
from open_universe_b200 import _native, program, testing, weights
from open_universe_b200.networks.universe.universe import signal_median as signal_median_host
from oracle import philox
from oracle import universe_oracle as uo
from oracle.emulator import Emulator

pytestmark = pytest.mark.gpu
SEED = 0x1234ABCD5678


def _plan(B, T, n_steps=8, aux=False, lanes=None, perturb=False):
    model, cfg = build_model("universepp_16k")
    if perturb:
        testing.perturb_snake_alpha(model)
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    mcfg = model.model_config()
    pm = weights.pack_model(sd, mcfg)
    pg = program.build_program(pm, mcfg, B, T, n_steps, aux=aux, lanes=lanes)
    plan = _native.NativePlan(_native.NativeModel(pm.blob.bytes(), 0), pg, use_graph=0)
    coefs = program.step_coefs(mcfg, n_steps, float(cfg.model.diffusion.epsilon))
    return model, cfg, sd, pm, pg, plan, coefs


def test_in_kernel_draws_match_cpu_restatement():
    B, T, n_steps = 4, 5000, 8
    model, cfg, sd, pm, pg, plan, coefs = _plan(B, T, n_steps, lanes=2)   # two lanes of two utterances
    assert getattr(pg, "lanes", None) is not None
    Lp = pg.Lp
    mix = torch.zeros(B, T, device="cuda")
    out = torch.zeros(B, T, device="cuda")
    with pytest.raises(RuntimeError):   # noise == NULL before a seed was set
        plan.run_ops(0, 1, mix.data_ptr(), None, out.data_ptr(), None, coefs, False, None)
    plan.set_noise_seed(SEED)
    want = philox.noise(SEED, n_steps, B, Lp)
    sigma0 = float(program.sampler_tables(model.model_config(), n_steps)[0][0])
    pick = [(0.0, 0.0, 1.0)] * n_steps   # x <- z_{n+1}
    seen = np.zeros((n_steps, B, Lp), np.float32)
    for lane, lp in enumerate(pg.lanes):
        base = 0 if lane == 0 else pg.lane_ops[0]
        names = pg.names[base : base + len(lp.ops)]
        pre = "" if lane == 0 else "L1."
        i = base + names.index(pre + "init_x")
        plan.run_ops(i, i + 1, mix.data_ptr(), None, out.data_ptr(), None, coefs, False, None)
        x = read_f32(plan, lp.x, lp.B * Lp).view(lp.B, Lp).numpy()
        b0 = 0 if lane == 0 else pg.lanes[0].B
        seen[0, b0 : b0 + lp.B] = x / sigma0
        for n in range(n_steps - 1):
            i = base + names.index(pre + f"s{n}.out")
            plan.run_ops(i, i + 1, mix.data_ptr(), None, out.data_ptr(), None, pick, False, None)
            seen[n + 1, b0 : b0 + lp.B] = read_f32(plan, lp.x, lp.B * Lp).view(lp.B, Lp).numpy()
    plan.status()
    # same Philox words; logf / sincospif differ from numpy's by a few ulp
    assert np.abs(seen - want).max() < 2e-5
    flat = seen.reshape(n_steps * B, Lp)
    assert abs(flat.mean()) < 1e-2 and abs(flat.std() - 1.0) < 1e-2
    c = np.corrcoef(flat)
    assert np.abs(c - np.eye(len(c))).max() < 6e-2   # independent across (step, utterance); 5k samples per row
    plan.close()


def test_seeded_enhance_matches_emulator_with_the_same_draws():
    B, T = 3, 4000
    model, cfg = build_model("universepp_16k")
    mcfg = model.model_config()
    pm = weights.pack_model({k: v.detach().float() for k, v in model.state_dict().items()}, mcfg)
    n_steps = int(cfg.model.diffusion.n_steps)
    pg = program.build_program(pm, mcfg, B, T, n_steps, lanes=1)
    g = torch.Generator().manual_seed(5)
    mix = 0.1 * torch.randn(B, T, generator=g)
    noise = torch.from_numpy(philox.noise(SEED, n_steps, B, pg.Lp))
    coefs = program.step_coefs(mcfg, n_steps, float(cfg.model.diffusion.epsilon))
    ref = Emulator(pm, pg).run(mix, noise, coefs)["out"]
    gm = model.to("cuda").eval()
    out = gm.enhance(mix.cuda(), noise_seed=SEED).cpu()
    assert uo.rel_l2(out, ref) < 1e-2 and uo.si_sdr(out, ref).min() > 40.0
    again = gm.enhance(mix.cuda(), noise_seed=SEED).cpu()
    assert torch.equal(out, again)                                    # a pure function of the seed
    other = gm.enhance(mix.cuda(), noise_seed=SEED + 1).cpu()
    assert uo.rel_l2(other, out) > 1e-3
    # the parity mode is untouched: torch draws, reference order
    o3 = gm.enhance(mix.cuda(), rng=torch.Generator(device="cuda").manual_seed(3)).cpu()
    o4 = gm.enhance(mix.cuda(), rng=torch.Generator(device="cuda").manual_seed(3)).cpu()
    assert torch.equal(o3, o4)


@pytest.mark.parametrize("warm", [None, 5])
def test_aux_record_matches_emulator(warm):
    """One AUX launch against the CPU emulator's op (oracle resample / snake / conv) on the GPU's own y_hat."""
    B, T, n_steps = 3, 6000, 8
    model, cfg, sd, pm, pg, plan, coefs = _plan(B, T, n_steps, aux=True, perturb=True)
    assert pg.idx_aux is not None
    g = torch.Generator().manual_seed(11)
    mix = 0.1 * torch.randn(B, T, generator=g)
    noise = torch.stack([torch.randn(B, pg.Lp, generator=g) for _ in range(n_steps)])
    dmix, dnoise, dout = mix.cuda(), noise.cuda(), torch.zeros(B, T, device="cuda")
    io = (dmix.data_ptr(), dnoise.data_ptr(), dout.data_ptr(), None, coefs, False, None)
    plan.run_ops(0, pg.idx_init_x, *io)
    i = pg.idx_aux if warm is None else pg.idx_aux + 1 + warm
    plan.run_ops(i, i + 1, *io)
    plan.status()
    x = read_f32(plan, pg.x, B * pg.Lp).view(B, pg.Lp)
    y_hat = read_planes(plan, pg.y_hat, B, pg.Lp)
    sig = uo.aux_to_wav(sd, y_hat)[:, 0, :]
    if warm is not None:
        sig = sig + float(program.sampler_tables(model.model_config(), n_steps)[0][warm]) * noise[warm]
    assert sig.std() > 1e-3 and uo.rel_l2(x, sig) < 1e-5
    # the in-kernel draw of the warm start
    if warm is not None:
        plan.set_noise_seed(SEED)
        plan.run_ops(i, i + 1, dmix.data_ptr(), None, dout.data_ptr(), None, coefs, False, None)
        plan.status()
        x2 = read_f32(plan, pg.x, B * pg.Lp).view(B, pg.Lp)
        z = torch.from_numpy(philox.noise(SEED, warm + 1, B, pg.Lp)[warm])
        want = uo.aux_to_wav(sd, y_hat)[:, 0, :] + float(program.sampler_tables(model.model_config(), n_steps)[0][warm]) * z
        assert uo.rel_l2(x2, want) < 1e-4
    plan.close()


def test_seeded_warm_start_runs_through_the_public_call():
    fix = load_fixture("pp16k_warm")
    model, _ = build_model(fix["model"], fix["seed_weights"])
    model = model.to("cuda").eval()
    mix, _ = golden_inputs(fix, model.tot_ds)
    kw = dict(fix["kwargs"])
    a = model.enhance(mix.cuda(), noise_seed=9, **kw)
    b = model.enhance(mix.cuda(), noise_seed=9, **kw)
    assert a.shape == mix.shape and torch.isfinite(a).all() and torch.equal(a, b)


@pytest.mark.parametrize("E", [2, 3, 4, 7])
def test_signal_median_kernel_matches_reference_algorithm(E):
    g = torch.Generator().manual_seed(E)
    B, S = 3, 20011
    sig = torch.randn(E, B, 1, S, generator=g)
    sig[1, 0] += 0.2   # make the vote less even
    want = signal_median_host(sig)
    got = _native.signal_median(sig.cuda()).cpu()
    assert got.shape == want.shape and torch.equal(got, want)
