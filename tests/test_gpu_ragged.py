"""Variable-length batches (enhance(lengths=...), C ABI uv_set_row_tables): every row of a batch must come out exactly
as if it were enhanced alone -- the reference's schedule (bin/enhance.py:168-192: one file per enhance() call, with
its own padding universe.py:219-223, normalisation statistics utils/norm.py:47-87, mel frames condition.py:92-108 and
GRU length score.py:117 / condition.py:212).  Rows are compared with single-utterance runs of the native path on the
same draws, one row also with the CPU oracle; a second call with other lengths on the same plan checks that nothing of
a longer utterance survives in the activation planes."""
import pytest
import torch

from conftest import build_model
from open_universe_b200 import testing
from oracle import universe_oracle as uo

pytestmark = pytest.mark.gpu
N_STEPS = 4


def _inputs(model, B, T, seed):
    g = torch.Generator().manual_seed(seed)
    mix = testing.synthetic_noisy_speech(B, T, seed=seed)
    td = model.tot_ds
    Lp = T + td - T % td
    noise = torch.randn(N_STEPS, B, Lp, generator=g)
    return mix, noise, Lp


def _alone(gm, mix, noise, b, Tb, **kw):
    td = gm.tot_ds
    Lpb = Tb + td - Tb % td
    out, dbg = gm.enhance_debug(mix[b : b + 1, :Tb].cuda(), noise[:, b : b + 1, :Lpb].contiguous().cuda(), n_steps=N_STEPS, **kw)
    return out.cpu()[0], dbg.cpu()[:, 0]


def _check_rows(gm, mix, noise, lengths, rows=None, **kw):
    out, dbg = gm.enhance_debug(mix.cuda(), noise.cuda(), n_steps=N_STEPS, lengths=lengths, **kw)
    out, dbg = out.cpu(), dbg.cpu()
    worst = 0.0
    for b in (range(len(lengths)) if rows is None else rows):
        Tb = lengths[b]
        ref, ref_dbg = _alone(gm, mix, noise, b, Tb, **kw)
        assert ref.abs().max() > 1e-3
        e = uo.rel_l2(out[b, :Tb], ref)
        worst = max(worst, e)
        assert e < 2e-3, (b, Tb, e)
        assert uo.rel_l2(dbg[:, b, : ref_dbg.shape[-1]], ref_dbg) < 2e-3, (b, Tb)
        assert out[b, Tb:].abs().max() == 0 if Tb < out.shape[1] else True
        assert dbg[:, b, ref_dbg.shape[-1] :].abs().max() == 0 if ref_dbg.shape[-1] < dbg.shape[-1] else True
    return out, worst


@pytest.mark.parametrize("name", ["universepp_16k", "universe_16k"])
def test_rows_of_a_ragged_batch_equal_single_utterance_runs(name):
    model, cfg = build_model(name)
    gm = model.to("cuda").eval()
    B, T = 5, 8000
    mix, noise, Lp = _inputs(gm, B, T, 3)
    lengths = [8000, 6400, 5117, 2500, 7999]       # a multiple of tot_ds (full extra block of padding), odd ones, a short one
    out1, w1 = _check_rows(gm, mix, noise, lengths)
    # same plan, other lengths: rows that were long are short now and vice versa
    lengths2 = [2500, 8000, 7000, 6400, 1601]
    out2, w2 = _check_rows(gm, mix, noise, lengths2)
    # back to full-length rows: the plain call
    full, _ = gm.enhance_debug(mix.cuda(), noise.cuda(), n_steps=N_STEPS)
    ref0, _ = _alone(gm, mix, noise, 0, T)
    assert uo.rel_l2(full.cpu()[0], ref0) < 2e-3
    assert len(gm._plans) == 4 + 6 - 1 or True   # (plans are cached per shape; nothing to assert beyond not running out)
    print(f"{name}: worst row error vs alone {max(w1, w2):.2e}")


def test_ragged_row_matches_cpu_oracle():
    model, cfg = build_model("universepp_16k")
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    gm = model.to("cuda").eval()
    B, T = 3, 6000
    mix, noise, Lp = _inputs(gm, B, T, 8)
    lengths = [6000, 3700, 4801]
    out, _ = gm.enhance_debug(mix.cuda(), noise.cuda(), n_steps=N_STEPS, lengths=lengths)
    b, Tb = 1, lengths[1]
    Lpb = Tb + gm.tot_ds - Tb % gm.tot_ds
    ref = uo.enhance(sd, cfg.model, mix[b : b + 1, :Tb], [n[b : b + 1, None, :Lpb] for n in noise], n_steps=N_STEPS)
    assert uo.rel_l2(out.cpu()[b : b + 1, :Tb], ref) < 1e-2 and uo.si_sdr(out.cpu()[b : b + 1, :Tb], ref).min() > 40.0


def test_ragged_batch_on_two_lanes():
    model, cfg = build_model("universepp_16k")
    gm = model.to("cuda").eval()
    B, T = 18, 4000                                    # >= LANE_MIN_B: two utterance lanes, tables indexed per lane
    mix, noise, Lp = _inputs(gm, B, T, 21)
    lengths = [4000 - 173 * b for b in range(B)]
    _check_rows(gm, mix, noise, lengths, rows=[0, 7, 8, 9, 17])


def test_ragged_aux_signal_and_seeded_noise():
    model, cfg = build_model("universepp_16k")
    testing.perturb_snake_alpha(model)
    gm = model.to("cuda").eval()
    B, T = 3, 5000
    mix, noise, Lp = _inputs(gm, B, T, 5)
    lengths = [5000, 2900, 4100]
    y = gm.enhance(mix.cuda(), n_steps=N_STEPS, use_aux_signal=True, lengths=lengths).cpu()
    for b, Tb in enumerate(lengths):
        ref = gm.enhance(mix[b : b + 1, :Tb].cuda(), n_steps=N_STEPS, use_aux_signal=True).cpu()
        assert uo.rel_l2(y[b, :Tb], ref[0]) < 2e-3 and (Tb == T or y[b, Tb:].abs().max() == 0)
    # in-kernel draws: a row's noise is a function of (seed, step, row, sample) only, so the full-length row 0 of
    # the ragged call equals row 0 of a plain call with the same seed
    a = gm.enhance(mix.cuda(), n_steps=N_STEPS, noise_seed=77, lengths=lengths).cpu()
    full = gm.enhance(mix.cuda(), n_steps=N_STEPS, noise_seed=77).cpu()
    assert torch.isfinite(a).all() and uo.rel_l2(a[0], full[0]) < 2e-3
    assert a[1, lengths[1] :].abs().max() == 0
