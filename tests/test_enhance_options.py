"""Host-side logic of the enhance() keyword surface that does not touch the networks: the known-score test
hook (reference universe.py:269-298) against the golden output of the real reference, and the
signal_median ensemble statistic (utils/stats.py:22-66)."""
import torch

from conftest import build_model, golden_inputs, load_fixture
from open_universe_b200.networks.universe.universe import signal_median
from oracle import universe_oracle as uo


def test_known_score_path_matches_reference_golden_on_cpu():
    fix = load_fixture("pp16k_target")
    model, _ = build_model(fix["model"], fix["seed_weights"])
    mix, noise = golden_inputs(fix, model.tot_ds)
    kw = fix["kwargs"]
    x = model._enhance_known_score(mix[:, None, :], fix["target"][:, None, :], 8, float(model.diff_kwargs.epsilon),
                                   kw["fake_score_snr"], torch.Generator().manual_seed(5), kw["keep_rms"], noise)
    assert uo.rel_l2(x[:, 0], fix["out"]) < 1e-5


def test_known_score_converges_to_target():
    """With a (nearly) exact score the sampler must land on the normalised target (closed form: the last step
    maps x to target up to the score noise)."""
    model, _ = build_model("universepp_16k", 7)
    g = torch.Generator().manual_seed(0)
    mix = 0.1 * torch.randn(2, 1, 4000, generator=g)
    target = 0.05 * torch.randn(2, 1, 4000, generator=g)
    x = model._enhance_known_score(mix, target, 8, 1.3, 300.0, g, False, None)
    tp, pad = model.pad(target)   # the reference normalises the padded signals
    t = tp - tp.mean(dim=(1, 2), keepdim=True)
    t = model.unpad(t * (10 ** (-26 / 20) / t.std(dim=(1, 2), keepdim=True)), pad)
    assert uo.rel_l2(x, t) < 1e-3


def test_signal_median_picks_a_whole_member():
    g = torch.Generator().manual_seed(1)
    sig = torch.randn(5, 3, 1, 257, generator=g)
    out = signal_median(sig)
    assert out.shape == sig.shape[1:]
    for b in range(3):
        assert any(torch.equal(out[b], sig[e, b]) for e in range(5))
    # one member: returned as is
    assert torch.equal(signal_median(sig[:1]), sig[0])
    # two members are exact copies of a third one placed in the middle of the sort order everywhere
    base = torch.randn(1, 2, 1, 64, generator=g)
    ens = torch.cat([base - 1.0, base, base + 1.0], dim=0)
    assert torch.equal(signal_median(ens), signal_median(ens.clone()))
