"""Drop-in boundary on the host side: state-dict layout, EMA parameter order, config handling,
load_model / add_enhance_arguments contracts (reference inference_utils/, universe.py:130-133)."""
import argparse
import json
import os

import pytest
import torch
import yaml

from conftest import GOLDEN, build_model, load_cfg
from open_universe_b200 import config, inference_utils, networks, testing  # noqa: F401
from open_universe_b200.networks.universe import layers


@pytest.mark.parametrize("name", ["universepp_16k", "universe_16k", "universepp_24k"])
def test_state_dict_layout_matches_reference(name):
    with open(os.path.join(GOLDEN, "state_dict_layout.json")) as f:
        ref = json.load(f)[name]
    model, _ = build_model(name)
    mine = [[k, list(v.shape)] for k, v in model.state_dict().items()]
    assert mine == ref["state_dict"]
    assert [list(p.shape) for p in model.model_parameters()] == ref["model_parameters"]


def test_mel_buffers_match_torchaudio():
    ta = pytest.importorskip("torchaudio")
    m = ta.transforms.MelSpectrogram(sample_rate=24000, n_mels=80, n_fft=640, hop_length=160, center=False)
    assert torch.allclose(layers.mel_filterbank(321, 80, 24000), m.mel_scale.fb, atol=1e-6)
    assert torch.allclose(layers.hann_window(640), m.spectrogram.window, atol=1e-6)


def test_seed_weights_deterministic():
    a, _ = build_model("universe_16k", seed=3)
    b, _ = build_model("universe_16k", seed=3)
    for (k, v), (_, w) in zip(a.state_dict().items(), b.state_dict().items()):
        assert torch.equal(v, w), k


def test_config_interpolation_and_float_coercion(tmp_path):
    node = {"model": {"a": {"x": 3, "r": [2, 4]}, "b": {"x": "${model.a.x}", "r": "${model.a.r}", "s": "5e-4", "m": "${datamodule.foo}"}}}
    cfg = config.resolve(node)
    assert cfg.model.b.x == 3 and cfg.model.b.r == [2, 4] and cfg.model.b.s == 5e-4 and cfg.model.b.m is None


def test_unknown_target_raises():
    with pytest.raises(ValueError):
        config.instantiate(config.to_attr({"_target_": "open_universe.networks.foo.Bar"}))


def _write_ckpt(tmp_path, name, with_ema):
    model, cfg = build_model(name)
    d = tmp_path / "exp" / "checkpoints"
    d.mkdir(parents=True)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    sd["loss_mpd.fake.weight"] = torch.zeros(3)  # training-only keys must be ignored
    data = {"state_dict": sd}
    shadow = None
    if with_ema:
        shadow = [p.detach().clone() + 0.5 for p in model.model_parameters()]
        data["ema"] = dict(decay=0.999, num_updates=10, shadow_params=shadow, collected_params=None)
    torch.save(data, d / "last.ckpt")
    hy = tmp_path / "exp" / ".hydra"
    hy.mkdir()
    with open(hy / "config.yaml", "w") as f:
        yaml.safe_dump(json.loads(json.dumps(cfg)), f)
    return d / "last.ckpt", model, shadow


def test_load_model_applies_ema_in_parameter_order(tmp_path):
    ckpt, src, shadow = _write_ckpt(tmp_path, "universepp_16k", with_ema=True)
    model, cfg = inference_utils.load_model(ckpt, device="cpu", return_config=True)
    assert not model.training and cfg.model.fs == 16000
    for p, s in zip(model.model_parameters(), shadow):
        assert torch.equal(p, s)


def test_load_model_without_ema_and_missing_config(tmp_path):
    ckpt, src, _ = _write_ckpt(tmp_path, "universe_16k", with_ema=False)
    model = inference_utils.load_model(ckpt, device="cpu")
    for (k, v), (_, w) in zip(model.state_dict().items(), src.state_dict().items()):
        assert torch.equal(v, w), k
    os.remove(tmp_path / "exp" / ".hydra" / "config.yaml")
    with pytest.raises(ValueError):
        inference_utils.load_model(ckpt, device="cpu")


def test_load_model_nonexistent_path_raises():
    with pytest.raises(Exception):
        inference_utils.load_model("no-such-org/no-such-model:rev")


def test_add_enhance_arguments():
    model, _ = build_model("universepp_16k")
    parser = inference_utils.add_enhance_arguments(model, argparse.ArgumentParser())
    args = parser.parse_args(["--n_steps", "4", "--keep_rms", "1"])
    assert args.n_steps == 4 and args.epsilon == 1.3 and args.ensemble_stat is None and args.keep_rms is True
    with pytest.raises(ValueError):
        inference_utils.add_enhance_arguments(object(), argparse.ArgumentParser())


def test_enhance_fails_loudly_without_gpu_and_validates_input():
    model, _ = build_model("universepp_16k")
    model.eval()
    with pytest.raises(ValueError):
        model.enhance(torch.zeros(1, 1, 1, 100))
    with pytest.raises(NotImplementedError):
        model.enhance(torch.zeros(1, 100), ensemble=2, ensemble_stat="bogus")
    with pytest.raises(RuntimeError, match="no CPU / PyTorch fallback"):
        model.enhance(torch.zeros(1, 1600))
    with pytest.raises(RuntimeError):
        model.get_score_model()(None, None, None)


def test_reference_import_paths_resolve_to_this_package():
    """`from open_universe import inference_utils` (reference README.md:27-49) and the `_target_` paths of the model
    configs resolve through the alias package to the very same module objects."""
    import open_universe
    import open_universe_b200
    from open_universe import inference_utils
    from open_universe.networks.universe import Universe, UniverseGAN

    assert inference_utils is open_universe_b200.inference_utils
    assert UniverseGAN is open_universe_b200.networks.universe.UniverseGAN and Universe is open_universe_b200.networks.universe.Universe
    assert hasattr(inference_utils, "load_model") and hasattr(inference_utils, "add_enhance_arguments")
    assert open_universe.networks is open_universe_b200.networks
