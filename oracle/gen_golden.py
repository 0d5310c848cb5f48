"""Generates tests/golden/* by running the REAL reference (oracle/ref_harness.py) in this
container.  Run once here: `python -m oracle.gen_golden`; the GPU box only reads the fixtures.

Each fixture holds the reference's own enhance() output and per-step network outputs for a
model whose weights are rebuilt deterministically at test time by
`open_universe_b200.testing.seed_weights`, plus the state-dict key/shape/parameter-order
listing of the reference modules.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from open_universe_b200 import config, networks, testing  # noqa: E402,F401
from oracle import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

CASES = [
    # name, model, B, T, kwargs
    ("pp16k", "universepp_16k", 2, 3000, {}),
    ("pp16k_opts", "universepp_16k", 1, 2072, dict(n_steps=5, epsilon=1.7, keep_rms=True)),
    ("orig16k", "universe_16k", 2, 3000, {}),
    ("pp24k", "universepp_24k", 2, 3500, {}),
    # enhance() keyword surface off the default path: known-score test hook (target + practically noiseless fake
    # score, so the CPU generator of the reference does not matter) and the signal_median ensemble statistic
    ("pp16k_target", "universepp_16k", 2, 3000, dict(fake_score_snr=200.0, keep_rms=True, with_target=True)),
    ("pp16k_sigmed", "universepp_16k", 1, 2072, dict(n_steps=4, ensemble=3, ensemble_stat="signal_median")),
    # signal decoupling layer paths (universe.py:313-331, universe_gan.py:147-151); snake alpha perturbed
    ("pp16k_aux", "universepp_16k", 2, 3000, dict(use_aux_signal=True, snake_alpha=True)),
    ("pp16k_warm", "universepp_16k", 2, 3000, dict(warm_start=5, keep_rms=True, snake_alpha=True)),
    # BASELINE config length (4 s clips: 401 bottleneck frames = GRU steps).  To keep the fixture small only the
    # final waveforms and the network outputs of utterance 1 at steps 0, 3 and 7 are stored (fp32).
    ("pp16k_4s", "universepp_16k", 2, 64000, dict(keep_steps=[0, 3, 7], keep_row=1)),
]


def make_inputs(B, T, tot_ds, n_steps, seed=1235):
    g = torch.Generator().manual_seed(seed)
    mix = 0.1 * torch.randn(B, T, generator=g) + 0.01
    mix[0] *= 30.0  # makes the peak normalisation (universe.py:356-357) fire for utterance 0
    Lp = T + (tot_ds - T % tot_ds)
    gn = torch.Generator().manual_seed(seed + 1)
    noise = torch.stack([torch.randn(B, Lp, generator=gn) for _ in range(n_steps)])
    return mix, noise


def main():
    os.makedirs(OUT, exist_ok=True)
    listing = {}
    only = set(sys.argv[1:])   # optional: regenerate only the named cases
    for case, name, B, T, kw in CASES:
        if only and case not in only:
            continue
        ref, cfg = rh.build_reference_model(name)
        mine = config.instantiate(config.to_attr(cfg))
        testing.seed_weights(mine, seed=7)
        kw = dict(kw)
        if kw.pop("snake_alpha", False):
            testing.perturb_snake_alpha(mine)
        sd = mine.state_dict()
        missing, unexpected = ref.load_state_dict(sd, strict=False)
        assert not unexpected and all(k.startswith("loss_") for k in missing), (missing[:5], unexpected[:5])
        tot_ds = ref.tot_ds
        n_steps = kw.get("n_steps", cfg["diffusion"]["n_steps"])
        with_target = kw.pop("with_target", False)
        keep_steps, keep_row = kw.pop("keep_steps", None), kw.pop("keep_row", None)
        mix, noise = make_inputs(B * (kw.get("ensemble") or 1), T, tot_ds, n_steps)
        mix = mix[:B]
        extra = {}
        if with_target:
            gt = torch.Generator().manual_seed(99)
            extra["target"] = 0.05 * torch.randn(B, T, generator=gt)
            kw["target"] = extra["target"][:, None, :]   # the reference wants the target in the mixture's 3-D shape
            kw["rng"] = torch.Generator().manual_seed(5)
        out, steps = rh.run_reference_enhance(ref, mix, [n[:, None, :] for n in noise], **kw)
        kw.pop("target", None), kw.pop("rng", None)
        st = torch.stack([s[:, 0] for s in steps]).contiguous() if steps else torch.zeros(0)
        if keep_steps is not None:
            st = st[keep_steps][:, keep_row : keep_row + 1].clone()
            extra.update(keep_steps=keep_steps, keep_row=keep_row)
        fix = dict(model=name, B=B, T=T, kwargs=kw, seed_weights=7, seed_inputs=1235, with_target=with_target, **extra,
                   out=out.contiguous(), steps=st)
        torch.save(fix, os.path.join(OUT, f"enhance_{case}.pt"))
        print(case, "out rms", out.pow(2).mean().sqrt().item(), "max", out.abs().max().item(),
              "net rms", [round(s.pow(2).mean().sqrt().item(), 4) for s in steps])
        if name not in listing:
            rsd = {k: v for k, v in ref.state_dict().items() if not k.startswith("loss_")}
            listing[name] = dict(
                state_dict=[[k, list(v.shape)] for k, v in rsd.items()],
                model_parameters=[list(p.shape) for p in ref.model_parameters()],
            )
    if only:
        return
    with open(os.path.join(OUT, "state_dict_layout.json"), "w") as f:
        json.dump(listing, f)
    # config files of the three named models, resolved (the reference yaml cannot travel either)
    for name in rh.MODEL_YAMLS:
        with open(os.path.join(OUT, f"config_{name}.json"), "w") as f:
            json.dump(dict(model=rh.load_model_cfg(name)), f, indent=1)


if __name__ == "__main__":
    main()
