"""Generates tests/golden/cli_flow.pt by running the REAL reference's file flow in this container (TEST INFRASTRUCTURE).

The reference enhances a directory file by file (open_universe/bin/enhance.py:168-192):
    audio, fs = torchaudio.load(path); audio = resample(audio, fs, model.fs)        # :77-80, :183-188
    enh = model.enhance(audio, **enhance_kwargs); enh = resample(enh, model.fs, fs)  # :189-190
    torchaudio.save(output_path, enh.cpu(), fs)                                      # :192
(a C-channel file is a batch of C utterances).  This script runs exactly those steps on three synthetic "files" --
mono at the model rate, stereo at the model rate, mono at 22.05 kHz (resampled in and out) -- with the real reference
model (oracle/ref_harness.py) and an injected noise sequence per file, and records inputs, draws and outputs.
`resample` below restates the reference's three-line helper (:77-80) around the same torchaudio call.
Run once here: `python -m oracle.gen_golden_cli`; the GPU box only reads the fixture."""
import os
import sys

import torch
import torchaudio

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from open_universe_b200 import config, networks, testing  # noqa: E402,F401
from oracle import ref_harness as rh  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "cli_flow.pt")

FILES = [  # name, fs, channels, samples
    ("a_mono16k.wav", 16000, 1, 3200),
    ("sub/b_stereo16k.wav", 16000, 2, 3000),
    ("sub/c_mono22k.wav", 22050, 1, 4410),
]


def resample(audio, fs, target_fs):   # reference bin/enhance.py:77-80
    if fs != target_fs:
        audio = torchaudio.functional.resample(audio, fs, target_fs)
    return audio


def main():
    name = "universepp_16k"
    ref, cfg = rh.build_reference_model(name)
    mine = config.instantiate(config.to_attr(cfg))
    testing.seed_weights(mine, seed=7)
    missing, unexpected = ref.load_state_dict(mine.state_dict(), strict=False)
    assert not unexpected and all(k.startswith("loss_") for k in missing)
    n_steps = cfg["diffusion"]["n_steps"]
    g = torch.Generator().manual_seed(4321)
    files = []
    for fname, fs, ch, n in FILES:
        audio = 0.1 * torch.randn(ch, n, generator=g) + 0.005
        x = resample(audio, fs, ref.fs)
        T = x.shape[-1]
        Lp = T + (ref.tot_ds - T % ref.tot_ds)
        noise = torch.stack([torch.randn(ch, Lp, generator=g) for _ in range(n_steps)])
        enh, _ = rh.run_reference_enhance(ref, x, [z[:, None, :] for z in noise])
        out = resample(enh, ref.fs, fs)
        files.append(dict(name=fname, fs=fs, audio=audio, noise=noise, out=out.contiguous(), model_rate_len=T))
        print(fname, tuple(audio.shape), "->", tuple(x.shape), "->", tuple(out.shape), "rms", out.pow(2).mean().sqrt().item())
    torch.save(dict(model=name, seed_weights=7, files=files), OUT)


if __name__ == "__main__":
    main()
