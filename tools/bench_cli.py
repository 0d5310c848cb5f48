"""Throughput of the file pipeline (open_universe_b200/bin/enhance.py) on synthetic WAV files: writes a random-init
UNIVERSE++ 16 kHz checkpoint and N noisy clips to a scratch directory, runs the CLI twice (the first run pays the
weight packing, plan creation and graph capture) and prints audio-seconds per wall-second including file IO.
--spread S: every file gets its own length, uniform in [seconds - S, seconds] (a corpus like Voicebank, where no two files
are equally long): the default schedule batches them with per-row lengths, --equal-length-batches shows what identical-
length batching gives on the same files.
Usage: python tools/bench_cli.py [--files 256] [--seconds 4] [--batch-size 64] [--spread 0] [--equal-length-batches]"""
import argparse
import json
import os
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch
import yaml
from scipy.io import wavfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from open_universe_b200 import config, networks, testing  # noqa: E402,F401
from open_universe_b200.bin import enhance as cli  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--files", type=int, default=256)
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--spread", type=float, default=0.0)
    ap.add_argument("--equal-length-batches", action="store_true")
    args = ap.parse_args()
    cfg = bench.load_model_cfg("universepp_16k")
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        (tmp / "exp" / "checkpoints").mkdir(parents=True)
        (tmp / "exp" / ".hydra").mkdir()
        torch.save({"state_dict": model.state_dict()}, tmp / "exp" / "checkpoints" / "last.ckpt")
        with open(tmp / "exp" / ".hydra" / "config.yaml", "w") as f:
            yaml.safe_dump(json.loads(json.dumps(cfg)), f)
        src = tmp / "noisy"
        src.mkdir()
        n = int(args.seconds * 16000)
        x = testing.synthetic_noisy_speech(args.files, n).numpy().astype(np.float32)
        g = np.random.default_rng(3)
        for i in range(args.files):
            ni = n - int(g.uniform(0, args.spread) * 16000) - (i if args.spread > 0 else 0)   # (distinct lengths)
            wavfile.write(str(src / f"clip{i:04d}.wav"), 16000, x[i, : max(ni, 1600)])
        argv = [str(src), str(tmp / "enh"), "--model", str(tmp / "exp" / "checkpoints" / "last.ckpt"),
                "--batch-size", str(args.batch_size)]
        t0 = time.perf_counter()
        m = cli.inference_utils.load_model(argv[3], device="cuda:0")
        t_load = time.perf_counter() - t0
        files, rel, is_dir = cli.find_files(src)
        outs = [cli.output_path_for(p, rel, is_dir, tmp / "enh") for p in files]
        rng = torch.Generator(device="cuda:0").manual_seed(1)
        for k in range(2):
            st = cli.enhance_files(m, files, outs, dict(rng=rng), batch_size=args.batch_size, device="cuda:0",
                                   ragged=not args.equal_length_batches)
            torch.cuda.synchronize()
            print(f"run {k}: {st['files']} files x {args.seconds:g} s (spread {args.spread:g} s) in {st['batches']} batches: {st['seconds']:.2f} s wall "
                  f"-> {st['audio_seconds'] / st['seconds']:.0f} audio-s/s (file read + resample + enhance + write"
                  f"{'; includes packing, plan creation and graph capture' if k == 0 else ''})")
        print(f"load_model: {t_load:.2f} s")


if __name__ == "__main__":
    main()
