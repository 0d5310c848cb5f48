"""Enhance a file or a directory of audio files -- the caller users actually run
(reference: open_universe/bin/enhance.py:24-192).

Same command line as the reference (`input output --model --hf-token --model-strict --seed --device` plus the
keywords of `model.enhance`, added by `inference_utils.add_enhance_arguments`), same output layout (the folder
structure is retained), same per-file arithmetic: a file of C channels is a batch of C utterances
(enhance.py:187-190), resampled to the model rate and back with `torchaudio.functional.resample`.

What differs is the schedule.  The reference enhances one file at a time (B = #channels, a few thousand eager
kernel launches each).  Here the files stream through in windows (decoded ahead by a thread pool, written back by
writer threads); the utterances of a window are grouped into batches of SIMILAR length and enhanced with
`model.enhance(x, lengths=...)`: every row keeps the padding, normalisation statistics, mel frames and GRU length of
its own file (universe.py:219-223,272), so a batch leaves every utterance's result what the file-by-file loop gives
(`--equal-length-batches` restores batches of identical length only); under `torchrun` the files are dealt whole to the ranks (one process per GPU, no
collective: the outputs are files), and each batch is one replay of the native CUDA graph.  `--batch-size` bounds the batch.

The noise of a batch is drawn by `enhance()` from the seeded generator in batch order, so results are reproducible
for a given seed, input set and batch size, but they are not the reference's file-by-file draws.

    python -m open_universe_b200.bin.enhance noisy/ enhanced/ --model exp/universepp/checkpoints/last.ckpt
    torchrun --nproc-per-node 8 -m open_universe_b200.bin.enhance noisy/ enhanced/ --model ...
"""
import argparse
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch

from open_universe_b200 import _native, inference_utils, sharding

AUDIO_SUFFIXES = [".wav", ".mp3", ".flac"]   # reference enhance.py:33


def handle_help():
    """Defers --help until the model's own arguments are known (reference enhance.py:36-57)."""
    help_idx = -1
    for flag in ("--help", "-h"):
        if flag in sys.argv:
            help_idx = sys.argv.index(flag)
    if "--model" not in sys.argv:
        return False
    if help_idx >= 0:
        sys.argv.pop(help_idx)
        return True
    return False


def find_files(path):
    """reference enhance.py:60-74"""
    path = Path(path)
    if not path.is_dir():
        return [path], path.parent, False
    files = sorted(p for p in path.rglob("*") if p.suffix in AUDIO_SUFFIXES)
    return files, path, True


# torchaudio's file IO needs a codec backend (torchcodec), absent from this image: the first failed attempt is
# remembered, because a failing torchaudio.load / save costs ~0.5 ms per file (7x the WAV read itself; measured: the
# decode of a 64-file window took longer than its batch on the device).  None = untried, True = works, False = skip.
_TORCHAUDIO_IO = {"load": None, "save": None}


def load_audio(path):
    """-> (float32 tensor [channels, samples], fs).  torchaudio.load when its codec backend exists (the reference's
    call, enhance.py:183), else WAV through scipy (this image has no torchcodec)."""
    if _TORCHAUDIO_IO["load"] is not False:
        try:
            import torchaudio

            audio, fs = torchaudio.load(str(path))
            _TORCHAUDIO_IO["load"] = True
            return audio.float(), int(fs)
        except (ImportError, RuntimeError, OSError):
            if _TORCHAUDIO_IO["load"] is None:
                _TORCHAUDIO_IO["load"] = False
    from scipy.io import wavfile

    if Path(path).suffix.lower() != ".wav":
        raise RuntimeError(f"{path}: no decoder for this format here (torchaudio's codec backend is missing); WAV works")
    fs, data = wavfile.read(str(path))
    if data.dtype == np.int16:
        x = data.astype(np.float32) / 32768.0
    elif data.dtype == np.int32:
        x = data.astype(np.float32) / 2147483648.0
    elif data.dtype == np.uint8:
        x = (data.astype(np.float32) - 128.0) / 128.0
    else:
        x = data.astype(np.float32, copy=False)   # float32 files: no copy
    if x.ndim == 1:
        return torch.from_numpy(x)[None], int(fs)
    return torch.from_numpy(np.ascontiguousarray(x.T)), int(fs)


def save_audio(path, audio, fs):
    """audio: float tensor [channels, samples] (reference: torchaudio.save, enhance.py:192)."""
    if _TORCHAUDIO_IO["save"] is not False:
        try:
            import torchaudio

            torchaudio.save(str(path), audio, fs)
            _TORCHAUDIO_IO["save"] = True
            return
        except (ImportError, RuntimeError, OSError):
            if _TORCHAUDIO_IO["save"] is None:
                _TORCHAUDIO_IO["save"] = False
    from scipy.io import wavfile

    a = audio.numpy().astype(np.float32, copy=False)
    wavfile.write(str(path), fs, a[0] if a.shape[0] == 1 else np.ascontiguousarray(a.T))


def resample(audio, fs, target_fs):
    """reference enhance.py:77-80"""
    if fs != target_fs:
        import torchaudio

        audio = torchaudio.functional.resample(audio, fs, target_fs)
    return audio


def output_path_for(path, rel_path, dir_proc, output):
    """reference enhance.py:173-181"""
    output = Path(output)
    if dir_proc:
        return output / Path(path).relative_to(rel_path)
    return output / Path(path).name if output.is_dir() else output


def enhance_files(model, files, out_paths, enhance_kwargs, batch_size=64, device=None, rank=0, world=1, progress=False,
                  window=None, io_threads=4, ragged=True):
    """Enhances this rank's share of `files` into `out_paths`.  Returns a dict of counters.
    Files are dealt to ranks whole (largest first, round robin: balanced without any exchange); every channel of a
    file is one utterance; utterances of similar (ragged=False: identical) resampled length share a batch.

    The files stream through in windows of `window` files: a pool of `io_threads` threads decodes window k + 1 while
    the device works on window k, and finished files are copied back and written by writer threads, so neither the
    host nor the device ever holds more than about two windows and file IO overlaps the device work (the reference
    is strictly serial: load, enhance, save, one file at a time; enhance.py:168-192).  Batches are formed inside a
    window, in ascending length."""
    from concurrent.futures import ThreadPoolExecutor

    device = device if device is not None else next(model.parameters()).device
    if window is None:
        window = max(64, batch_size)   # small enough that decoding window k + 1 overlaps the device work on window k
    t0 = time.perf_counter()
    order = sorted(range(len(files)), key=lambda i: (-os.path.getsize(files[i]), str(files[i])))
    mine = [fi for k, fi in enumerate(order) if k % world == rank]
    windows = [mine[k : k + window] for k in range(0, len(mine), window)]
    loader = ThreadPoolExecutor(max(1, io_threads))
    prefetch = ThreadPoolExecutor(1)   # runs load_window (which fans out over `loader`) one window ahead
    writer = ThreadPoolExecutor(max(2, io_threads))
    stats = dict(files=0, utterances=0, batches=0, audio_seconds=0.0)

    use_cuda = torch.device(device).type == "cuda"
    copy_stream = torch.cuda.Stream(device) if use_cuda else None
    down_stream = torch.cuda.Stream(device) if use_cuda else None
    stage = {}   # slot -> (pinned host buffer, event behind its last upload)
    out_stage = {}   # slot -> (pinned host buffer, future of the writer that reads it)
    native_io = os.path.exists(_native.LIB_PATH) and os.environ.get("UV_NATIVE_WAV", "1") != "0"

    def load_window(wi):
        """Decodes window wi and uploads it: -> ([(file index, audio on the device [C, n], fs)], (flat device buffer,
        event) or None).  WAV files are decoded by the native batch reader (csrc/wav_io.cuh: host threads, no
        interpreter lock) straight into the window's staging buffer; other formats go through load_audio.  On a CUDA
        device the buffer is pinned and the whole window goes up as ONE copy on a copy stream (this runs in the
        prefetch thread, a window ahead): a per-file `.to(device)` from pageable memory is synchronous -- it made the
        main thread wait for the device to finish the previous window before it could prepare the next one."""
        w = windows[wi]
        paths = [files[fi] for fi in w]
        meta = [None] * len(w)
        info = None
        if native_io:
            wavs = [k for k, p in enumerate(paths) if Path(p).suffix.lower() == ".wav"]
            if wavs:
                info, probed = _native.wav_probe([paths[k] for k in wavs], io_threads)
                for j, k in enumerate(wavs):
                    meta[k] = probed[j]
        other = {k: load_audio(paths[k]) for k in range(len(w)) if meta[k] is None}
        shapes = [(meta[k][0], meta[k][1]) if meta[k] is not None else tuple(other[k][0].shape) for k in range(len(w))]
        rates = [meta[k][2] if meta[k] is not None else other[k][1] for k in range(len(w))]
        offs = [0]
        for c, n in shapes:
            offs.append(offs[-1] + c * n)
        total = offs[-1]
        slot = wi % 3
        buf, ev = stage.get(slot, (None, None))
        if ev is not None:
            ev.synchronize()
        if buf is None or buf.numel() < total:
            buf = torch.empty(max(total, 1), dtype=torch.float32)
            if use_cuda:
                buf = buf.pin_memory()
        if info is not None:
            ok = _native.wav_read([paths[k] for k in wavs], info, [offs[k] for k in wavs], buf, io_threads)
            for j, k in enumerate(wavs):
                if meta[k] is not None and not ok[j]:
                    raise RuntimeError(f"{paths[k]}: read error")
        for k, (a, _) in other.items():
            buf[offs[k] : offs[k + 1]].copy_(a.reshape(-1))
        if not use_cuda:
            stage[slot] = (None, None)   # the views below alias the buffer: a fresh one per window
            return [(fi, buf[offs[k] : offs[k + 1]].view(shapes[k]), rates[k]) for k, fi in enumerate(w)], None
        with torch.cuda.device(device), torch.cuda.stream(copy_stream):
            dev = torch.empty(max(total, 1), dtype=torch.float32, device=device)
            dev[:total].copy_(buf[:total], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        stage[slot] = (buf, ev)
        return [(fi, dev[offs[k] : offs[k + 1]].view(shapes[k]), rates[k]) for k, fi in enumerate(w)], (dev, ev)

    def write_window(items, flat, ev):
        """items: [(path, (channels, frames), fs, offset)] into the flat host tensor `flat` (valid once ev is done)."""
        if ev is not None:
            ev.synchronize()
        for d in {Path(p).parent for p, _, _, _ in items}:
            d.mkdir(exist_ok=True, parents=True)
        wavs = [it for it in items if native_io and Path(it[0]).suffix.lower() == ".wav"]
        if wavs:
            ok = _native.wav_write([it[0] for it in wavs], [it[1] for it in wavs], [it[2] for it in wavs], [it[3] for it in wavs],
                                   flat, io_threads)
            if not all(ok):
                raise RuntimeError(f"{wavs[ok.index(False)][0]}: write error")
        for path, shape, fs, off in items:
            if not (native_io and Path(path).suffix.lower() == ".wav"):
                save_audio(path, flat[off : off + shape[0] * shape[1]].view(shape).clone(), fs)

    if progress:
        from tqdm import tqdm

        bar = tqdm(total=len(mine))
    pending = []
    nxt = prefetch.submit(load_window, 0) if windows else None
    with torch.no_grad():
        for wi in range(len(windows)):
            loaded, h2d = nxt.result()
            nxt = prefetch.submit(load_window, wi + 1) if wi + 1 < len(windows) else None
            if h2d is not None:   # the upload ran on the copy stream: order it in front of this stream's work
                torch.cuda.current_stream(device).wait_event(h2d[1])
                h2d[0].record_stream(torch.cuda.current_stream(device))
            utts = []        # (file index, channel, tensor [T] on the device at the model rate)
            fs_of = {}
            for fi, audio, fs in loaded:
                fs_of[fi] = fs
                a = resample(audio, fs, model.fs)
                for c in range(a.shape[0]):
                    utts.append((fi, c, a[c]))
            lens = [u[2].shape[-1] for u in utts]
            if ragged:
                batches = sharding.ragged_batches(lens, list(range(len(utts))), batch_size)
            else:
                batches = [(b, lens[b[0]]) for b in sharding.length_buckets(lens, list(range(len(utts))), batch_size)]
            done = {}
            for batch, T in batches:
                if all(lens[i] == T for i in batch):
                    y = model.enhance(torch.stack([utts[i][2] for i in batch]), **enhance_kwargs)
                else:   # rows of different lengths: each keeps its own geometry (csrc RowGeo)
                    x = torch.zeros((len(batch), T), dtype=utts[batch[0]][2].dtype, device=device)
                    for k, i in enumerate(batch):
                        x[k, : lens[i]] = utts[i][2]
                    y = model.enhance(x, lengths=[lens[i] for i in batch], **enhance_kwargs)
                for k, i in enumerate(batch):
                    done[i] = y[k, : lens[i]]
                stats["audio_seconds"] += sum(lens[i] for i in batch) / model.fs
            by_file = {}
            for i, (fi, c, _) in enumerate(utts):
                by_file.setdefault(fi, []).append(i)
            # the window's results go down as ONE copy into a pinned buffer on a copy stream; a writer thread encodes
            # the files from it (native batch writer for WAV) once the copy has landed
            enh = {fi: resample(torch.stack([done[i] for i in idx]), model.fs, fs_of[fi]) for fi, idx in by_file.items()}
            items, off = [], 0
            for fi, e in enh.items():
                items.append((str(out_paths[fi]), tuple(e.shape), fs_of[fi], off))
                off += e.numel()
            if items:
                flat_dev = torch.cat([e.reshape(-1).float() for e in enh.values()])
                slot = wi % 3
                hbuf, fut = out_stage.get(slot, (None, None))
                if fut is not None:
                    fut.result()
                if hbuf is None or hbuf.numel() < off:
                    hbuf = torch.empty(max(off, 1), dtype=torch.float32)
                    if use_cuda:
                        hbuf = hbuf.pin_memory()
                ev = None
                if use_cuda:
                    main = torch.cuda.current_stream(device)
                    ready = torch.cuda.Event()
                    ready.record(main)
                    with torch.cuda.stream(down_stream):
                        down_stream.wait_event(ready)
                        hbuf[:off].copy_(flat_dev, non_blocking=True)
                        ev = torch.cuda.Event()
                        ev.record(down_stream)
                    flat_dev.record_stream(down_stream)
                else:
                    hbuf[:off].copy_(flat_dev)
                fut = writer.submit(write_window, items, hbuf, ev)
                out_stage[slot] = (hbuf, fut)
                pending.append(fut)
            stats["files"] += len(by_file)
            stats["utterances"] += len(done)
            stats["batches"] += len(batches)
            if progress:
                bar.update(len(by_file))
            # bound what is in flight: at most the previous windows may still be waiting for their writer
            while len(pending) > 2:
                pending.pop(0).result()
    for f in pending:
        f.result()
    prefetch.shutdown()
    loader.shutdown()
    writer.shutdown()
    stats["seconds"] = time.perf_counter() - t0
    return stats


def main(argv=None):
    if argv is not None:
        sys.argv = [sys.argv[0]] + list(argv)
    parser = argparse.ArgumentParser(description="Enhance a file or a directory of audio files")
    parser.add_argument("input", type=Path, help="Path to an audio file or a folder of audio files")
    parser.add_argument("output", type=Path,
                        help="Output path for the enhanced files. In the case of a folder, the orignal structure is retained.")
    parser.add_argument("--model", type=str, default="line-corporation/open-universe:plusplus",
                        help="A checkpoint path or a model stored in the model zoo")
    parser.add_argument("--hf-token", type=str, help="Huggingface access token")
    parser.add_argument("--model-strict", action="store_true", help="Use strict policy to load the model.")
    parser.add_argument("--seed", type=int, default=1028282, help="Set a deterministic seed to get reproducible results")
    parser.add_argument("--device", type=str, default="cuda:0", help="cuda:X (there is no CPU path in this implementation)")
    parser.add_argument("--batch-size", type=int, default=64, help="Utterances of equal length enhanced together")
    parser.add_argument("--equal-length-batches", action="store_true", help="Only batch utterances of identical length")
    parser.add_argument("--window", type=int, default=None, help="Files decoded ahead and batched together (default: max(64, batch size))")
    parser.add_argument("--io-threads", type=int, default=4, help="Decoder threads")
    parser.add_argument("--precision", type=str, default=None, choices=["bf16", "fp32"],
                        help="bf16 (default): tensor-core tier, rel-L2 <= 1e-2 against the reference; fp32: parity tier, <= 1e-4, ~4.5x slower")
    requires_help = handle_help()
    args, _ = parser.parse_known_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    device = args.device
    if world > 1:
        device = f"cuda:{int(os.environ.get('LOCAL_RANK', '0'))}"
    if not device.startswith("cuda"):
        raise ValueError(f"open_universe_b200 runs on a CUDA sm_100 device only; provided {device}")
    if not torch.cuda.is_available():
        raise RuntimeError("no CUDA device: open_universe_b200 has no CPU fallback")

    model = inference_utils.load_model(args.model, device=device, strict=args.model_strict, hf_token=args.hf_token)
    if args.precision is not None:
        model.precision = args.precision
    rng = torch.Generator(device=device)
    rng.manual_seed(args.seed + rank)
    inference_utils.add_enhance_arguments(model, parser)
    if requires_help:
        sys.argv.append("--help")
    args = parser.parse_args()
    enhance_kwargs = {}
    for group in parser._action_groups:
        if group.title == "enhance":
            enhance_kwargs = {a.dest: getattr(args, a.dest, None) for a in group._group_actions}
    enhance_kwargs["rng"] = rng

    files, rel_path, dir_proc = find_files(args.input)
    outs = [output_path_for(p, rel_path, dir_proc, args.output) for p in files]
    stats = enhance_files(model, files, outs, enhance_kwargs, batch_size=args.batch_size, device=device, rank=rank,
                          world=world, progress=dir_proc and rank == 0, window=args.window, io_threads=args.io_threads,
                          ragged=not args.equal_length_batches)
    print(f"[rank {rank}] {stats['files']} files, {stats['utterances']} utterances in {stats['batches']} batches, "
          f"{stats['audio_seconds']:.1f} s of audio in {stats['seconds']:.2f} s "
          f"({stats['audio_seconds'] / max(stats['seconds'], 1e-9):.1f} audio-s/s incl. file IO and plan creation)")
    return stats


if __name__ == "__main__":
    main()
