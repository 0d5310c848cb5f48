"""The file pipeline around enhance() (open_universe_b200/bin/enhance.py; reference open_universe/bin/enhance.py):
file discovery, output layout, WAV IO, equal-length batching, per-rank dealing on the CPU with a stand-in model;
end to end on the GPU against direct model.enhance calls with the same generator."""
import json
import os

import numpy as np
import pytest
import torch
import yaml
from scipy.io import wavfile

from conftest import GOLDEN, build_model
from open_universe_b200 import sharding
from open_universe_b200.bin import enhance as cli


def _write_wavs(root, specs):
    """specs: list of (relative path, fs, channels, samples, dtype)"""
    g = np.random.default_rng(0)
    paths = []
    for rel, fs, ch, n, dt in specs:
        p = root / rel
        p.parent.mkdir(parents=True, exist_ok=True)
        x = 0.1 * g.standard_normal((n, ch)).astype(np.float32)
        if dt == "int16":
            x = np.round(x * 32767).astype(np.int16)
        wavfile.write(str(p), fs, x[:, 0] if ch == 1 else x)
        paths.append(p)
    return paths


class _HalfModel(torch.nn.Module):
    """Stand-in with the attributes the pipeline reads: halves the signal, records the batches it saw."""

    fs = 16000

    def __init__(self):
        super().__init__()
        self.w = torch.nn.Parameter(torch.zeros(1))
        self.batches = []

    def enhance(self, mix, **kw):
        self.batches.append(tuple(mix.shape))
        self.lengths = getattr(self, "lengths", []) + [kw.get("lengths")]
        return 0.5 * mix


SPECS = [("a.wav", 16000, 1, 4000, "float32"), ("sub/b.wav", 16000, 1, 4000, "int16"), ("sub/c.wav", 16000, 2, 4000, "float32"),
         ("sub/deep/d.wav", 8000, 1, 3000, "float32"), ("e.wav", 16000, 1, 5000, "float32"), ("notes.txt", 0, 0, 0, "")]


def _make_tree(tmp_path):
    src = tmp_path / "noisy"
    _write_wavs(src, [s for s in SPECS if s[0].endswith(".wav")])
    (src / "notes.txt").write_text("not audio")
    return src


def test_find_files_and_output_layout(tmp_path):
    src = _make_tree(tmp_path)
    files, rel, is_dir = cli.find_files(src)
    assert is_dir and rel == src
    assert sorted(str(p.relative_to(src)) for p in files) == ["a.wav", "e.wav", "sub/b.wav", "sub/c.wav", "sub/deep/d.wav"]
    out = tmp_path / "enh"
    assert cli.output_path_for(src / "sub" / "b.wav", rel, True, out) == out / "sub" / "b.wav"   # structure retained
    f1, rel1, d1 = cli.find_files(src / "a.wav")
    assert f1 == [src / "a.wav"] and not d1
    assert cli.output_path_for(f1[0], rel1, False, tmp_path / "x.wav") == tmp_path / "x.wav"
    (tmp_path / "outdir").mkdir()
    assert cli.output_path_for(f1[0], rel1, False, tmp_path / "outdir") == tmp_path / "outdir" / "a.wav"


def test_wav_io_roundtrip(tmp_path):
    (p16, pf) = _write_wavs(tmp_path, [("i16.wav", 16000, 2, 100, "int16"), ("f32.wav", 24000, 1, 77, "float32")])
    a, fs = cli.load_audio(p16)
    assert a.shape == (2, 100) and fs == 16000 and a.dtype == torch.float32 and a.abs().max() <= 1.0
    b, fs = cli.load_audio(pf)
    assert b.shape == (1, 77) and fs == 24000
    cli.save_audio(tmp_path / "o.wav", b, fs)
    c, fs2 = cli.load_audio(tmp_path / "o.wav")
    assert fs2 == fs and torch.equal(b, c)


@pytest.mark.parametrize("world", [1, 2])
def test_enhance_files_batches_equal_lengths_and_deals_whole_files(tmp_path, world):
    src = _make_tree(tmp_path)
    files, rel, is_dir = cli.find_files(src)
    out = tmp_path / "enh"
    outs = [cli.output_path_for(p, rel, is_dir, out) for p in files]
    seen, n_files = [], 0
    for rank in range(world):
        model = _HalfModel()
        st = cli.enhance_files(model, files, outs, {}, batch_size=3, device="cpu", rank=rank, world=world, ragged=False)
        n_files += st["files"]
        seen += model.batches
        for B, T in model.batches:
            assert B <= 3
    assert n_files == 5
    # 16 kHz: four utterances of 4000 samples (a, b, both channels of c), one of 5000; the 8 kHz file becomes 6000
    assert sorted(T for B, T in seen for _ in range(B)) == [4000, 4000, 4000, 4000, 5000, 6000]
    for p, o in zip(files, outs):
        x, fs = cli.load_audio(p)
        y, fs_o = cli.load_audio(o)
        assert fs_o == fs and y.shape == x.shape                       # original rate and channel count restored
        if fs == 16000:
            assert torch.allclose(y, 0.5 * x, atol=1e-6)


def test_enhance_files_ragged_batches_carry_row_lengths(tmp_path):
    """Default schedule: utterances of similar length share a batch, every row with its own length (enhance(lengths=))."""
    src = tmp_path / "noisy"
    specs = [(f"f{k:02d}.wav", 16000, 1, 9000 + 37 * k, "float32") for k in range(20)] + [("short.wav", 16000, 1, 900, "float32")]
    files = _write_wavs(src, specs)
    outs = [tmp_path / "out" / p.name for p in files]
    model = _HalfModel()
    st = cli.enhance_files(model, files, outs, {}, batch_size=8, device="cpu")
    assert st["files"] == 21 and st["utterances"] == 21 and st["batches"] == 4          # 8 + 8 + 4 similar ones, the short one alone
    for (B, T), lens in zip(model.batches, model.lengths):
        assert B <= 8 and T % 8000 == 0 or lens is None
        if lens is not None:
            assert len(lens) == B and max(lens) <= T and min(lens) >= 0.75 * max(lens)
    for p, o in zip(files, outs):
        assert torch.allclose(cli.load_audio(o)[0], 0.5 * cli.load_audio(p)[0], atol=1e-6)


def test_ragged_batches_bound_the_wasted_tail():
    lengths = [16000, 15000, 9000, 8900, 8800, 2000, 16000, 100]
    batches = sharding.ragged_batches(lengths, list(range(8)), 3, quantum=4000, max_waste=0.25)
    assert sorted(i for b, _ in batches for i in b) == list(range(8))
    for b, T in batches:
        ls = [lengths[i] for i in b]
        assert len(b) <= 3 and T % 4000 == 0 and T >= max(ls) and T - max(ls) < 4000 and min(ls) >= 0.75 * max(ls)


def test_length_buckets_never_mix_lengths():
    lengths = [5, 7, 5, 5, 7, 9, 5]
    batches = sharding.length_buckets(lengths, list(range(7)), 2)
    assert sorted(i for b in batches for i in b) == list(range(7))
    for b in batches:
        assert len(b) <= 2 and len({lengths[i] for i in b}) == 1


@pytest.mark.gpu
def test_cli_end_to_end_equals_direct_enhance_calls(tmp_path):
    """python -m open_universe_b200.bin.enhance on a small tree == model.enhance on the same equal-length batches,
    in the same order, with the same generator (the native path itself is parity-tested in test_gpu_parity.py)."""
    model, cfg = build_model("universepp_16k")
    d = tmp_path / "exp" / "checkpoints"
    d.mkdir(parents=True)
    torch.save({"state_dict": {k: v.clone() for k, v in model.state_dict().items()}}, d / "last.ckpt")
    (tmp_path / "exp" / ".hydra").mkdir()
    with open(tmp_path / "exp" / ".hydra" / "config.yaml", "w") as f:
        yaml.safe_dump(json.loads(json.dumps(cfg)), f)
    src = tmp_path / "noisy"
    specs = [("a.wav", 16000, 1, 3000, "float32"), ("s/b.wav", 16000, 1, 3000, "float32"), ("s/c.wav", 16000, 1, 2000, "float32")]
    _write_wavs(src, specs)
    out = tmp_path / "enh"
    st = cli.main([str(src), str(out), "--model", str(d / "last.ckpt"), "--seed", "5", "--n_steps", "4", "--batch-size", "8",
                   "--equal-length-batches"])
    assert st["files"] == 3 and st["batches"] == 2
    gm = model.to("cuda").eval()
    rng = torch.Generator(device="cuda").manual_seed(5)
    files = {rel: cli.load_audio(src / rel)[0].cuda() for rel, *_ in specs}
    # batches in ascending length; inside a batch: files largest first, then by path (the dealing order)
    ref_c = gm.enhance(files["s/c.wav"], n_steps=4, rng=rng)
    ref_ab = gm.enhance(torch.cat([files["a.wav"], files["s/b.wav"]]), n_steps=4, rng=rng)
    got = {rel: cli.load_audio(out / rel)[0] for rel, *_ in specs}
    assert torch.allclose(got["s/c.wav"], ref_c.cpu(), atol=1e-6)
    assert torch.allclose(got["a.wav"], ref_ab[0:1].cpu(), atol=1e-6)
    assert torch.allclose(got["s/b.wav"], ref_ab[1:2].cpu(), atol=1e-6)


def test_windows_stream_every_file_once_with_bounded_batches(tmp_path):
    """The streaming schedule (windows decoded ahead, writer threads): every file written exactly once, batches never
    mix lengths or cross a window, results independent of the window size."""
    src = tmp_path / "noisy"
    specs = [(f"f{k:03d}.wav", 16000, 1, 2000 + 100 * (k % 7), "float32") for k in range(37)]
    files = _write_wavs(src, specs)
    for window in (5, 64):
        outs = [tmp_path / f"out{window}" / p.name for p in files]
        model = _HalfModel()
        st = cli.enhance_files(model, files, outs, {}, batch_size=4, device="cpu", window=window, io_threads=3, ragged=False)
        assert st["files"] == 37 and st["utterances"] == 37
        assert all(B <= 4 for B, T in model.batches)
        assert sum(B for B, T in model.batches) == 37
        for p, o in zip(files, outs):
            assert torch.allclose(cli.load_audio(o)[0], 0.5 * cli.load_audio(p)[0], atol=1e-6)


@pytest.mark.gpu
def test_cli_300_files_of_300_distinct_lengths(tmp_path):
    """A corpus where every file has its own length (the reference loops file by file, enhance.py:168-192): one plan
    per length goes through the bounded plan cache without running out of memory, and every output equals the direct
    one-by-one model.enhance call with the same generator state."""
    model, cfg = build_model("universepp_16k")
    d = tmp_path / "exp" / "checkpoints"
    d.mkdir(parents=True)
    torch.save({"state_dict": {k: v.clone() for k, v in model.state_dict().items()}}, d / "last.ckpt")
    (tmp_path / "exp" / ".hydra").mkdir()
    with open(tmp_path / "exp" / ".hydra" / "config.yaml", "w") as f:
        yaml.safe_dump(json.loads(json.dumps(cfg)), f)
    src = tmp_path / "noisy"
    specs = [(f"d{k % 5}/f{k:03d}.wav", 16000, 1, 1600 + 53 * k, "float32") for k in range(300)]
    files = _write_wavs(src, specs)
    out = tmp_path / "enh"
    st = cli.main([str(src), str(out), "--model", str(d / "last.ckpt"), "--seed", "9", "--n_steps", "2", "--batch-size", "8",
                   "--window", "64", "--equal-length-batches"])
    assert st["files"] == 300 and st["batches"] == 300
    gm = model.to("cuda").eval()
    rng = torch.Generator(device="cuda").manual_seed(9)
    found, _, _ = cli.find_files(src)
    order = sorted(range(len(found)), key=lambda i: (-os.path.getsize(found[i]), str(found[i])))
    worst = 0.0
    for w0 in range(0, 300, 64):
        win = order[w0 : w0 + 64]
        for fi in sorted(win, key=lambda i: os.path.getsize(found[i])):   # batches of a window in ascending length
            x = cli.load_audio(found[fi])[0].cuda()
            ref = gm.enhance(x, n_steps=2, rng=rng).cpu()
            got = cli.load_audio(out / found[fi].relative_to(src))[0]
            worst = max(worst, (got - ref).abs().max().item())
    assert len(gm._plans) <= gm.plan_cache_max
    assert worst < 1e-6, worst


@pytest.mark.gpu
def test_cli_ragged_corpus_equals_direct_ragged_calls(tmp_path):
    """The default schedule on a corpus where every file has its own length: a handful of batches instead of one per
    file, each output equal to model.enhance(x, lengths=...) on the same batch with the same generator state (the
    per-row geometry itself is tested against file-by-file runs in test_gpu_ragged.py)."""
    model, cfg = build_model("universepp_16k")
    d = tmp_path / "exp" / "checkpoints"
    d.mkdir(parents=True)
    torch.save({"state_dict": {k: v.clone() for k, v in model.state_dict().items()}}, d / "last.ckpt")
    (tmp_path / "exp" / ".hydra").mkdir()
    with open(tmp_path / "exp" / ".hydra" / "config.yaml", "w") as f:
        yaml.safe_dump(json.loads(json.dumps(cfg)), f)
    src = tmp_path / "noisy"
    specs = [(f"f{k:03d}.wav", 16000, 1, 6000 + 61 * k, "float32") for k in range(40)]
    files = _write_wavs(src, specs)
    out = tmp_path / "enh"
    st = cli.main([str(src), str(out), "--model", str(d / "last.ckpt"), "--seed", "9", "--n_steps", "2", "--batch-size", "16",
                   "--window", "64"])
    assert st["files"] == 40 and st["batches"] <= 4
    gm = model.to("cuda").eval()
    rng = torch.Generator(device="cuda").manual_seed(9)
    found, _, _ = cli.find_files(src)
    order = sorted(range(len(found)), key=lambda i: (-os.path.getsize(found[i]), str(found[i])))
    audio = [cli.load_audio(found[fi])[0][0] for fi in order]          # utterances in the window's order
    lens = [a.shape[-1] for a in audio]
    worst = 0.0
    for batch, T in sharding.ragged_batches(lens, list(range(40)), 16):
        x = torch.zeros(len(batch), T)
        for k, i in enumerate(batch):
            x[k, : lens[i]] = audio[i]
        ref = gm.enhance(x.cuda(), n_steps=2, rng=rng, lengths=[lens[i] for i in batch]).cpu()
        for k, i in enumerate(batch):
            got = cli.load_audio(out / found[order[i]].relative_to(src))[0][0]
            worst = max(worst, (got - ref[k, : lens[i]]).abs().max().item())
            assert ref[k, lens[i]:].abs().max() == 0 if lens[i] < T else True
    assert worst < 1e-6, worst


def test_native_batch_wav_io_matches_the_python_path(tmp_path):
    """uv_wav_probe / uv_wav_read / uv_wav_write (csrc/wav_io.cuh) against scipy + load_audio: int16 / int32 / uint8 /
    float32, mono and stereo, odd chunk sizes; unsupported files are reported, not mangled."""
    import numpy as np
    from scipy.io import wavfile

    from open_universe_b200 import _native
    from open_universe_b200.bin import enhance as cli

    g = np.random.default_rng(0)
    specs = [("f32m", np.float32, 1, 4001), ("f32s", np.float32, 2, 3000), ("i16m", np.int16, 1, 70001), ("i16s", np.int16, 2, 5),
             ("i32m", np.int32, 1, 777), ("u8s", np.uint8, 3, 1234)]
    paths = []
    for name, dt, ch, n in specs:
        x = g.uniform(-0.9, 0.9, size=(n, ch))
        if dt == np.int16:
            d = (x * 32767).astype(np.int16)
        elif dt == np.int32:
            d = (x * 2147483647).astype(np.int32)
        elif dt == np.uint8:
            d = (x * 127 + 128).astype(np.uint8)
        else:
            d = x.astype(np.float32)
        p = tmp_path / f"{name}.wav"
        wavfile.write(str(p), 16000 + 1000 * ch, d[:, 0] if ch == 1 else d)
        paths.append(p)
    bad = tmp_path / "bad.wav"
    bad.write_bytes(b"RIFFxxxxWAVEjunk")
    f64 = tmp_path / "f64.wav"
    wavfile.write(str(f64), 8000, g.uniform(-1, 1, 100))   # float64: not decoded natively
    info, probed = _native.wav_probe(paths + [bad, f64], 3)
    assert probed[-1] is None and probed[-2] is None
    assert [p[:2] for p in probed[:-2]] == [(ch, n) for _, _, ch, n in specs]
    offs = [0]
    for ch, n, _ in probed[:-2]:
        offs.append(offs[-1] + ch * n)
    buf = torch.full((offs[-1] + 7,), 9.0)
    ok = _native.wav_read(paths + [bad, f64], info, offs[:-1] + [offs[-1], offs[-1]], buf, 3)
    assert ok[: len(specs)] == [True] * len(specs) and ok[-2:] == [False, False]
    assert (buf[offs[-1]:] == 9.0).all()
    cli._TORCHAUDIO_IO["load"] = None
    for k, p in enumerate(paths):
        ref, fs = cli.load_audio(p)
        got = buf[offs[k] : offs[k + 1]].view(ref.shape)
        assert fs == probed[k][2] and torch.equal(got, ref), p
    # write: planar [C][n] out of one flat buffer -> files scipy reads back bit-exactly
    outs = [tmp_path / "o" / f"{name}.wav" for name, *_ in specs]
    (tmp_path / "o").mkdir()
    shapes = [(ch, n) for _, _, ch, n in specs]
    okw = _native.wav_write(outs, shapes, [p[2] for p in probed[:-2]], offs[:-1], buf, 2)
    assert all(okw)
    for k, o in enumerate(outs):
        fs, d = wavfile.read(str(o))
        assert d.dtype == np.float32 and fs == probed[k][2]
        d = d[:, None] if d.ndim == 1 else d
        assert np.array_equal(d.T, buf[offs[k] : offs[k + 1]].view(shapes[k]).numpy())
    assert _native.wav_write([tmp_path / "nodir" / "x.wav"], [(1, 4)], [8000], [0], buf, 1) == [False]


# ---------------------------------------------------------------- the reference's own file flow (oracle/gen_golden_cli.py)
def _flow_fixture():
    return torch.load(os.path.join(GOLDEN, "cli_flow.pt"), weights_only=False)


def test_oracle_reproduces_the_reference_file_flow():
    """load -> resample -> enhance -> resample of the REAL reference on three files (mono / stereo / 22.05 kHz),
    recorded by oracle/gen_golden_cli.py, against the CPU oracle run through the same steps (bin/enhance.py:77-80,183-190)."""
    from open_universe_b200.bin import enhance as cli
    from oracle import universe_oracle as uo

    fix = _flow_fixture()
    model, cfg = build_model(fix["model"], fix["seed_weights"])
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    for f in fix["files"]:
        x = cli.resample(f["audio"], f["fs"], model.fs)
        assert x.shape[-1] == f["model_rate_len"]
        enh = uo.enhance(sd, cfg.model, x, [z[:, None, :] for z in f["noise"]])
        out = cli.resample(enh, model.fs, f["fs"])
        assert out.shape == f["out"].shape
        assert uo.rel_l2(out, f["out"]) < 1e-5, f["name"]


class _ReplayModel:
    """Stand-in handed to enhance_files: forwards to the native model with the fixture's draws, matched to the rows of
    whatever batch the schedule forms (the CLI draws its own noise per batch; the fixture's draws are per file)."""

    def __init__(self, model, rows):
        self.model, self.rows, self.fs, self.calls = model, rows, model.fs, 0

    def parameters(self):
        return self.model.parameters()

    def enhance(self, x, lengths=None, **kw):
        B, T = x.shape
        lengths = list(lengths) if lengths is not None else [T] * B
        td = self.model.tot_ds
        Lp = T + (td - T % td)
        n_steps = self.rows[0][1].shape[0]
        noise = torch.zeros(n_steps, B, Lp)
        for b in range(B):
            head = x[b, :16].cpu()
            k = min(range(len(self.rows)), key=lambda i: (self.rows[i][0][:16] - head).abs().sum().item())
            z = self.rows[k][1]
            assert z.shape[-1] == lengths[b] + (td - lengths[b] % td)
            noise[:, b, : z.shape[-1]] = z
        self.calls += 1
        out, _ = self.model.enhance_debug(x, noise.to(x.device), lengths=None if all(v == T for v in lengths) else lengths)
        return out


@pytest.mark.gpu
@pytest.mark.parametrize("precision,tol", [("bf16", 1e-2), ("fp32", 1e-4)])
def test_cli_reproduces_the_reference_file_flow(tmp_path, precision, tol):
    """The three files of the reference-flow fixture as WAV files through enhance_files (native WAV IO, one ragged batch
    with per-row lengths, resampling in and out) against what the reference's file-by-file loop produced."""
    import numpy as np
    from scipy.io import wavfile

    from open_universe_b200.bin import enhance as cli
    from oracle import universe_oracle as uo

    fix = _flow_fixture()
    model, cfg = build_model(fix["model"], fix["seed_weights"])
    model = model.to("cuda").eval()
    model.precision = precision
    src, dst = tmp_path / "in", tmp_path / "out"
    rows = []
    for f in fix["files"]:
        p = src / f["name"]
        p.parent.mkdir(parents=True, exist_ok=True)
        a = f["audio"].numpy().astype(np.float32)
        wavfile.write(str(p), f["fs"], a[0] if a.shape[0] == 1 else np.ascontiguousarray(a.T))
        x = cli.resample(f["audio"], f["fs"], model.fs)
        for c in range(x.shape[0]):
            rows.append((x[c], f["noise"][:, c]))
    files, rel, is_dir = cli.find_files(src)
    outs = [cli.output_path_for(p, rel, is_dir, dst) for p in files]
    proxy = _ReplayModel(model, rows)
    st = cli.enhance_files(proxy, files, outs, {}, batch_size=8, device="cuda:0")
    assert st["files"] == 3 and st["utterances"] == 4 and proxy.calls == st["batches"]
    for f in fix["files"]:
        fs, d = wavfile.read(str(dst / f["name"]))
        got = torch.from_numpy(np.ascontiguousarray((d[:, None] if d.ndim == 1 else d).T))
        assert fs == f["fs"] and got.shape == f["out"].shape
        err, sdr = uo.rel_l2(got, f["out"]), uo.si_sdr(got, f["out"]).min().item()
        print(f"{precision} CLI vs reference file flow, {f['name']}: rel-L2 {err:.2e}, SI-SDR {sdr:.1f} dB")
        assert err < tol and sdr > 40.0, f["name"]
