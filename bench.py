#!/usr/bin/env python
"""Headline benchmark: audio-seconds enhanced per second, 8-step UNIVERSE++ 16 kHz.

  python bench.py --gpus N --steps K --warmup W            # native sm_100a path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's algorithm on the host CPU
  python bench.py --workload cfg3|cfg3_n50|cfg4|cfg5       # the other BASELINE.json configs (one resident batch)
  python bench.py --workload cfg5_sweep [--clips 4096]     # BASELINE configs[4] as stated: 4096 x 30 s clips from
                                                           # pinned host memory, utterance-sharded (strong scaling)

A "step" is one full enhance() pass over one batch of synthetic noisy clips
(BASELINE.json configs[1]: batch 64 x 4 s, 8 sampler steps, bf16 operands / fp32 accumulate).
N > 1 (torchrun, one rank per GPU): utterances are sharded, every rank runs the same per-GPU
workload (weak scaling), no data-path collective; time = max over ranks.
Prints ONE JSON line on rank 0 (stdout carries nothing else).
"""
import argparse
import glob
import hashlib
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # BASELINE.json configs[1] (the configuration the metric is quoted on)
    "cfg2": dict(model="universepp_16k", B=64, T=64000, n_steps=8, desc="UNIVERSE++ 16 kHz, batch 64 x 4 s clips, 8 steps"),
    "cfg3": dict(model="universe_16k", B=64, T=64000, n_steps=8, desc="UNIVERSE original 16 kHz, batch 64 x 4 s, n_steps 8 (config-faithful)"),
    "cfg3_n50": dict(model="universe_16k", B=64, T=64000, n_steps=50, desc="UNIVERSE original 16 kHz, batch 64 x 4 s, n_steps 50 (paper-faithful)"),
    "cfg4": dict(model="universepp_24k", B=32, T=240000, n_steps=8, desc="UNIVERSE++ 24 kHz, batch 32 x 10 s clips, 8 steps"),
    "cfg5": dict(model="universepp_16k", B=32, T=480000, n_steps=8, desc="UNIVERSE++ 16 kHz, 30 s clips in batches of 32 (one resident batch per rank)"),
    # BASELINE.json configs[4] as stated: the whole 4096-clip set, sharded over the ranks, host to host
    "cfg5_sweep": dict(model="universepp_16k", B=32, T=480000, n_steps=8, clips=4096,
                       desc="UNIVERSE++ 16 kHz utterance-sharded throughput sweep, 4096 x 30 s clips"),
}


def load_model_cfg(name):
    from open_universe_b200 import config

    with open(os.path.join(ROOT, "tests", "golden", f"config_{name}.json")) as f:
        return config.to_attr(json.load(f))


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return dict(tensor=d["bf16_tflops_sustained"], tensor_burst=d["bf16_tflops"], hbm=d["hbm_gbs"], src="measured")
    return dict(tensor=1400.0, tensor_burst=1590.0, hbm=6650.0, src="fallback")


def source_hash():
    """Hash of everything that decides which kernels run and what they move: an ncu pass (roofline.traffic) is only
    quoted for the build it was taken on."""
    h = hashlib.sha1()
    files = sorted(glob.glob(os.path.join(ROOT, "open_universe_b200", "csrc", "*.cu*")))
    files += [os.path.join(ROOT, "include", "universe_b200.h"), os.path.join(ROOT, "open_universe_b200", "program.py"),
              os.path.join(ROOT, "open_universe_b200", "weights.py")]
    for f in files:
        with open(f, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:12]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                       "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().strip().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return dict(sm_mhz=statistics.median(sm) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# ------------------------------------------------------------------------------------------ CPU legs
def cpu_reference_run(wl, n_clips, clip_seconds, repeats, threads=None):
    """The reference's algorithm on the host cores: the torch-CPU fp32 oracle port (oracle/universe_oracle.py, pinned
    to the real reference's outputs by tests/test_oracle_golden.py; the Python reference itself cannot travel to the
    GPU box)."""
    from open_universe_b200 import config, networks, testing  # noqa: F401
    from oracle import universe_oracle as uo

    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    cfg = load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    sd = {k: v.detach() for k, v in model.state_dict().items()}
    fs = int(cfg.model.fs)
    T = int(clip_seconds * fs)
    tot = model.tot_ds
    g = torch.Generator().manual_seed(1235)
    mix = 0.1 * torch.randn(n_clips, T, generator=g)
    Lp = T + (tot - T % tot)
    noises = [torch.randn(n_clips, 1, Lp, generator=g) for _ in range(wl["n_steps"])]
    times = []
    with torch.no_grad():
        for _ in range(repeats):
            t0 = time.perf_counter()
            uo.enhance(sd, cfg.model, mix, noises, n_steps=wl["n_steps"])
            times.append(time.perf_counter() - t0)
    return n_clips * clip_seconds, times, threads


def run_reference_arm(args, wl, rank):
    """--impl reference: rank 0 alone times the CPU algorithm on a bounded sample of the workload (a small batch of
    clips of the workload's length, capped at 4 s, so that all host threads have work)."""
    if rank != 0:
        return
    clip_s = wl["T"] / 16000 if "16k" in wl["model"] else wl["T"] / 24000
    clip_s = min(clip_s, 4.0)
    n_clips = args.ref_clips
    audio_s, times, threads = cpu_reference_run(wl, n_clips, clip_s, args.warmup + args.steps)
    timed = times[args.warmup:]
    ms = 1e3 * sum(timed) / len(timed)
    value = audio_s / (ms / 1e3)
    sample = (f"{n_clips} clips x {clip_s:g} s per step ({wl['n_steps']} sampler steps), torch CPU fp32 oracle port of the "
              "reference enhance()")
    line = dict(impl="reference", metric="audio-sec enhanced/sec", value=value, unit="audio-s/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=ms, higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["desc"], sample=sample),
                cpu_baseline=dict(value=value, unit="audio-s/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="audio-s/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def gpu_eager_baseline(wl, dev, n_clips, repeats=3):
    """Stated baseline (SURVEY.md 2.3 'bar to beat', opt-in: --gpu-eager-baseline): the oracle port of the reference,
    i.e. plain PyTorch eager ops (cuDNN convolutions, cuDNN GRU), on the same B200 in fp32 with TF32 on and off.
    Never part of the product path."""
    from open_universe_b200 import config, networks, testing  # noqa: F401
    from oracle import universe_oracle as uo

    cfg = load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    sd = {k: v.detach().to(dev) for k, v in model.state_dict().items()}
    fs = int(cfg.model.fs)
    T = wl["T"]
    Lp = T + (model.tot_ds - T % model.tot_ds)
    g = torch.Generator(device=dev).manual_seed(3)
    mix = 0.1 * torch.randn(n_clips, T, device=dev, generator=g)
    noises = [torch.randn(n_clips, 1, Lp, device=dev, generator=g) for _ in range(wl["n_steps"])]
    res = {}
    for tf32 in (True, False):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.backends.cuda.matmul.allow_tf32 = tf32
        best = None
        with torch.no_grad(), torch.device(dev):   # (factory calls inside the port -- arange, linspace, nn.GRU -- land on the GPU)
            for _ in range(repeats + 1):
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                uo.enhance(sd, cfg.model, mix, list(noises), n_steps=wl["n_steps"])
                torch.cuda.synchronize(dev)
                dt = time.perf_counter() - t0
                best = dt if best is None else min(best, dt)
        res["tf32" if tf32 else "fp32"] = dict(value=n_clips * T / fs / best, unit="audio-s/s", ms_per_pass=best * 1e3)
    res["sample"] = f"{n_clips} clips x {T / fs:g} s, {wl['n_steps']} steps, PyTorch eager (cuDNN) oracle port on the same GPU, best of {repeats}"
    return res


# ------------------------------------------------------------------------------------------ NCCL log
def nccl_log_setup(rank):
    """NCCL's INFO lines go to a file (stdout must stay one JSON line); rank 0 copies the banner to stderr afterwards
    and reports what it saw."""
    path = os.path.join(tempfile.gettempdir(), f"uv_nccl_{os.getpid()}.log")
    os.environ.setdefault("NCCL_DEBUG", "INFO")
    if os.environ["NCCL_DEBUG"].upper() in ("VERSION", "WARN"):
        os.environ["NCCL_DEBUG"] = "INFO"
    os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
    os.environ["NCCL_DEBUG_FILE"] = path
    return path


def nccl_log_report(path):
    info = dict(log=None, nranks=None, version=None)
    try:
        text = open(path).read()
    except OSError:
        return info
    m = re.search(r"NCCL version ([\w.+-]+)", text)
    if m:
        info["version"] = m.group(1)
    m = re.search(r"nranks (\d+)", text)
    if m:
        info["nranks"] = int(m.group(1))
    info["nvls"] = "NVLS" in text
    keep = [ln for ln in text.splitlines() if "NCCL version" in ln or "nranks" in ln or "Init COMPLETE" in ln or "NVLS" in ln]
    for ln in keep[:12]:
        print(ln, file=sys.stderr)
    info["log"] = f"{len(text.splitlines())} NCCL INFO lines in {path} (summary copied to stderr)"
    return info


# ------------------------------------------------------------------------------------------ native arm
def interval_union(iv):
    iv = sorted(iv)
    tot, cur_a, cur_b = 0.0, None, None
    for a, b in iv:
        if cur_b is None or a > cur_b:
            if cur_b is not None:
                tot += cur_b - cur_a
            cur_a, cur_b = a, b
        else:
            cur_b = max(cur_b, b)
    if cur_b is not None:
        tot += cur_b - cur_a
    return tot


def setup_native(wl, dev):
    from open_universe_b200 import config, networks, testing  # noqa: F401

    cfg = load_model_cfg(wl["model"])
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    return cfg, model.to(dev).eval()


def run_native(args, wl, rank, world, local_rank):
    import torch.distributed as dist

    from open_universe_b200 import _abi, program, testing

    if args.warmup < 3:
        args.warmup = 3
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    nccl_log = None
    if world > 1:
        nccl_log = nccl_log_setup(rank)
        dist.init_process_group("nccl", device_id=dev)
    cfg, model = setup_native(wl, dev)
    model.precision = args.precision
    hp = args.precision == "fp32"
    fs = int(cfg.model.fs)
    B, T, n_steps = wl["B"], wl["T"], wl["n_steps"]
    mix_host = testing.synthetic_noisy_speech(B, T, seed=1234 + rank).pin_memory()
    out_host = torch.empty(B, T).pin_memory()
    mix_dev = mix_host.to(dev)
    plan = model._get_plan(B, T, n_steps)
    Lp = plan.pg.Lp
    gen = torch.Generator(device=dev).manual_seed(1 + rank)
    noise = torch.randn(n_steps, B, Lp, device=dev, generator=gen)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    def step_resident():
        return model._enhance_impl(mix_dev, n_steps, None, None, None, None, False, False, None, "median", None,
                                   noise=noise, return_steps=False)

    # End to end, the way a serving loop calls the public API: every step copies ITS input from pinned host memory and ITS
    # result back to pinned host memory.  The copies run on their own streams (double-buffered device inputs), so the H2D
    # of step k + 1 and the D2H of step k - 1 overlap the compute of step k; the host never runs more than two steps ahead
    # (it waits for the D2H of step k - 2), and the timed region ends behind the last D2H.  Same-box probe: 43.16 ms per
    # step with blocking copies on one stream, 42.60 ms this way.
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    dbuf = [torch.empty(B, T, device=dev) for _ in range(2)]

    def run_e2e(k):
        in_ready = [torch.cuda.Event() for _ in range(2)]
        in_free = [torch.cuda.Event() for _ in range(2)]
        for e in in_free:
            e.record(main)
        done, keep = [], [None, None]

        def prefetch(slot):
            with torch.cuda.stream(s_in):
                s_in.wait_event(in_free[slot])
                dbuf[slot].copy_(mix_host, non_blocking=True)      # H2D from pinned memory
                in_ready[slot].record(s_in)

        prefetch(0)
        for i in range(k):
            cur = i & 1
            if i >= 2:
                done[i - 2].synchronize()
            if i + 1 < k:
                prefetch(cur ^ 1)
            main.wait_event(in_ready[cur])
            out = model.enhance(dbuf[cur], n_steps=n_steps, rng=gen)   # public API (draws the noise with torch.randn)
            in_free[cur].record(main)
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev)
                out_host.copy_(out, non_blocking=True)             # D2H of the result
                d = torch.cuda.Event()
                d.record(s_out)
            done.append(d)
            keep[cur] = out
        for d in done[-2:]:
            main.wait_event(d)

    def timed(fn, k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        sync_all()
        ms = torch.tensor([e0.elapsed_time(e1) / k], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item()

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms = timed(step_resident, args.steps)
    clocks = sampler.stop() if sampler else None
    run_e2e(3)
    ms_e2e = timed(lambda: run_e2e(args.steps), 1) / args.steps

    # the dominant kernel INSIDE the timed schedule: uv_trace_ops runs the pass the way the graph runs it (both
    # utterance lanes on their own streams, same fork / join / stagger edges) with an event pair around every
    # launch; the union of the kernel's intervals over both lanes is its busy time, scaled to the graph-timed step
    coefs = program.step_coefs(model.model_config(), n_steps, float(cfg.model.diffusion.epsilon))
    out_dev = torch.empty(B, T, device=dev)
    io = (mix_dev.data_ptr(), noise.data_ptr(), out_dev.data_ptr(), None, coefs, False)
    # (how the two lanes' launches interleave differs from pass to pass, and with it the union of a kernel's intervals:
    # one warm-up trace, then the pass with the median GEMM share of five)
    ops = plan.pg.ops

    def kind_of(op):
        # the bf16 tier's one GEMM on f32 planes (the mel front end's DFT) runs in gemm_hp_tc_kernel, not in gemm_tc_kernel
        if op.type == _abi.OP_GEMM and op.a.dtype == _abi.DT_F32 and not hp:
            return "GEMM_DFT"
        return _abi.OP_NAMES[op.type]

    plan.trace_ops(*io)
    traces = []
    for _ in range(5):
        t0s, t1s = plan.trace_ops(*io)
        live = [(op, a, b) for op, a, b in zip(ops, t0s, t1s) if op.step != -2]
        total = max(b for _, _, b in live) - min(a for _, a, _ in live)
        busy, summed = {}, {}
        for kind in sorted({kind_of(op) for op, _, _ in live}):
            iv = [(a, b) for op, a, b in live if kind_of(op) == kind]
            busy[kind] = interval_union(iv)
            summed[kind] = sum(b - a for a, b in iv)
        traces.append((busy.get("GEMM", 0.0) / total, total, busy, summed))
    traces.sort(key=lambda t: t[0])
    gemm_share_runs = [round(t[0], 3) for t in traces]
    _, trace_total, by_kind_busy, by_kind_sum = traces[len(traces) // 2]
    torch.cuda.synchronize(dev)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    audio_s = B * T / fs * world
    fl = program.algorithmic_flops(model.model_config(), Lp, n_steps)
    pk = peaks()
    n_gemm = sum(1 for op in ops if kind_of(op) == "GEMM")
    gemm_share = by_kind_busy.get("GEMM", 0.0) / trace_total
    gemm_ms = gemm_share * ms                                  # <= ms_per_step by construction
    # the ConvBlocks of the 64-column levels run in their own fused kernel (csrc/convblock_pair.cuh): their FLOPs and
    # bytes are accounted to that kernel, not to gemm_tc_kernel.  Per launch, reference formulation: 2 C C (5 + 3 + 3)
    # per sample; bytes = the tensors the launch has to move once (x, condition, outputs; bf16)
    fused = [op for op in ops if op.type == _abi.OP_CONVBLOCK and op.step != -2]
    fused_flops = sum(2.0 * op.C * op.C * 11 * (op.M_valid * 64 // op.C) * (op.m_rows // op.Lh_A) for op in fused)
    fused_bytes = sum(128.0 * op.M_valid * (op.m_rows // op.Lh_A) * (1 + op.n_out + (1 if op.add1.off >= 0 else 0) + (1 if op.i2 else 0))
                      for op in fused)
    gemm_flops = fl["gemm_executed"] * B - fused_flops
    achieved = gemm_flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    # measured DRAM bytes per launch of the same kernel: an ncu pass over this command (tools/launch_summary.py),
    # quoted only if it was taken on this very build
    traffic, traffic_src, traffic_note = None, None, None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    build = source_hash()
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if hp:
            traffic_note = "no ncu pass of the fp32 tier's kernel"
        elif tj.get("kernel") == "gemm_tc_kernel" and tj.get("build") == build:
            traffic, traffic_src = tj["traffic_bytes_per_launch"], tj.get("source")
        else:
            traffic_note = f"profiles/roofline_traffic.json is from build {tj.get('build')}, this is {build}: not quoted"
    whole = fl["total"] * B * world / (ms * 1e-3) / 1e12
    roofline = dict(bound="tensor", kernel="gemm_hp_tc_kernel" if hp else "gemm_tc_kernel", achieved=achieved, peak=pk["tensor"], unit="TFLOP/s",
                    frac=achieved / pk["tensor"], traffic=traffic, traffic_unit="bytes/launch (dram read + write, ncu)",
                    traffic_source=traffic_src, traffic_note=traffic_note, build=build,
                    peak_source=pk["src"] + " sustained bf16 (cuBLAS, seconds-long loop under the power cap)",
                    launches_per_step=n_gemm, kernel_ms_per_step=gemm_ms, share_of_step=gemm_share, share_of_step_in_5_traces=gemm_share_runs,
                    # its share of the SUM of all launch durations of the traced pass: the figure that compares with the
                    # serialised ncu launch list (profiles/r2_launches_final.txt), where nothing overlaps
                    share_of_summed_launch_time=by_kind_sum.get("GEMM", 0.0) / max(sum(by_kind_sum.values()), 1e-9),
                    accounting=("kernel_ms_per_step = union of the kernel's launch intervals over both utterance lanes in the "
                                "two-stream schedule (uv_trace_ops, CUDA events) / traced pass x graph-timed ms_per_step; "
                                "achieved = algorithmic FLOPs of the layers this kernel executes (reference formulation, hoisted "
                                "1x1 condition projections counted once, the fused ConvBlocks of the 64-column levels excluded: "
                                "see roofline.fused_convblock) / that time"),
                    flops_per_clip=dict(gemm_executed=fl["gemm_executed"], gemm_reference=fl["gemm"], total_reference=fl["total"],
                                        in_gemm_kernel=gemm_flops / B, in_fused_convblocks=fused_flops / B),
                    whole_step_tflops=whole, whole_step_frac=whole / pk["tensor"],
                    whole_step_note="every dense FLOP of the pass in the reference's formulation / the whole timed step")
    if hp:
        roofline["note"] = ("fp32 parity tier: every product is 3 bf16 MMAs (x_hi w_hi + x_hi w_lo + x_lo w_hi), so the ceiling of "
                            "`achieved` (useful FLOPs) is peak / 3; activations are f32 planes (twice the bytes of the bf16 tier)")
    cb_ms = by_kind_busy.get("CONVBLOCK", 0.0) / trace_total * ms
    if fused and cb_ms > 0:
        gbs = fused_bytes / (cb_ms * 1e-3) / 1e9
        roofline["fused_convblock"] = dict(kernel="convblock_pair_kernel", bound="hbm", achieved=gbs, peak=pk["hbm"], unit="GB/s",
                                           frac=gbs / pk["hbm"], launches_per_step=len(fused), kernel_ms_per_step=cb_ms,
                                           algorithmic_bytes_per_step=fused_bytes, tflops=fused_flops / (cb_ms * 1e-3) / 1e12,
                                           note="bytes = x + condition + outputs of every launch, once (bf16); the kernel keeps "
                                                "c1 / c2 in shared memory; its MMAs (N = 64) are shared-memory bound (DESIGN.md)")
    if traffic and gemm_ms > 0:
        gbs = traffic * n_gemm / (gemm_ms * 1e-3) / 1e9
        roofline["hbm_view"] = dict(achieved=gbs, peak=pk["hbm"], unit="GB/s", frac=gbs / pk["hbm"],
                                    note="same launches seen as HBM traffic: the C <= 64 levels are HBM bound")
    line = dict(metric="audio-sec enhanced/sec", value=audio_s / (ms * 1e-3), unit="audio-s/s", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=ms, higher_is_better=True, scaling="weak", vs_baseline=None,
                dtype="fp32 (split-bf16 x3 on the tensor cores, fp32 accumulate; rel-L2 <= 1e-4 tier)" if hp else "bf16",
                data="synthetic",
                config=dict(workload=wl["desc"], model=wl["model"], batch_per_gpu=B, clip_samples=T, n_steps=n_steps,
                            weights="random-init (seeded)", sharding=f"utterances x{world}, no collective",
                            lanes=f"{len(plan.pg.lanes)} utterance lane(s) per GPU (independent halves of the batch on two captured streams)",
                            l2="per-step working set (GBs of activations) >> 126 MB L2; no explicit flush",
                            arena_gb=round(plan.arena_bytes / 1e9, 2),
                            timing="CUDA events on the launch stream, max over ranks"),
                clocks=clocks,
                e2e=dict(value=audio_s / (ms_e2e * 1e-3), unit="audio-s/s", ms_per_step=ms_e2e,
                         h2d_bytes_per_step=B * T * 4, d2h_bytes_per_step=B * T * 4,
                         note="model.enhance() per step on pinned host buffers; each step's H2D / D2H run on copy streams and "
                              "overlap the neighbouring steps' compute (host at most two steps ahead); all inside the timed region"),
                gpu_launches=plan.num_launches * args.steps,
                roofline=roofline,
                busy_ms_in_schedule={k: round(v / trace_total * ms, 3) for k, v in sorted(by_kind_busy.items(), key=lambda kv: -kv[1])},
                traced_pass_ms=round(trace_total, 3))
    if world > 1:
        line["nccl"] = nccl_log_report(nccl_log)
    if world == 1 and not args.no_cpu_baseline:
        a_s, times, threads = cpu_reference_run(WORKLOADS["cfg2"], 1, 2.0, 3)
        line["cpu_baseline"] = dict(value=a_s / min(times), unit="audio-s/s", cores=threads, kind="port",
                                    sample="1 clip x 2 s, 8 steps (BASELINE configs[0]), best of 3, torch CPU fp32 oracle port")
    if world == 1 and args.gpu_eager_baseline:
        line["gpu_eager_baseline"] = gpu_eager_baseline(wl, dev, args.eager_clips)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_sweep(args, wl, rank, world, local_rank):
    """BASELINE configs[4] as stated: `clips` 30 s clips in pinned host memory, dealt to the ranks in contiguous
    shards (sharding.shard_bounds), enhanced in batches of B through the public API and written back to a pinned host
    buffer at the clip's original index (the 'final host gather' is rank-local: every rank owns its slice of the
    output; nothing is pickled through the process group).  H2D of batch k + 1 and D2H of batch k - 1 run on their own
    streams under the compute of batch k.  Strong scaling: the job is fixed, `value` = all audio seconds / wall time of
    the slowest rank, every copy included."""
    import torch.distributed as dist

    from open_universe_b200 import sharding

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    nccl_log = None
    if world > 1:
        nccl_log = nccl_log_setup(rank)
        dist.init_process_group("nccl", device_id=dev)
    cfg, model = setup_native(wl, dev)
    fs = int(cfg.model.fs)
    B, T, n_steps = wl["B"], wl["T"], wl["n_steps"]
    n_clips = args.clips or wl["clips"]
    lo, hi = sharding.shard_bounds(n_clips, rank, world)
    mine = hi - lo
    # synthetic clips of the shard: a base block of distinct clips repeated with per-clip gains (the cost of the path
    # does not depend on the data; generating 7.9 GB of filtered noise would only lengthen the setup)
    from open_universe_b200 import testing
    base = testing.synthetic_noisy_speech(min(mine, 16), T, seed=77 + rank)
    clips = torch.empty(mine, T).pin_memory()
    for k in range(mine):
        clips[k].copy_(base[k % base.shape[0]] * (0.5 + (k % 7) / 7.0))
    result = torch.zeros(mine, T).pin_memory()
    batches = [(k, min(k + B, mine)) for k in range(0, mine, B)]
    gen = torch.Generator(device=dev).manual_seed(5 + rank)
    s_in, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    main = torch.cuda.current_stream(dev)
    dbuf = [torch.empty(B, T, device=dev) for _ in range(2)]
    in_ready = [torch.cuda.Event() for _ in range(2)]
    in_free = [torch.cuda.Event() for _ in range(2)]

    def run_all(n_batches=None):
        todo = batches if n_batches is None else batches[:n_batches]
        outs = [None, None]
        out_done = [None, None]
        for i, (a, b) in enumerate(todo[:1]):
            with torch.cuda.stream(s_in):
                dbuf[0][: b - a].copy_(clips[a:b], non_blocking=True)
                in_ready[0].record(s_in)
        for i, (a, b) in enumerate(todo):
            cur = i & 1
            if i + 1 < len(todo):   # prefetch the next batch while this one computes
                na, nb = todo[i + 1]
                with torch.cuda.stream(s_in):
                    if i >= 1:
                        s_in.wait_event(in_free[cur ^ 1])
                    dbuf[cur ^ 1][: nb - na].copy_(clips[na:nb], non_blocking=True)
                    in_ready[cur ^ 1].record(s_in)
            main.wait_event(in_ready[cur])
            if out_done[cur] is not None:
                main.wait_event(out_done[cur])   # the result tensor of two batches ago has left the device
            y = model.enhance(dbuf[cur][: b - a], n_steps=n_steps, rng=gen)
            in_free[cur].record(main)
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev)
                result[a:b].copy_(y, non_blocking=True)
                done = torch.cuda.Event()
                done.record(s_out)
            outs[cur], out_done[cur] = y, done
        torch.cuda.synchronize(dev)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    run_all(min(len(batches), max(3, args.warmup)))   # warm-up: plan creation, graph capture, allocator
    sampler = ClockSampler(local_rank) if rank == 0 else None
    times = []
    for _ in range(max(1, args.steps)):
        sync_all()
        t0 = time.perf_counter()
        run_all()
        sync_all()
        times.append(time.perf_counter() - t0)
    clocks = sampler.stop() if sampler else None
    sec = torch.tensor([statistics.median(times)], device=dev)
    if world > 1:
        dist.all_reduce(sec, op=dist.ReduceOp.MAX)
    ok = bool(torch.isfinite(result).all()) and float(result.abs().max()) > 0
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    sec = sec.item()
    audio_s = n_clips * T / fs
    plan = next(iter(model._plans.values()))
    line = dict(metric="audio-sec enhanced/sec", value=audio_s / sec, unit="audio-s/s", n_gpus=world, steps=max(1, args.steps),
                warmup=max(3, args.warmup), ms_per_step=sec * 1e3, higher_is_better=True, scaling="strong", vs_baseline=None,
                dtype="bf16", data="synthetic",
                config=dict(workload=wl["desc"], model=wl["model"], clips=n_clips, clip_samples=T, batch=B, n_steps=n_steps,
                            step="one pass over the whole clip set (host to host)", warmup_unit="batches",
                            sharding=f"contiguous shards of {mine} clips per rank x{world}, no collective, rank-local pinned output",
                            pipeline="H2D of batch k+1 / D2H of batch k-1 on their own streams under the compute of batch k",
                            timing="wall clock around the whole job between barriers + device syncs, median of steps, max over ranks",
                            outputs_finite=ok, arena_gb=round(plan.arena_bytes / 1e9, 2)),
                clocks=clocks,
                e2e=dict(value=audio_s / sec, unit="audio-s/s", ms_per_step=sec * 1e3,
                         h2d_bytes_per_step=n_clips * T * 4, d2h_bytes_per_step=n_clips * T * 4),
                gpu_launches=plan.num_launches * len(batches) * max(1, args.steps))
    if world > 1:
        line["nccl"] = nccl_log_report(nccl_log)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"],
                    help="fp32: the rel-L2 <= 1e-4 parity tier (f32 activations, split-bf16 tensor-core GEMMs); default bf16 = the headline")
    ap.add_argument("--clips", type=int, default=None, help="cfg5_sweep: size of the clip set (default 4096)")
    ap.add_argument("--ref-clips", type=int, default=4, help="--impl reference: clips per CPU step")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gpu-eager-baseline", action="store_true",
                    help="also time the oracle port with PyTorch eager ops (cuDNN) on the GPU: a stated baseline, off by default")
    ap.add_argument("--eager-clips", type=int, default=16)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    sweep = "clips" in wl
    if args.steps is None:
        args.steps = 1 if sweep else 10
    if args.impl == "reference":
        run_reference_arm(args, wl, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (native arm) needs a CUDA sm_100 device; there is no CPU fallback")
    if sweep:
        run_sweep(args, wl, rank, world, local_rank)
    else:
        run_native(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
