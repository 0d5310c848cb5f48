#!/usr/bin/env python
"""Headline benchmark: audio-seconds enhanced per second, 8-step UNIVERSE++ 16 kHz.

  python bench.py --gpus N --steps K --warmup W            # native sm_100a path (this repo)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's algorithm on the host CPU

A "step" is one full enhance() pass over one batch of synthetic noisy clips
(BASELINE.json configs[1]: batch 64 x 4 s, 8 sampler steps, bf16 operands / fp32 accumulate).
N > 1 (torchrun, one rank per GPU): utterances are sharded, every rank runs the same per-GPU
workload (weak scaling), no data-path collective; time = max over ranks.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
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
    "cfg5": dict(model="universepp_16k", B=32, T=480000, n_steps=8, desc="UNIVERSE++ 16 kHz, 30 s clips in batches of 32 (utterance-sharded sweep)"),
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


def cpu_reference_run(wl, n_clips, clip_seconds, repeats, threads=None):
    """The reference's algorithm on the host cores: the torch-CPU fp32 oracle port
    (oracle/universe_oracle.py; the Python reference itself cannot travel to the GPU box)."""
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
    """--impl reference: rank 0 alone times the CPU algorithm on a bounded sample of the workload."""
    if rank != 0:
        return
    clip_s = wl["T"] / 16000 if "16k" in wl["model"] else wl["T"] / 24000
    clip_s = min(clip_s, 4.0)
    n_clips = 1
    audio_s, times, threads = cpu_reference_run(wl, n_clips, clip_s, args.warmup + args.steps)
    timed = times[args.warmup:]
    ms = 1e3 * sum(timed) / len(timed)
    value = audio_s / (ms / 1e3)
    sample = f"{n_clips} clip x {clip_s:g} s per step ({wl['n_steps']} sampler steps), torch CPU fp32 oracle port of the reference enhance()"
    line = dict(impl="reference", metric="audio-sec enhanced/sec", value=value, unit="audio-s/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=ms, higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["desc"], sample=sample),
                cpu_baseline=dict(value=value, unit="audio-s/s", cores=threads, kind="port", sample=sample),
                e2e=dict(value=value, unit="audio-s/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0)
    print(json.dumps(line), flush=True)


def run_native(args, wl, rank, world, local_rank):
    import torch.distributed as dist

    from open_universe_b200 import config, networks, program, testing  # noqa: F401

    if args.warmup < 3:
        args.warmup = 3
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        # stdout carries exactly one JSON line: keep NCCL's version / info banner out of it
        # (VERSION and every level above it print the banner; no NCCL_DEBUG at all prints nothing)
        os.environ.pop("NCCL_DEBUG", None)
        if os.environ.get("UV_NCCL_DEBUG"):
            os.environ["NCCL_DEBUG"] = os.environ["UV_NCCL_DEBUG"]
        dist.init_process_group("nccl", device_id=dev)
    cfg = load_model_cfg(wl["model"])
    fs = int(cfg.model.fs)
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=0)
    model = model.to(dev).eval()
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

    def step_e2e():
        m = mix_host.to(dev, non_blocking=True)          # H2D from pinned memory
        out = model.enhance(m, n_steps=n_steps, rng=gen)   # public API (draws the noise with torch.randn)
        out_host.copy_(out)                              # D2H of the result (blocking)
        return out_host

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
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    # per-kernel device times of one eager pass (CUDA events on the launching stream)
    coefs = program.step_coefs(model.model_config(), n_steps, float(cfg.model.diffusion.epsilon))
    out_dev = torch.empty(B, T, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    prof = plan.profile_ops(mix_dev.data_ptr(), noise.data_ptr(), out_dev.data_ptr(), None, coefs, False, stream)
    from open_universe_b200 import _abi
    by_kind = {}
    for op, t in zip(plan.pg.ops, prof):
        k = _abi.OP_NAMES[op.type]
        by_kind[k] = by_kind.get(k, 0.0) + t
    # two utterance lanes: the timed launches are half-batch ones whose start-up / tail costs overlap the other lane
    # inside the graph but are paid in full by a launch timed alone; also time the same kernel on whole-batch
    # launches (single-lane program of the same workload) so that both views are on record
    gemm_ms_1lane = None
    if rank == 0 and len(plan.pg.lanes) > 1 and not args.no_single_lane:
        from open_universe_b200 import _native
        pg1 = program.build_program(model._packed, model.model_config(), B, T, n_steps, lanes=1)
        plan1 = _native.NativePlan(model._native_model, pg1, use_graph=False)
        for _ in range(2):
            prof1 = plan1.profile_ops(mix_dev.data_ptr(), noise.data_ptr(), out_dev.data_ptr(), None, coefs, False, stream)
        gemm_ms_1lane = sum(t for op, t in zip(pg1.ops, prof1) if op.type == _abi.OP_GEMM)
        n_gemm_1lane = sum(1 for op in pg1.ops if op.type == _abi.OP_GEMM)
        plan1.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    audio_s = B * T / fs * world
    fl = program.algorithmic_flops(model.model_config(), Lp, n_steps)
    pk = peaks()
    gemm_ms = by_kind.get("GEMM", 0.0)
    n_gemm = sum(1 for op in plan.pg.ops if op.type == _abi.OP_GEMM)
    achieved = fl["gemm"] * B / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    eager_ms = sum(by_kind.values())
    # measured DRAM bytes per launch of the same kernel (ncu pass over this command, profiles/roofline_traffic.json)
    traffic, traffic_src = None, None
    tpath = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("kernel") == "gemm_tc_kernel":
            traffic, traffic_src = tj["traffic_bytes_per_launch"], tj.get("source")
    roofline = dict(bound="tensor", kernel="gemm_tc_kernel", achieved=achieved, peak=pk["tensor"], unit="TFLOP/s",
                    frac=achieved / pk["tensor"], traffic=traffic, traffic_unit="bytes/launch (dram read + write, ncu)",
                    traffic_source=traffic_src,
                    hbm_view=(dict(achieved=traffic * n_gemm / (gemm_ms * 1e-3) / 1e9, peak=pk["hbm"], unit="GB/s",
                                   frac=traffic * n_gemm / (gemm_ms * 1e-3) / 1e9 / pk["hbm"],
                                   note="same launches seen as HBM traffic: the C <= 64 levels are HBM bound, the rest tensor bound")
                              if traffic and gemm_ms > 0 else None),
                    peak_source=pk["src"] + " sustained bf16",
                    launches_per_step=n_gemm, kernel_ms_per_step=gemm_ms, share_of_step=gemm_ms / eager_ms if eager_ms else None,
                    whole_step_tflops=fl["total"] * B * world / (ms * 1e-3) / 1e12)
    if gemm_ms_1lane:
        a1 = fl["gemm"] * B / (gemm_ms_1lane * 1e-3) / 1e12
        roofline["whole_batch_launches"] = dict(
            achieved=a1, frac=a1 / pk["tensor"], launches_per_step=n_gemm_1lane, kernel_ms_per_step=gemm_ms_1lane,
            note="same kernel, same workload as ONE utterance lane (whole-batch launches, eager, CUDA events); the headline "
                 "fields above time the half-batch launches of the two-lane program one at a time, which counts the "
                 "start-up / tail of every launch that the other lane hides inside the graph")
    line = dict(metric="audio-sec enhanced/sec", value=audio_s / (ms * 1e-3), unit="audio-s/s", n_gpus=world, steps=args.steps,
                warmup=args.warmup, ms_per_step=ms, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16",
                data="synthetic",
                config=dict(workload=wl["desc"], model=wl["model"], batch_per_gpu=B, clip_samples=T, n_steps=n_steps,
                            weights="random-init (seeded)", sharding=f"utterances x{world}, no collective",
                            lanes=f"{len(plan.pg.lanes)} utterance lane(s) per GPU (independent halves of the batch on two captured streams)",
                            l2="per-step working set (GBs of activations) >> 126 MB L2; no explicit flush",
                            timing="CUDA events on the launch stream, max over ranks"),
                clocks=clocks,
                e2e=dict(value=audio_s / (ms_e2e * 1e-3), unit="audio-s/s", ms_per_step=ms_e2e,
                         h2d_bytes_per_step=B * T * 4, d2h_bytes_per_step=B * T * 4),
                gpu_launches=plan.num_launches * args.steps,
                roofline=roofline,
                breakdown_ms={k: round(v, 3) for k, v in sorted(by_kind.items(), key=lambda kv: -kv[1])})
    if world == 1 and not args.no_cpu_baseline:
        a_s, times, threads = cpu_reference_run(WORKLOADS["cfg2"], 1, 2.0, 3)
        line["cpu_baseline"] = dict(value=a_s / min(times), unit="audio-s/s", cores=threads, kind="port",
                                    sample="1 clip x 2 s, 8 steps (BASELINE configs[0]), best of 3, torch CPU fp32 oracle port")
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-single-lane", action="store_true", help="skip the whole-batch-launch view of the roofline")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (native arm) needs a CUDA sm_100 device; there is no CPU fallback")
    run_native(args, wl, rank, world, local_rank)


if __name__ == "__main__":
    main()
