"""Drop-in `Universe` / `UniverseGAN` model objects (reference
open_universe/networks/universe/universe.py:44-375, universe_gan.py:62-151).

They keep the reference's constructor signature, state-dict layout, EMA handling and the
`enhance()` keyword surface, but own no PyTorch compute: `enhance()` compiles the weights
into a packed blob + op program (weights.py, program.py) and replays one CUDA graph of
hand-written sm_100a kernels through the C ABI (_native.py).
"""
import collections
import itertools
import logging
import math
import os
from typing import Optional

import torch

from ... import _native, program, weights
from ...config import AttrDict, instantiate, to_attr
from ...ema import ExponentialMovingAverage
from .layers import SignalDecouplingP

log = logging.getLogger(__name__)


class Universe(torch.nn.Module):
    def __init__(self, fs, normalization_norm, score_model, condition_model, diffusion, losses=None, training=None,
                 validation=None, optimizer=None, scheduler=None, grad_clipper=None, transform=None,
                 normalization_kwargs={}, with_noise_target=False, detach_cond=False, edm=None):
        super().__init__()
        self.fs = fs
        self.normalization_norm = normalization_norm
        self.normalization_kwargs = to_attr(dict(normalization_kwargs or {}))
        self.with_noise_target = with_noise_target
        self.detach_cond = detach_cond
        self.opt_kwargs, self.schedule_kwargs, self.grad_clip_kwargs = optimizer, scheduler, grad_clipper
        self.diff_kwargs = to_attr(dict(diffusion))
        self.losses_kwargs = to_attr(losses or {})
        self.val_kwargs = to_attr(validation or {})
        self.train_kwargs = to_attr(training or {})
        if transform is not None:
            raise NotImplementedError("`transform` (STFT-domain models) is outside the supported enhance() path")
        score_model, condition_model = to_attr(dict(score_model)), to_attr(dict(condition_model))
        self._score_cfg, self._cond_cfg = score_model, condition_model
        if edm is not None:
            self.edm_kwargs = to_attr(dict(edm))
            self._edm_model = instantiate(score_model)
            self.with_edm = True
        else:
            self.edm_kwargs = None
            self.score_model = instantiate(score_model)
            self.with_edm = False
        self.condition_model = instantiate(condition_model)
        self.n_channels = score_model.get("n_channels", 32)
        rate_factors = score_model.get("rate_factors", [2, 4, 4, 5])
        self.n_stages = len(rate_factors)
        self.latent_n_channels = 2**self.n_stages * self.n_channels
        self.tot_ds = math.prod(int(r) for r in rate_factors)
        self.init_losses(score_model, condition_model, self.losses_kwargs, self.train_kwargs)

        self.ema_decay = float(self.train_kwargs.get("ema_decay", 0.0) or 0.0)
        self._error_loading_ema = False
        self.ema = ExponentialMovingAverage(self.model_parameters(), decay=self.ema_decay) if self.ema_decay > 0.0 else None

        self._packed = None      # weights.PackedModel
        self._native_model = None
        self._plans = collections.OrderedDict()
        self.use_cuda_graph = True
        # plan cache: least recently used plans are closed (arena, graph) when the cache holds more than
        # plan_cache_bytes of arenas or plan_cache_max plans (a directory of files of many lengths must not run the
        # device out of memory); the graph of a plan is captured at its second use
        self.plan_cache_bytes = int(float(os.environ.get("UV_PLAN_CACHE_GB", "48")) * 2**30)
        self.plan_cache_max = int(os.environ.get("UV_PLAN_CACHE_MAX", "32"))
        # "bf16" (default): bf16 operands, fp32 accumulation on the tensor cores (north_star gate rel-L2 <= 1e-2);
        # "fp32": the parity tier (rel-L2 <= 1e-4): f32 activations and weights, fp32 FMA kernels (csrc/hp_ops.cuh)
        self.precision = os.environ.get("UV_PRECISION", "bf16")

    # ------------------------------------------------------------------ reference API surface
    def init_losses(self, score_model, condition_model, losses, training):
        """Training losses (MDN heads, reference universe.py:138-168) are not part of inference;
        their checkpoint keys (loss_*) are ignored by load_state_dict below."""

    def get_score_model(self):
        return self._edm_model if self.with_edm else self.score_model

    def model_parameters(self):
        return itertools.chain(self.get_score_model().parameters(), self.condition_model.parameters())

    def model_config(self):
        """The resolved `model` config node this object was built from (what the compiler needs)."""
        cfg = dict(fs=self.fs, normalization_norm=self.normalization_norm, normalization_kwargs=dict(self.normalization_kwargs),
                   score_model=dict(self._score_cfg), condition_model=dict(self._cond_cfg), diffusion=dict(self.diff_kwargs))
        if self.with_edm:
            cfg["edm"] = dict(self.edm_kwargs)
        return cfg

    def load_state_dict(self, state_dict, strict=True, assign=False):
        # training-only sub-modules of the reference (MDN / GAN losses) are not instantiated here
        sd = {k: v for k, v in state_dict.items() if not k.startswith("loss_")}
        self._invalidate()
        return super().load_state_dict(sd, strict=strict, assign=assign)

    def train(self, mode=True, no_ema=False):
        """Same EMA protocol as the reference (universe.py:841-862): eval() swaps the EMA shadow
        parameters in, train() restores the raw ones."""
        res = super().train(mode)
        if getattr(self, "ema", None) is None:
            return res
        if not self._error_loading_ema:
            if mode is False and not no_ema:
                self.ema.store(self.model_parameters())
                self.ema.copy_to(self.model_parameters())
            elif self.ema.collected_params is not None:
                self.ema.restore(self.model_parameters())
        self._invalidate()
        return res

    def eval(self, no_ema=False):
        return self.train(False, no_ema=no_ema)

    def to(self, *args, **kwargs):
        if getattr(self, "ema", None) is not None:
            self.ema.to(*args, **kwargs)
        self._invalidate()
        return super().to(*args, **kwargs)

    def aux_to_wav(self, y_aux):
        return y_aux

    @staticmethod
    def _resample(x, kernel, orig, new):
        """torchaudio `Resample` with its precomputed filter bank (functional._apply_sinc_resample_kernel): pad
        (width, width + orig), strided conv with the `new` phase filters, interleave the phases, crop."""
        width = (kernel.shape[-1] - orig) // 2
        shape = x.shape
        w = torch.nn.functional.pad(x.reshape(-1, shape[-1]), (width, width + orig))
        y = torch.nn.functional.conv1d(w[:, None], kernel.to(w), stride=orig)
        y = y.transpose(1, 2).reshape(w.shape[0], -1)[..., : math.ceil(new * shape[-1] / orig)]
        return y.reshape(shape[:-1] + y.shape[-1:])

    def pad(self, x, pad=None):
        if pad is None:
            pad = self.tot_ds - x.shape[-1] % self.tot_ds
        return torch.nn.functional.pad(x, (pad // 2, pad - pad // 2)), pad

    def unpad(self, x, pad):
        return x[..., pad // 2 : -(pad - pad // 2)]

    def forward(self, xt, sigma, cond):
        raise RuntimeError("Universe has no PyTorch forward in open_universe_b200: use enhance()")

    # ------------------------------------------------------------------ native plumbing
    def _invalidate(self):
        for p in getattr(self, "_plans", {}).values():
            p.close()
        self._plans = collections.OrderedDict()
        if getattr(self, "_native_model", None) is not None:
            self._native_model.close()
        self._native_model = None
        self._packed = None

    def _device(self):
        return next(self.parameters()).device

    def _get_plan(self, B, T, n_steps, aux=False):
        dev = self._device()
        if dev.type != "cuda":
            raise RuntimeError(
                "open_universe_b200 runs only on a CUDA sm_100 (B200) device; the model is on "
                f"'{dev}'. There is no CPU / PyTorch fallback: move it with model.to('cuda')."
            )
        if self.precision not in program.PRECISIONS:
            raise ValueError(f"precision must be one of {program.PRECISIONS}, got {self.precision!r}")
        hp = self.precision == "fp32"
        if self._packed is not None and hp and not self._packed.hp:
            self._invalidate()   # the blob on the device lacks the fp32 weight images: pack again with them
        if self._packed is None:
            sd = {k: v.detach().float().cpu() for k, v in self.state_dict().items()}
            self._packed = weights.pack_model(sd, self.model_config(), hp=hp)
            self._native_model = _native.NativeModel(self._packed.blob.bytes(), dev.index or 0)
        key = (B, T, n_steps, aux) + (("fp32",) if hp else ())
        plan = self._plans.get(key)
        if plan is not None:
            self._plans.move_to_end(key)
            return plan
        pg = program.build_program(self._packed, self.model_config(), B, T, n_steps, aux=aux, precision=self.precision)
        # make room first: the new arena must fit next to the ones that stay
        held = sum(p.arena_bytes for p in self._plans.values())
        while self._plans and (held + pg.arena > self.plan_cache_bytes or len(self._plans) >= self.plan_cache_max):
            _, old = self._plans.popitem(last=False)
            held -= old.arena_bytes
            old.close()
        with torch.cuda.device(dev):
            plan = _native.NativePlan(self._native_model, pg, use_graph=2 if self.use_cuda_graph else 0)
        self._plans[key] = plan
        return plan

    # ------------------------------------------------------------------ enhance (reference universe.py:231-375)
    def enhance(
        self,
        mix,
        n_steps: Optional[int] = None,
        epsilon: Optional[float] = None,
        target: Optional[torch.Tensor] = None,
        fake_score_snr: Optional[float] = None,
        rng: Optional[torch.Generator] = None,
        use_aux_signal: Optional[bool] = False,
        keep_rms: Optional[bool] = False,
        ensemble: Optional[int] = None,
        ensemble_stat: Optional[str] = "median",
        warm_start: Optional[int] = None,
        noise_seed: Optional[int] = None,
        lengths=None,
    ) -> torch.Tensor:
        """Reference signature (universe.py:231-245) plus two trailing keywords of this port.
        noise_seed = None draws the sampler noise with torch.randn(generator=rng) exactly like the reference (same
        draws, same order); noise_seed = k lets the kernels draw it themselves (Philox keyed by k: same distribution,
        another stream, no [n_steps, B, Lp] noise tensor; `rng` is then unused).
        lengths = [T_0, ..., T_{B-1}]: row b of `mix` holds T_b valid samples (right-padded with anything) and is
        enhanced exactly like a file of that length on its own -- its own padding, normalisation statistics, mel
        frames and GRU length -- which is how the reference's file loop (bin/enhance.py:168-192) treats a corpus
        whose files all differ in length.  Row b of the result is valid up to T_b and zero after it.  In parity
        mode row b consumes the first Lp_b values of its rows of the [B, 1, Lp] draws."""
        out = self._enhance_impl(mix, n_steps, epsilon, target, fake_score_snr, rng, use_aux_signal, keep_rms,
                                 ensemble, ensemble_stat, warm_start, noise=None, return_steps=False, noise_seed=noise_seed,
                                 lengths=lengths)
        return out

    def enhance_debug(self, mix, noise, n_steps=None, epsilon=None, keep_rms=False, lengths=None):
        """Parity entry point: runs enhance() with an injected noise sequence `noise`
        ([n_steps, B, Lp] on the model device) and also returns the per-step network outputs."""
        return self._enhance_impl(mix, n_steps, epsilon, None, None, None, False, keep_rms, None, "median", None,
                                  noise=noise, return_steps=True, lengths=lengths)

    def _enhance_impl(self, mix, n_steps, epsilon, target, fake_score_snr, rng, use_aux_signal, keep_rms, ensemble,
                      ensemble_stat, warm_start, noise, return_steps, noise_seed=None, lengths=None):
        if epsilon is None:
            epsilon = self.diff_kwargs.epsilon
        if n_steps is None:
            n_steps = self.diff_kwargs.n_steps
        n_steps, epsilon = int(n_steps), float(epsilon)
        x_ndim = mix.ndim
        if x_ndim == 1:
            mix = mix[None, None, :]
        elif x_ndim == 2:
            mix = mix[:, None, :]
        elif x_ndim > 3:
            raise ValueError("The input should have at most 3 dimensions")
        if mix.shape[1] != 1:
            raise ValueError("open_universe_b200 supports single-channel conditioners (input_channels == 1)")
        if ensemble is not None and ensemble_stat not in ("mean", "median", "signal_median"):
            raise NotImplementedError()   # reference universe.py:359-367: only looked at when an ensemble is asked for
        aux = bool(use_aux_signal) or warm_start is not None
        if warm_start is not None and not 0 <= int(warm_start) < n_steps:
            raise ValueError("warm_start must be a step index in [0, n_steps)")
        if self.normalization_norm not in (2, "2"):
            raise NotImplementedError("only normalization_norm == 2 is supported")
        if self.diff_kwargs.get("schedule", "geometric") != "geometric":
            raise NotImplementedError()
        if n_steps < 2:
            raise ValueError("n_steps must be at least 2")

        if lengths is not None:
            lengths = [int(v) for v in lengths]
            if len(lengths) != mix.shape[0]:
                raise ValueError("lengths: one value per row of mix")
            if target is not None:
                raise NotImplementedError("lengths with the known-score test hook")
        if ensemble is not None:
            mix_shape = mix.shape
            mix = torch.stack([mix] * ensemble, dim=0).view((-1,) + mix_shape[1:])
            if lengths is not None:
                lengths = lengths * ensemble

        B, T = mix.shape[0], mix.shape[-1]
        dev = self._device()
        if target is not None:
            # "perfect score" test hook (reference universe.py:269-298): the networks are not evaluated at all
            if dev.type != "cuda":
                raise RuntimeError("open_universe_b200 runs only on a CUDA device")
            if target.ndim == 1:
                target = target[None, None, :]
            elif target.ndim == 2:
                target = target[:, None, :]
            if ensemble is not None:
                target = torch.stack([target] * ensemble, dim=0).view((-1,) + target.shape[1:])
            x = self._enhance_known_score(mix.to(dev), target.to(dev), n_steps, epsilon, fake_score_snr, rng, keep_rms, noise)
            return self._finish(x, x_ndim, ensemble, ensemble_stat, mix_shape if ensemble is not None else None)
        plan = self._get_plan(B, T, n_steps, aux=aux)
        plan.set_lengths(lengths)   # (None: back to full-length rows)
        Lp = plan.pg.Lp
        mix_f = mix.reshape(B, T).to(device=dev, dtype=torch.float32).contiguous()
        seeded = noise is None and noise_seed is not None
        if seeded:
            plan.set_noise_seed(noise_seed)   # the kernels draw z[step][utterance][sample] themselves
        elif noise is None:
            # same draws, same order as the reference (universe.py:39-41,326,338): n_steps tensors of (B,1,Lp)
            # (use_aux_signal draws nothing, warm_start = n draws for the steps from n on: universe.py:317-331)
            first = n_steps if use_aux_signal else (int(warm_start) if warm_start is not None else 0)
            noise = (torch.empty if first == 0 else torch.zeros)((n_steps, B, Lp), dtype=torch.float32, device=dev)
            for i in range(first, n_steps):   # drawn in place: randn(out=) is empty(...).normal_(), the same draws
                torch.randn((B, 1, Lp), dtype=torch.float32, device=dev, generator=rng, out=noise[i].view(B, 1, Lp))
        else:
            noise = noise.to(device=dev, dtype=torch.float32).reshape(n_steps, B, Lp).contiguous()
        out = torch.empty((B, T), dtype=torch.float32, device=dev)
        dbg = torch.empty((n_steps, B, Lp), dtype=torch.float32, device=dev) if return_steps else None
        coefs = program.step_coefs(self.model_config(), n_steps, epsilon)
        stream = torch.cuda.current_stream(dev).cuda_stream
        io = (mix_f.data_ptr(), None if seeded else noise.data_ptr(), out.data_ptr(), dbg.data_ptr() if dbg is not None else None,
              coefs, bool(keep_rms), stream)
        with torch.cuda.device(dev):
            if not aux:
                plan.enhance(*io)
            else:
                self._enhance_aux(plan, io, B, Lp, n_steps, noise, use_aux_signal, warm_start)
        x = self._finish(out.to(mix.dtype)[:, None, :], x_ndim, ensemble, ensemble_stat, mix_shape if ensemble is not None else None)
        if return_steps:
            return x, dbg
        return x

    def _enhance_aux(self, plan, io, B, Lp, n_steps, noise, use_aux_signal, warm_start):
        """use_aux_signal / warm_start (reference universe.py:313-331): front end + conditioner (including the
        decoder tail that yields y_hat) run natively, `aux_to_wav` maps y_hat to a waveform with a few device-side
        torch ops (off the hot path), which either is the result (use_aux_signal) or, plus sigma_n z, the state
        the native sampler resumes from at step n = warm_start.  The draws are consumed like the reference's:
        noise[warm_start] initialises x, noise[warm_start + 1 ...] are the z of the following steps."""
        pg = plan.pg
        plan.run_ops(0, pg.idx_init_x, *io)
        if getattr(pg, "idx_aux", None) is not None:
            # native: one AUX launch = aux_to_wav(y_hat) [+ sigma_n z_n] into the sampler state (csrc/aux_ops.cuh)
            if use_aux_signal:
                plan.run_ops(pg.idx_aux, pg.idx_aux + 1, *io)
            else:
                n0 = int(warm_start)
                plan.run_ops(pg.idx_aux + 1 + n0, pg.idx_aux + 2 + n0, *io)
                plan.run_ops(pg.step_first_op[n0], pg.idx_post, *io)
            plan.run_ops(pg.idx_post, pg.idx_post + 1, *io)
            return
        # a model without the signal_decoupling_layer (aux_to_wav is the identity on a 1-channel y_hat): host-side ops
        if noise is None:
            raise NotImplementedError("noise_seed with use_aux_signal / warm_start needs the signal_decoupling_layer")
        t = pg.y_hat
        C = t.nch * 8
        planes = plan.arena_view(t.off, t.nch * t.P * 8, noise.device).view(t.nch, t.P, 8)
        rows = planes[:, t.G : t.G + B * t.Lh].reshape(t.nch, B, t.Lh, 8)[:, :, :Lp]        # [chunk][b][l][8]
        y_hat = rows.permute(1, 0, 3, 2).reshape(B, C, Lp)
        sig = self.aux_to_wav(y_hat)[:, 0, :]
        x = plan.arena_view(pg.x, B * Lp, noise.device).view(B, Lp)
        if use_aux_signal:
            x.copy_(sig)
        else:
            n0 = int(warm_start)
            sigma = program.sampler_tables(self.model_config(), n_steps)[0]
            x.copy_(sig + float(sigma[n0]) * noise[n0])
            plan.run_ops(pg.step_first_op[n0], pg.idx_post, *io)
        plan.run_ops(pg.idx_post, pg.idx_post + 1, *io)

    @staticmethod
    def _finish(x, x_ndim, ensemble, ensemble_stat, mix_shape):
        """Ensemble statistic and rank restoration (reference universe.py:359-373)."""
        if ensemble is not None:
            x = x.view((-1,) + tuple(mix_shape))
            if ensemble_stat == "mean":
                x = x.mean(dim=0)
            elif ensemble_stat == "median":
                x = x.median(dim=0).values
            elif x.is_cuda:
                x = _native.signal_median(x)   # uv_signal_median: two launches, no sort of the whole ensemble
            else:
                x = signal_median(x)
        if x_ndim == 1:
            x = x[0, 0]
        elif x_ndim == 2:
            x = x[:, 0, :]
        return x

    def _enhance_known_score(self, mix, target, n_steps, epsilon, fake_score_snr, rng, keep_rms, noise):
        """enhance() with `target` given (reference universe.py:269-298,321-357): the score is the analytic
        -(x - target) / sigma^2 plus white noise at `fake_score_snr` dB (default 5), so the sampler follows a
        closed form.  Off the hot path: a handful of elementwise device ops, no network evaluation.
        `noise`: optional injected list of the n_steps sampler draws (unit variance), as in enhance_debug."""
        mix_len = mix.shape[-1]
        mix_rms = mix.square().mean(dim=(-2, -1), keepdim=True).sqrt()
        mix, pad = self.pad(mix)
        target, _ = self.pad(target, pad=pad)
        kw = dict(self.normalization_kwargs or {})
        level = 10.0 ** (float(kw.get("level_db", 0.0)) / 20.0)

        def norm2(sig):   # utils/norm.py:22-23,47-87 with norm = 2, zero_mean
            sig = sig - sig.mean(dim=(1, 2), keepdim=True)
            return sig * (level / sig.std(dim=(1, 2), keepdim=True).clamp(min=1e-5))

        if kw.get("ref", "noisy") == "both":
            target = norm2(target)
        else:
            mean = mix.mean(dim=(1, 2), keepdim=True)
            gain = level / (mix - mean).std(dim=(1, 2), keepdim=True).clamp(min=1e-5)
            target = (target - mean) * gain
        score_snr = 5.0 if fake_score_snr is None else float(fake_score_snr)
        sig = program.sampler_tables(self.model_config(), n_steps)[0].to(device=mix.device, dtype=mix.dtype)
        delta_t = 1.0 / (n_steps - 1)
        gamma = (float(self.diff_kwargs.sigma_max) / float(self.diff_kwargs.sigma_min)) ** -delta_t
        eta = 1.0 - gamma ** epsilon
        beta = math.sqrt(1.0 - gamma ** (2.0 * (epsilon - 1.0)))

        def draw(i):
            if noise is not None:
                return noise[i].to(mix).reshape(target.shape)
            return torch.randn(target.shape, dtype=mix.dtype, device=mix.device, generator=rng)

        def score(x, s):
            true = -(x - target) / s ** 2
            rms = true.square().mean().sqrt() * 10.0 ** (-score_snr / 20.0)
            return true + torch.randn(true.shape, dtype=true.dtype, device=true.device, generator=rng) * rms

        x = draw(0) * sig[0]
        for n in range(n_steps - 1):
            sc = score(x, sig[n])
            x = x + sig[n] ** 2 * eta * sc + beta * sig[n + 1] * draw(n + 1)
        x = x + sig[-1] ** 2 * score(x, sig[-1])
        x = self.unpad(x, pad)
        x = torch.nn.functional.pad(x, (0, mix_len - x.shape[-1]))
        if keep_rms:
            x = x * (mix_rms / x.square().mean(dim=(-2, -1), keepdim=True).sqrt().clamp(min=1e-5))
        scale = x.abs().max(dim=-1, keepdim=True).values
        return torch.where(scale > 1.0, x / scale, x)


def signal_median(signal):
    """Ensemble member that is "the median" for the most samples (reference utils/stats.py:22-66): per sample,
    the sorted position holding the member whose index is closest to n/2; the member index that this position
    equals most often over the samples is returned whole.  signal: (ensemble, batch, ...) -> (batch, ...)."""
    shape = signal.shape
    flat = signal.flatten(start_dim=2)
    n = flat.shape[0]
    order = flat.sort(dim=0).indices
    pos = (order - n / 2).abs().min(dim=0).indices                     # (batch, samples)
    counts = torch.nn.functional.one_hot(pos, n).sum(dim=1)            # (batch, n)
    select = counts.argmax(dim=1)
    out = flat[select, torch.arange(flat.shape[1], device=flat.device)]
    return out.reshape(shape[1:])


class UniverseGAN(Universe):
    """UNIVERSE++ (reference universe_gan.py:62-151): for inference it only adds the
    `signal_decoupling_layer` parameters (used by use_aux_signal / warm_start) to the state dict
    and to the EMA parameter list."""

    def __init__(self, fs, normalization_norm, score_model, condition_model, diffusion, losses=None, training=None,
                 validation=None, optimizer=None, scheduler=None, grad_clipper=None, transform=None,
                 normalization_kwargs={}, detach_cond=False, edm=None):
        super().__init__(fs, normalization_norm, score_model, condition_model, diffusion, losses, training, validation,
                         optimizer, scheduler, grad_clipper, transform=transform, normalization_kwargs=normalization_kwargs,
                         detach_cond=detach_cond, edm=edm)

    def init_losses(self, score_model, condition_model, losses, training):
        if losses.get("use_signal_decoupling", False):
            self.signal_decoupling_layer = SignalDecouplingP(self.n_channels, losses.get("signal_decoupling_act", None))
        else:
            self.signal_decoupling_layer = None

    def aux_to_wav(self, y_aux):
        """signal_decoupling_layer (reference universe_gan.py:119-128,147-151; blocks.py:209-231): alias-free snake
        (2x sinc up-sampling, x + sin^2(a x) / a with a = exp(alpha), 2x down-sampling; bigvgan alias_free_act.py:8-30,
        snake.py:53-64) or PReLU, then Conv1d(C -> 1, k3, same)."""
        sdl = getattr(self, "signal_decoupling_layer", None)
        if sdl is None:
            return y_aux
        x = y_aux.float()
        act = getattr(sdl.prelu, "act", None)
        if act is not None:
            x = self._resample(x, act.upsample.kernel, 1, 2)
            a = torch.exp(act.act.alpha.to(x))[None, :, None]
            x = x + (1.0 / (a + 1e-9)) * torch.sin(x * a) ** 2
            x = self._resample(x, act.downsample.kernel, 2, 1)
        elif hasattr(sdl.prelu, "weight"):
            x = torch.nn.functional.prelu(x, sdl.prelu.weight.to(x))
        return torch.nn.functional.conv1d(x, sdl.conv.weight.to(x), sdl.conv.bias.to(x), padding="same")

    def model_parameters(self):
        params = itertools.chain(self.get_score_model().parameters(), self.condition_model.parameters())
        if getattr(self, "signal_decoupling_layer", None) is not None:
            params = itertools.chain(params, self.signal_decoupling_layer.parameters())
        return params
