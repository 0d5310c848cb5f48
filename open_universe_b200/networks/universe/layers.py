"""Parameter containers that reproduce the reference's state-dict layout and
parameter registration order (needed for `load_state_dict` on reference checkpoints and
for the positional EMA overlay: reference universe.py:130-133, universe_gan.py:138-145).

These modules hold weights only.  There is no PyTorch forward path: all arithmetic runs in
the native sm_100a kernels (csrc/), driven by `Universe.enhance`.
"""
import math

import torch
from torch import nn


def _p(*shape):
    return nn.Parameter(torch.empty(*shape))


class WConv(nn.Module):
    """Conv1d / ConvTranspose1d / Linear weights; with old-style weight norm the parameters
    are (bias,) weight_g, weight_v with g of shape (w.shape[0], 1, ...) (reference blocks.py:40-46)."""

    def __init__(self, w_shape, bias=True, weight_norm=False, bias_len=None):
        super().__init__()
        w_shape = tuple(w_shape)
        bias_len = w_shape[0] if bias_len is None else bias_len
        self.weight_norm = weight_norm
        if weight_norm:
            if bias:
                self.bias = _p(bias_len)
            self.weight_g = _p(w_shape[0], *([1] * (len(w_shape) - 1)))
            self.weight_v = _p(*w_shape)
        else:
            self.weight = _p(*w_shape)
            if bias:
                self.bias = _p(bias_len)
        self.reset()

    def reset(self):
        w = self.weight_v if self.weight_norm else self.weight
        fan_in = w[0].numel() if w.ndim > 1 else w.numel()
        bound = 1.0 / math.sqrt(max(fan_in, 1))
        with torch.no_grad():
            w.uniform_(-bound, bound)
            if self.weight_norm:
                dims = tuple(range(1, w.ndim))
                self.weight_g.copy_(w.pow(2).sum(dim=dims, keepdim=True).sqrt())
            if hasattr(self, "bias"):
                self.bias.uniform_(-bound, bound)


class PReLUw(nn.Module):
    def __init__(self):
        super().__init__()
        self.weight = nn.Parameter(torch.full((1,), 0.25))


class LowPass(nn.Module):
    """BinomialAntiAlias buffer (reference blocks.py:66-72,123-126): unit-RMS binomial taps."""

    def __init__(self, kernel_size):
        super().__init__()
        row = torch.tensor([math.comb(kernel_size - 1, i) for i in range(kernel_size)], dtype=torch.float64)
        row = row / row.square().mean().sqrt()
        row = row.float()
        self.register_buffer("weights", row / row.square().mean().sqrt())


class PReLUConv(nn.Module):
    """State of the reference's PReLU_Conv (blocks.py:137-207)."""

    def __init__(self, cin, cout, k, transpose=False, weight_norm=False, antialias=False):
        super().__init__()
        bias = True
        if antialias:
            self.bias = nn.Parameter(torch.zeros(cout))
            self.low_pass_filter = LowPass(2 * k + 1)
            bias = False
        self.prelu = PReLUw()
        w_shape = (cin, cout, k) if transpose else (cout, cin, k)
        self.conv = WConv(w_shape, bias=bias, weight_norm=weight_norm, bias_len=cout)


class ConvBlockP(nn.Module):
    """State of the reference's ConvBlock (blocks.py:240-329)."""

    def __init__(self, c, rate=None, direction="none", weight_norm=False, antialias=False):
        super().__init__()
        if direction == "down":
            self.rate_change_conv = PReLUConv(c, 2 * c, rate, weight_norm=weight_norm, antialias=antialias)
        elif direction == "up":
            self.rate_change_conv = PReLUConv(2 * c, c, rate, transpose=True, weight_norm=weight_norm, antialias=antialias)
        self.conv1 = PReLUConv(c, c, 5, weight_norm=weight_norm)
        self.conv2 = PReLUConv(c, c, 3, weight_norm=weight_norm)
        self.conv3 = PReLUConv(c, c, 3, weight_norm=weight_norm)


class LinearPReLU(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.prelu = PReLUw()
        self.lin = WConv((o, i))


class SigmaBlockP(nn.Module):
    """reference sigma_block.py:36-48."""

    def __init__(self, n_rff=32, n_dim=256, scale=16):
        super().__init__()
        self.register_buffer("freq", scale * torch.zeros(n_rff).normal_())
        self.layer1 = LinearPReLU(2 * n_rff, 4 * n_rff)
        self.layer2 = LinearPReLU(4 * n_rff, 8 * n_rff)
        self.layer3 = LinearPReLU(8 * n_rff, n_dim)


class SimpleTimeEmbeddingP(nn.Module):
    """reference sigma_block.py:66-71."""

    def __init__(self, n_dim=256):
        super().__init__()
        self.weight = nn.Parameter(torch.zeros((1, 1)))
        self.bias = nn.Parameter(torch.zeros((1, 1)))
        self.n_dim = n_dim


def hann_window(n):
    return (0.5 - 0.5 * torch.cos(2.0 * math.pi * torch.arange(n, dtype=torch.float64) / n)).float()


def mel_filterbank(n_freqs, n_mels, sample_rate, f_min=0.0, f_max=None):
    """HTK-scale triangular filterbank without normalisation: the buffer that
    torchaudio.transforms.MelSpectrogram registers (reference condition.py:75-81)."""
    f_max = float(sample_rate // 2) if f_max is None else f_max
    all_freqs = torch.linspace(0, sample_rate // 2, n_freqs)
    m_min = 2595.0 * math.log10(1.0 + f_min / 700.0)
    m_max = 2595.0 * math.log10(1.0 + f_max / 700.0)
    m_pts = torch.linspace(m_min, m_max, n_mels + 2)
    f_pts = 700.0 * (10 ** (m_pts / 2595.0) - 1.0)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts.unsqueeze(0) - all_freqs.unsqueeze(1)
    down = (-1.0 * slopes[:, :-2]) / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    return torch.max(torch.zeros(1), torch.min(down, up))


class _Buf(nn.Module):
    def __init__(self, name, value):
        super().__init__()
        self.register_buffer(name, value)


class MelSpecP(nn.Module):
    def __init__(self, n_fft, n_mels, sample_rate):
        super().__init__()
        self.spectrogram = _Buf("window", hann_window(n_fft))
        self.mel_scale = _Buf("fb", mel_filterbank(n_fft // 2 + 1, n_mels, sample_rate))


class MelAdapterP(nn.Module):
    """reference condition.py:68-90 (sample_rate is hard-coded to 24000 there, :76)."""

    def __init__(self, n_mels, output_channels, ds_factor, oversample=2, weight_norm=False):
        super().__init__()
        self.mel_spec = MelSpecP(oversample * ds_factor, n_mels, 24000)
        self.conv = WConv((output_channels, n_mels, 3), weight_norm=weight_norm)
        self.conv_block = ConvBlockP(output_channels, weight_norm=weight_norm)


def sinc_hann_kernel(orig, new, zeros=6, rolloff=0.99):
    """Polyphase windowed-sinc resampling filter bank of torchaudio 2.x `Resample(orig, new)` (defaults
    lowpass_filter_width 6, rolloff 0.99, sinc_interp_hann), restated from its published definition: phase i of
    `new`, taps j in [-width, width + orig): t = (j / orig - i / new) * base, base = min(orig, new) * rolloff,
    h = sinc(pi t) cos^2(pi t / (2 zeros)) base / orig on |t| <= zeros.  Returns (kernel [new, 1, 2 width + orig], width).
    The reference stores these as buffers of bigvgan's Activation1d (alias_free_act.py:21-22)."""
    base = min(orig, new) * rolloff
    width = math.ceil(zeros * orig / base)
    j = torch.arange(-width, width + orig, dtype=torch.float64)[None, None, :] / orig
    i = torch.arange(0, -new, -1, dtype=torch.float64)[:, None, None] / new
    t = ((i + j) * base).clamp(-zeros, zeros)
    window = torch.cos(t * math.pi / zeros / 2) ** 2
    t = t * math.pi
    k = torch.where(t == 0, torch.ones_like(t), torch.sin(t) / t) * window * (base / orig)
    return k.float(), width


class SnakeActP(nn.Module):
    """State of bigvgan AliasFreeSnake inside `signal_decoupling_layer`
    (reference bigvgan/snake.py, alias_free_act.py:8-30): alpha + the two resampling kernels."""

    def __init__(self, channels):
        super().__init__()
        self.act = nn.Module()
        self.act.alpha = nn.Parameter(torch.zeros(channels))
        self.upsample = _Buf("kernel", sinc_hann_kernel(1, 2)[0])      # [2, 1, 15]
        self.downsample = _Buf("kernel", sinc_hann_kernel(2, 1)[0])    # [1, 1, 28]


class SignalDecouplingP(nn.Module):
    """reference universe_gan.py:119-128: PReLU_Conv(C, 1, 3, act_type='snake')."""

    def __init__(self, channels, act):
        super().__init__()
        self.prelu = nn.Module()
        if act in ("snake", "snakebeta"):
            self.prelu.act = SnakeActP(channels)
        elif act == "prelu":
            self.prelu = PReLUw()
        self.conv = WConv((1, channels, 3))
