"""TEST INFRASTRUCTURE ONLY (CPU oracle) -- never imported by the product path.

A CPU fp32 restatement, in plain functional torch, of the reference's
UNIVERSE / UNIVERSE++ `enhance()` hot path, operating directly on a reference
state dict (same key names).  Each function cites the reference lines it
follows (paths relative to /root/reference/open_universe).  Pinned against the
real reference by `oracle/gen_golden.py` -> `tests/golden/*.pt` (the reference
itself has no tests or golden vectors: SURVEY.md section 4).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module.
"""
import math

import torch
import torch.nn.functional as F

SQRT1_2 = 1.0 / math.sqrt(2.0)


# ----------------------------------------------------------------- parameters
def get_weight(sd, prefix):
    """Effective conv/linear weight. Old-style weight norm w = g * v / ||v|| with the
    norm over all dims but 0 (networks/universe/blocks.py:40-46)."""
    if prefix + ".weight_g" in sd:
        g, v = sd[prefix + ".weight_g"], sd[prefix + ".weight_v"]
        dims = tuple(range(1, v.ndim))
        return g * v / v.pow(2).sum(dim=dims, keepdim=True).sqrt()
    return sd[prefix + ".weight"]


def get_bias(sd, prefix):
    return sd.get(prefix + ".bias")


# ----------------------------------------------------------------- blocks.py
def low_pass(x, taps):
    """BinomialAntiAlias.forward, blocks.py:128-134."""
    c = x.shape[1]
    w = taps[None, None, :].expand(c, 1, taps.shape[0])
    return F.conv1d(x, w, padding="same", groups=c)


def prelu_conv(sd, prefix, x, stride=1, transpose=False, same=False):
    """PReLU_Conv.forward, blocks.py:209-231 (act_type == 'prelu')."""
    if not transpose and x.shape[-1] % stride != 0:
        x = F.pad(x, (0, stride - x.shape[-1] % stride))
    x = F.prelu(x, sd[prefix + ".prelu.weight"])
    aa = prefix + ".low_pass_filter.weights" in sd
    w = get_weight(sd, prefix + ".conv")
    b = get_bias(sd, prefix + ".conv")
    if aa and not transpose:
        x = low_pass(x, sd[prefix + ".low_pass_filter.weights"])
    if transpose:
        x = F.conv_transpose1d(x, w, b, stride=stride)
    else:
        x = F.conv1d(x, w, b, stride=stride, padding="same" if same else 0)
    if aa and transpose:
        x = low_pass(x, sd[prefix + ".low_pass_filter.weights"])
    if prefix + ".bias" in sd:
        x = x + sd[prefix + ".bias"].reshape(1, -1, 1)
    return x


def film(x, y):
    """blocks.py:57-63."""
    c = x.shape[1]
    return y[:, :c, None] * x + y[:, c:, None]


def conv_block(sd, prefix, h, rate=None, direction="none", noise_cond=None, input_cond=None, res=None, length=None):
    """ConvBlock.forward, blocks.py:331-416. Returns (h_next, v_out, cond_out)."""
    if direction == "up":
        if length is not None and rate * h.shape[-1] < length:
            h = F.pad(h, (0, 1))
        h = prelu_conv(sd, prefix + ".rate_change_conv", h, stride=rate, transpose=True)
        if length is not None:
            h = F.pad(h, (0, length - h.shape[-1]))
    if res is not None:
        h = (h + res) * SQRT1_2
    cond_out = prelu_conv(sd, prefix + ".conv1", h, same=True)
    c = cond_out
    if input_cond is not None:
        c = (cond_out + input_cond) * SQRT1_2
    if noise_cond is not None:
        c = film(c, noise_cond)
    c = prelu_conv(sd, prefix + ".conv2", c, same=True)
    c = prelu_conv(sd, prefix + ".conv3", c, same=True)
    v_out = (h + c) * SQRT1_2
    if direction == "down":
        v_pad = v_out
        if h.shape[-1] % rate != 0:
            v_pad = F.pad(v_out, (0, rate - h.shape[-1] % rate))
        return prelu_conv(sd, prefix + ".rate_change_conv", v_pad, stride=rate), v_out, cond_out
    return v_out, v_out, cond_out


# ----------------------------------------------------------------- GRU
def gru_explicit(x, w_ih, w_hh, b_ih, b_hh, reverse=False):
    """One direction of torch.nn.GRU written out (gate order r, z, n;
    n = tanh(W_in x + b_in + r * (W_hn h + b_hn))). x: [B, T, I] -> [B, T, H]."""
    B, T, _ = x.shape
    H = w_hh.shape[1]
    gi = x @ w_ih.t() + b_ih
    h = x.new_zeros(B, H)
    out = x.new_zeros(B, T, H)
    steps = range(T - 1, -1, -1) if reverse else range(T)
    for t in steps:
        gh = h @ w_hh.t() + b_hh
        r = torch.sigmoid(gi[:, t, :H] + gh[:, :H])
        z = torch.sigmoid(gi[:, t, H : 2 * H] + gh[:, H : 2 * H])
        n = torch.tanh(gi[:, t, 2 * H :] + r * gh[:, 2 * H :])
        h = (1 - z) * n + z * h
        out[:, t] = h
    return out


def gru(sd, prefix, x, num_layers, explicit=False):
    """Bidirectional batch_first torch.nn.GRU (score.py:84-90, condition.py:173-179).
    x: [B, C, T] -> [B, C, T]."""
    x = x.transpose(1, 2)
    if explicit:
        for l in range(num_layers):
            outs = []
            for suf, rev in (("", False), ("_reverse", True)):
                outs.append(
                    gru_explicit(
                        x,
                        sd[f"{prefix}.weight_ih_l{l}{suf}"],
                        sd[f"{prefix}.weight_hh_l{l}{suf}"],
                        sd[f"{prefix}.bias_ih_l{l}{suf}"],
                        sd[f"{prefix}.bias_hh_l{l}{suf}"],
                        reverse=rev,
                    )
                )
            x = torch.cat(outs, dim=-1)
    else:
        C = x.shape[-1]
        m = torch.nn.GRU(C, C // 2, num_layers=num_layers, bidirectional=True, batch_first=True).to(x.device)
        m.load_state_dict({k[len(prefix) + 1 :]: v for k, v in sd.items() if k.startswith(prefix + ".")})
        with torch.no_grad():
            x, _ = m(x)
    return x.transpose(1, 2)


# ----------------------------------------------------------------- sigma_block.py
def time_embedding(sd, prefix, log10_sigma, simple, n_dim):
    """SimpleTimeEmbedding.forward sigma_block.py:73-78 / SigmaBlock.forward :50-57."""
    if simple:
        t = torch.arange(n_dim // 2)
        f = 0.5 * torch.sigmoid(sd[prefix + ".weight"] * log10_sigma[:, None] + sd[prefix + ".bias"])
        p = 2.0 * math.pi * f * t
        return torch.cat([torch.sin(p), torch.cos(p)], dim=-1)
    p = 2.0 * math.pi * sd[prefix + ".freq"][None, :] * log10_sigma[:, None]
    g = torch.cat([torch.sin(p), torch.cos(p)], dim=-1)
    for i in (1, 2, 3):
        g = F.linear(g, sd[f"{prefix}.layer{i}.lin.weight"], sd[f"{prefix}.layer{i}.lin.bias"])
        g = F.prelu(g, sd[f"{prefix}.layer{i}.prelu.weight"])
    return g


# ----------------------------------------------------------------- score.py
def score_network(sd, prefix, cfg, x, sigma, cond, explicit_gru=False):
    """ScoreNetwork.forward score.py:278-298 (+ ScoreEncoder :105-128, ScoreDecoder :197-211)."""
    rates = list(cfg["rate_factors"])
    simple = cfg.get("time_embedding") == "simple"
    g = time_embedding(sd, prefix + ".sigma_block", torch.log10(sigma), simple, cfg.get("noise_cond_dim", 512))
    x = F.conv1d(x, sd[prefix + ".input_conv.weight"], sd[prefix + ".input_conv.bias"], padding="same")
    n_ds = len(rates) + (1 if cfg.get("extra_conv_block") else 0)
    residuals, lengths = [], []
    for i in range(n_ds):
        nc = F.linear(g, get_weight(sd, f"{prefix}.encoder.cond_proj.{i}"), get_bias(sd, f"{prefix}.encoder.cond_proj.{i}"))
        lengths.append(x.shape[-1])
        if i < len(rates):
            x, res, _ = conv_block(sd, f"{prefix}.encoder.ds_modules.{i}", x, rates[i], "down", noise_cond=nc)
        else:
            x, res, _ = conv_block(sd, f"{prefix}.encoder.ds_modules.{i}", x, noise_cond=nc)
        residuals.append(res)
    x = gru(sd, prefix + ".encoder.gru", x, 1, explicit=explicit_gru)
    residuals, lengths = residuals[::-1], lengths[::-1]
    up_rates = ([None] if cfg.get("extra_conv_block") else []) + rates[::-1]
    for i, r in enumerate(up_rates):
        nc = F.linear(g, get_weight(sd, f"{prefix}.decoder.noise_cond_proj.{i}"), get_bias(sd, f"{prefix}.decoder.noise_cond_proj.{i}"))
        sc = F.conv1d(cond[i], get_weight(sd, f"{prefix}.decoder.signal_cond_proj.{i}"), get_bias(sd, f"{prefix}.decoder.signal_cond_proj.{i}"))
        x, _, _ = conv_block(
            sd, f"{prefix}.decoder.up_modules.{i}", x, r, "up" if r else "none",
            noise_cond=nc, input_cond=sc, res=residuals[i], length=lengths[i],
        )
    x = F.prelu(x, sd[prefix + ".prelu.weight"])
    x = prelu_conv(sd, prefix + ".output_conv", x, same=True)
    return x


# ----------------------------------------------------------------- condition.py
def mel_spec(sd, prefix, x, ds_factor, oversample):
    """MelAdapter.compute_mel_spec condition.py:92-108 with torchaudio's
    MelSpectrogram(center=False, power=2) written out; window and filterbank are
    the buffers stored in the state dict."""
    n_fft = oversample * ds_factor
    pad_tot = n_fft - ds_factor
    pl, pr = pad_tot // 2, pad_tot - pad_tot // 2
    r = x.shape[-1] % ds_factor
    x = F.pad(x, (pl, (ds_factor - r if r else 0) + pr))
    win = sd[prefix + ".mel_spec.spectrogram.window"]
    fb = sd[prefix + ".mel_spec.mel_scale.fb"]
    frames = x.squeeze(1).unfold(-1, n_fft, ds_factor)  # [B, F, n_fft]
    spec = torch.fft.rfft(frames * win, dim=-1).abs().pow(2)  # [B, F, n_fft/2+1]
    mel = (spec @ fb).transpose(1, 2)  # [B, n_mels, F]
    norm = (mel**2).sum(dim=-2, keepdim=True).mean(dim=-1, keepdim=True).sqrt()
    return mel / norm.clamp(min=1e-5)


def conditioner(sd, prefix, cfg, x, explicit_gru=False, return_all=False):
    """ConditionerNetwork.forward condition.py:346-377 (train=True): returns the list of
    conditions (and y_hat, h when return_all)."""
    rates = list(cfg["rate_factors"])
    tot_ds = math.prod(rates)
    extra = cfg.get("extra_conv_block")
    # MelAdapter.forward :110-114
    m = mel_spec(sd, prefix + ".input_mel", x, tot_ds, cfg.get("n_mel_oversample", 4))
    m = F.conv1d(m, get_weight(sd, prefix + ".input_mel.conv"), get_bias(sd, prefix + ".input_mel.conv"), padding="same")
    x_mel, _, _ = conv_block(sd, prefix + ".input_mel.conv_block", m)
    h = F.conv1d(x, get_weight(sd, prefix + ".input_conv"), get_bias(sd, prefix + ".input_conv"), padding="same")
    # ConditionerEncoder.forward :189-220
    outputs, lengths = [], []
    n_ds = len(rates) + (1 if extra else 0)
    for i in range(n_ds):
        lengths.append(h.shape[-1])
        if i < len(rates):
            h, res, _ = conv_block(sd, f"{prefix}.encoder.ds_modules.{i}", h, rates[i], "down")
        else:
            h, res, _ = conv_block(sd, f"{prefix}.encoder.ds_modules.{i}", h)
        if i < len(rates) - 1:
            st = math.prod(rates[i:])
            outputs.append(prelu_conv(sd, f"{prefix}.encoder.st_convs.{i}", res, stride=st))
    outputs.append(h)
    out = x_mel
    for o in outputs:
        out = out + o
    out = out * (1.0 / math.sqrt(len(outputs) + 1))
    out, _, _ = conv_block(sd, prefix + ".encoder.conv_block1", out)
    res = out
    out = gru(sd, prefix + ".encoder.gru", out, 2, explicit=explicit_gru)
    if cfg.get("encoder_gru_residual"):
        out = (out + res) * SQRT1_2
    out, _, _ = conv_block(sd, prefix + ".encoder.conv_block2", out)
    latent = out
    lengths = lengths[::-1]
    # ConditionerDecoder.forward :264-270
    y, _, _ = conv_block(sd, prefix + ".decoder.input_conv_block", out)
    up_rates = ([None] if extra else []) + rates[::-1]
    conds = []
    for i, r in enumerate(up_rates):
        y, _, c = conv_block(sd, f"{prefix}.decoder.up_modules.{i}", y, r, "up" if r else "none", length=lengths[i])
        conds.append(c)
    if return_all:
        return conds, y, latent
    return conds


# ----------------------------------------------------------------- universe.py
def sampler_constants(cfg, n_steps=None, epsilon=None):
    """universe.py:301-311, 380-386: (sigma[n], eta, beta)."""
    d = cfg["diffusion"]
    n_steps = d["n_steps"] if n_steps is None else n_steps
    epsilon = d["epsilon"] if epsilon is None else epsilon
    delta_t = 1.0 / (n_steps - 1)
    gamma = (d["sigma_max"] / d["sigma_min"]) ** -delta_t
    eta = 1 - gamma**epsilon
    beta = math.sqrt(1 - gamma ** (2 * (epsilon - 1.0)))
    time = torch.linspace(0, 1, n_steps).flip(dims=[0])
    sigma = d["sigma_min"] * (d["sigma_max"] / d["sigma_min"]) ** time
    return sigma, eta, beta


def normalize(mix, level_db):
    """utils/norm.py:47-87 with norm=2, zero_mean=True (target None)."""
    mean = mix.mean(dim=(1, 2), keepdim=True)
    mix = mix - mean
    gain = 10 ** (level_db / 20.0) / mix.std(dim=(1, 2), keepdim=True).clamp(min=1e-5)
    return mix * gain


def resample(x, kernel, orig, new):
    """torchaudio.transforms.Resample.forward with its stored filter bank (the kernels are buffers of the
    checkpoint: bigvgan/alias_free_act.py:21-22): pad (w, w + orig), conv with stride orig, interleave phases."""
    width = (kernel.shape[-1] - orig) // 2
    shape = x.shape
    w = F.pad(x.reshape(-1, shape[-1]), (width, width + orig))
    y = F.conv1d(w[:, None], kernel, stride=orig).transpose(1, 2).reshape(w.shape[0], -1)
    y = y[..., : math.ceil(new * shape[-1] / orig)]
    return y.reshape(shape[:-1] + y.shape[-1:])


def aux_to_wav(sd, y_aux, prefix="signal_decoupling_layer"):
    """UniverseGAN.aux_to_wav universe_gan.py:147-151 -> PReLU_Conv(C, 1, 3, act_type) blocks.py:209-231 with the
    alias-free snake of bigvgan/snake.py:53-64,131-160 + alias_free_act.py:24-30 (alpha in log scale)."""
    if prefix + ".conv.weight" not in sd:
        return y_aux
    x = y_aux
    if prefix + ".prelu.act.act.alpha" in sd:
        x = resample(x, sd[prefix + ".prelu.act.upsample.kernel"], 1, 2)
        a = torch.exp(sd[prefix + ".prelu.act.act.alpha"])[None, :, None]
        x = x + (1.0 / (a + 1e-9)) * torch.sin(x * a) ** 2
        x = resample(x, sd[prefix + ".prelu.act.downsample.kernel"], 2, 1)
    elif prefix + ".prelu.weight" in sd:
        x = F.prelu(x, sd[prefix + ".prelu.weight"])
    return F.conv1d(x, sd[prefix + ".conv.weight"], sd[prefix + ".conv.bias"], padding="same")


def enhance(sd, cfg, mix, noises, n_steps=None, epsilon=None, keep_rms=False, explicit_gru=False, return_steps=False,
            use_aux_signal=False, warm_start=None):
    """Universe.enhance universe.py:231-375 (target None, no ensemble) with the noise sequence injected.
    mix: [B, T]; noises: list of tensors [B, 1, Lp], consumed in the reference's order (n_steps of them on the
    default path; with warm_start = n the first one initialises x at step n). Returns [B, T]."""
    edm = cfg.get("edm") is not None
    sp = "_edm_model" if edm else "score_model"
    scfg = cfg["score_model"]
    tot_ds = math.prod(scfg["rate_factors"])
    mix = mix[:, None, :]
    mix_rms = mix.square().mean(dim=(-2, -1), keepdim=True).sqrt()
    T = mix.shape[-1]
    pad = tot_ds - T % tot_ds
    mix = F.pad(mix, (pad // 2, pad - pad // 2))
    level_db = cfg["normalization_kwargs"]["level_db"]
    mix = normalize(mix, level_db)
    sigma, eta, beta = sampler_constants(cfg, n_steps, epsilon)
    n = sigma.shape[0]
    B = mix.shape[0]
    cond, y_hat, _ = conditioner(sd, "condition_model", cfg["condition_model"], mix, explicit_gru=explicit_gru, return_all=True)

    def score_fn(x, s):
        sv = s.expand(B)
        if not edm:
            net = score_network(sd, sp, scfg, x, sv, cond, explicit_gru)
            return net, net
        # _edm_score_wrapper universe.py:197-209, _get_edm_weights :175-189
        sigma_data = 10.0 ** (cfg["edm"].get("data_level_db", level_db) / 20.0)
        sn = (s**2 + sigma_data**2) ** 0.5
        w_skip, w_in, w_out = sigma_data**2 / (s**2 + sigma_data**2), 1.0 / sn, s * sigma_data / sn
        net = score_network(sd, sp, scfg, w_in * x, cfg["edm"]["noise"] * sv, cond, explicit_gru)
        est = w_skip * x + w_out * net
        return (est - x) / s**2, net

    steps = []
    noises = list(noises)
    if use_aux_signal:                       # universe.py:317-319
        x = aux_to_wav(sd, y_hat)
    else:
        n0 = 0 if warm_start is None else int(warm_start)
        x = noises.pop(0) * sigma[n0]        # universe.py:324-331
        if warm_start is not None:
            x = aux_to_wav(sd, y_hat) + x
        for i in range(n0, n - 1):
            score, net = score_fn(x, sigma[i])
            steps.append(net)
            z = noises.pop(0) * sigma[i + 1]
            x = x + sigma[i] ** 2 * eta * score + beta * z
        score, net = score_fn(x, sigma[-1])
        steps.append(net)
        x = x + sigma[-1] ** 2 * score
    x = x[..., pad // 2 : -(pad - pad // 2)]
    if keep_rms:
        x_rms = x.square().mean(dim=(-2, -1), keepdim=True).sqrt().clamp(min=1e-5)
        x = x * (mix_rms / x_rms)
    scale = x.abs().max(dim=-1, keepdim=True).values
    x = torch.where(scale > 1.0, x / scale, x)
    x = x[:, 0, :]
    if return_steps:
        return x, steps
    return x


def si_sdr(est, ref):
    """10 log10(|a r|^2 / |e - a r|^2), a = <e, r> / |r|^2 (SURVEY.md 8d parity gate)."""
    a = (est * ref).sum(-1, keepdim=True) / ref.pow(2).sum(-1, keepdim=True)
    return 10 * torch.log10((a * ref).pow(2).sum(-1) / (est - a * ref).pow(2).sum(-1))


def rel_l2(a, b):
    return ((a - b).norm() / b.norm()).item()
