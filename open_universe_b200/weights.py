"""Weight folding and packing: reference state dict -> device weight blob.

All transformations are exact algebra done once at load time in fp32:
  * old-style weight norm  w = g * v / ||v||  (norm over all dims but 0;
    reference networks/universe/blocks.py:40-46),
  * the binomial anti-aliasing low-pass of PReLU_Conv (blocks.py:217-229)
    folded into the strided / transposed rate-change convolutions,
  * re-expression of every dense layer as a multi-tap GEMM
        D[row, n] = sum_{src s, shift q, c}  A_s[row + q, c] * W4[n, s, q, c]
    and packing of W4 into bf16 images [n_tile][half 2][s][group][q][KC/8][N_TILE/2][8]
    (the no-swizzle K-major UMMA layout consumed by csrc/gemm_tc.cuh; the GEMM kernel runs on SM pairs,
    tcgen05 cta_group::2, and the CTA of rank h holds columns [h N_TILE/2, (h+1) N_TILE/2) of every image),
  * GRU recurrent weights into per-CTA K-major UMMA slices (csrc/gru.cuh).
"""
import math
import os

import numpy as np
import torch


def eff_weight(sd, prefix):
    if prefix + ".weight_g" in sd:
        g, v = sd[prefix + ".weight_g"].float(), sd[prefix + ".weight_v"].float()
        dims = tuple(range(1, v.ndim))
        return g * v / v.pow(2).sum(dim=dims, keepdim=True).sqrt()
    return sd[prefix + ".weight"].float()


TILE_M = 256          # csrc/gemm_tc.cuh kTileM: rows of a tile of the SM pair
SUB_M = 128           # ... of which every CTA of the pair holds 128 (kSubM)
SMEM_BUDGET = 227 * 1024 - (4096 + 6144 + 640 + 640 + 16 + 512 + 1024)  # gemm_plan_pipeline(): 227 KB minus the fixed regions
# widest N tile.  256 columns run the pair's MMA at the same rate as 128 (tools/mma_bench.cu) and halve the A re-reads,
# but leave two (msub = 1) accumulator stages and halve the tile count of the short level-3 / 4 launches; measured on the
# whole model (tools/ab_msub.sh, eager pass, one lane): 128 / 256 = 46.97 / 47.50 ms with 256-row CTA tiles for N <= 128
NTILE_MAX = int(os.environ.get("UV_NTILE_MAX", "128"))


def _kc_options(c_in):
    cand = [64, 48, 32, 16] + ([c_in] if c_in <= 96 else [])
    return sorted({k for k in cand if c_in % k == 0 and k % 16 == 0}, reverse=True)


def pick_kc(c_in, n_total=None, n_tile=None, n_src=1, n_shift=1, reserve=0):
    """Channels per K group.  Largest slab that divides C_in, except that a layer whose whole weight
    set can stay resident in shared memory (single N tile) takes the largest KC that still leaves
    room for three A stages next to it (mirrors gemm_plan_pipeline in csrc/gemm_tc.cuh)."""
    opts = _kc_options(c_in)
    if n_total is not None and n_total == n_tile:
        total_b = (n_total * n_src * n_shift * c_in * 2 + n_tile * 32) // 2   # per CTA: half of the columns
        # (a weight set that leaves no room for a deep ring of full-width A slabs is streamed instead: with 200 KB of
        # resident weights the 256-channel layers fell back to KC = 16 and ran 40 % slower)
        for kc in (opts if total_b <= 120 * 1024 else []):
            a_stage = (kc // 8) * (SUB_M + n_shift - 1) * 16
            if total_b + 3 * a_stage + 256 + reserve <= SMEM_BUDGET:
                return kc
    return opts[0]


def pick_ntile(c_out):
    """Largest multiple of 16 that divides C_out and leaves two TMEM accumulator stages (<= 256 columns)."""
    for n in range(min(NTILE_MAX, c_out), 15, -16):
        if c_out % n == 0:
            return n
    raise ValueError(f"no N tile for C_out={c_out}")


MULTI_PHASE = os.environ.get("UV_MULTI_PHASE", "1") != "0"


def pick_ntile_up(c_out, up):
    """N tile of a transposed conv (N = up x C_out columns, phase major).  A tile may span several whole phases:
    the A slabs are then read once for k phases instead of once per phase and the MMAs are k times wider
    (csrc/gemm_tc.cuh steps the output row at every phase boundary of a tile)."""
    n = pick_ntile(c_out)
    if not MULTI_PHASE or n != c_out:
        return n
    k = max(d for d in range(1, up + 1) if up % d == 0 and c_out * d <= NTILE_MAX)
    return c_out * k


def to_bf16_bits(x):
    return x.contiguous().to(torch.bfloat16).view(torch.int16).numpy()


class Blob:
    """Append-only byte blob with 256-byte aligned entries."""

    def __init__(self):
        self.parts = []
        self.size = 0

    def add(self, arr):
        b = arr.tobytes() if isinstance(arr, np.ndarray) else bytes(arr)
        off = self.size
        pad = (-len(b)) % 256
        self.parts.append(b + b"\0" * pad)
        self.size += len(b) + pad
        return off

    def add_f32(self, t):
        return self.add(np.ascontiguousarray(t.detach().float().cpu().numpy(), dtype=np.float32))

    def bytes(self):
        return b"".join(self.parts)


# ---------------------------------------------------------------- W4 builders
def w4_plain(w):
    """Conv1d weight [Co, Ci, k] ('same', odd k) -> W4 [Co, 1, k, Ci], smin = -(k//2)."""
    return w.permute(0, 2, 1)[:, None].contiguous(), -(w.shape[2] // 2)


S2D = os.environ.get("UV_S2D", "1") != "0"


def w4_same_s2d(w):
    """'same' conv [Co, Ci, k <= 5] on tensors stored phase-split by 2 ("space to depth"): row m of phase plane ph is
    sample 2m + ph.  Output column n = ph Co + co (the two output phases of a row), source s = input phase plane,
    shift q in {-1, 0, 1}: input sample 2(m + q) + s, i.e. tap t = 2q + s - ph of the original kernel (zero where
    |t| exceeds the kernel).  -> W4 [2 Co, 2, 3, Ci], smin = -1.  A k5 conv is then 2 sources x 3 shifts = 6 K blocks
    for 2 output samples instead of 5 per sample, and the tensor core re-reads each A slab 3 times, not 5."""
    co, ci, k = w.shape
    pad = k // 2
    assert k % 2 == 1 and k <= 5
    w4 = torch.zeros(2 * co, 2, 3, ci)
    for ph in range(2):
        for s_ in range(2):
            for qi, q in enumerate((-1, 0, 1)):
                t = 2 * q + s_ - ph
                if abs(t) <= pad:
                    w4[ph * co : (ph + 1) * co, s_, qi, :] = w[:, :, t + pad]
    return w4, -1


def w4_down(w, lp=None):
    """Strided conv (k = stride = r), optionally preceded by the 'same' low-pass lp
    (2r+1 taps): y[m] = sum_tau W'[tau] v[r m + tau - r], W' = W (*) lp (3r taps).
    Source t = phase plane v[r m' + t], shift q - 1:  W4[co, t, q, ci] = W'[co, ci, r q + t]."""
    co, ci, r = w.shape
    if lp is None:
        return w.permute(0, 2, 1)[:, :, None].contiguous(), 0
    wf = torch.zeros(co, ci, 3 * r)
    for t in range(r):
        wf[:, :, t : t + 2 * r + 1] += w[:, :, t : t + 1] * lp[None, None, :]
    return wf.view(co, ci, 3, r).permute(0, 3, 2, 1).contiguous(), -1


def w4_up(w, lp=None):
    """ConvTranspose1d (k = stride = r), weight [Ci, Co, r], optionally followed by the
    low-pass.  Output row r m + t is column n = t * Co + co of a GEMM on x[m + d]:
    W4[t*Co + co, 0, d+1, ci] = sum_j lp[j] W[ci, co, t + j - r (d+1)]."""
    ci, co, r = w.shape
    if lp is None:
        return w.permute(2, 1, 0).reshape(r * co, 1, 1, ci).contiguous(), 0
    w4 = torch.zeros(r, co, 1, 3, ci)
    for t in range(r):
        for j in range(2 * r + 1):
            dq, tt = divmod(t + j, r)
            w4[t, :, 0, dq, :] += lp[j] * w[:, :, tt].t()
    return w4.view(r * co, 1, 3, ci).contiguous(), -1


def pack_w4(w4, kc, n_tile, bias_n=None):
    """(n_tile here is the PACKING tile: half an N tile of the kernel, see add_gemm.)
    [N, S, Q, Ci] fp32 -> per N tile: bf16 images [S][Ci/kc][Q][kc/8][n_tile][8] followed (bias_n
    given: fp32 [N]) by the bias image [2][n_tile][8] whose K columns 0 / 1 carry the hi / lo bf16
    halves of the bias (it meets the kernel's constant ones tile, csrc/gemm_tc.cuh)."""
    n, s, q, ci = w4.shape
    assert n % n_tile == 0 and ci % kc == 0 and kc % 16 == 0 and n_tile % 8 == 0
    x = w4.view(n // n_tile, n_tile, s, q, ci // kc, kc // 8, 8)
    x = x.permute(0, 2, 4, 3, 5, 1, 6).reshape(n // n_tile, -1)  # nt, (s, g, q, j, n, e)
    if bias_n is not None:
        b = bias_n.float().view(n // n_tile, n_tile)
        hi = b.to(torch.bfloat16).float()
        img = torch.zeros(n // n_tile, 2, n_tile, 8)
        img[:, 0, :, 0] = hi
        img[:, 0, :, 1] = b - hi
        x = torch.cat([x, img.reshape(n // n_tile, -1)], dim=1)
    return to_bf16_bits(x)


def pack_w4_split(w4, kc, n_tile):
    """[N, S, Q, Ci] fp32 -> the "bf16x3" images of the fp32 tier's tensor-core GEMM: w = hi + lo with hi = bf16(w),
    lo = bf16(w - hi) (16 mantissa bits together), per N tile and (source, K group, shift) one item
    [hi | lo][kc/8][n_tile][8] bf16 in the no-swizzle K-major UMMA layout (one bulk copy per item)."""
    n, s, q, ci = w4.shape
    assert n % n_tile == 0 and ci % kc == 0 and kc % 16 == 0 and n_tile % 16 == 0
    hi = w4.to(torch.bfloat16).float()
    lo = (w4 - hi).to(torch.bfloat16).float()
    x = torch.stack([hi, lo]).view(2, n // n_tile, n_tile, s, q, ci // kc, kc // 8, 8)   # h, nt, n, s, q, g, j, e
    x = x.permute(1, 3, 5, 4, 0, 6, 2, 7).contiguous()                                       # nt, s, g, q, h, j, n, e
    return to_bf16_bits(x.reshape(-1))


def unpack_w4(bits, n, s, q, ci, kc, n_tile, has_bias=False):
    """Inverse of pack_w4 (used by the CPU emulator): -> fp32 [N, S, Q, Ci] (and the bias [N] the
    tensor core sees, hi + lo, when has_bias)."""
    x = torch.from_numpy(np.ascontiguousarray(bits)).view(torch.bfloat16).float()
    x = x.view(n // n_tile, -1)
    n_w = s * q * ci * n_tile
    w = x[:, :n_w].reshape(n // n_tile, s, ci // kc, q, kc // 8, n_tile, 8)
    w = w.permute(0, 5, 1, 3, 2, 4, 6).reshape(n, s, q, ci).contiguous()
    if not has_bias:
        return w
    img = x[:, n_w:].reshape(n // n_tile, 2, n_tile, 8)
    return w, (img[:, 0, :, 0] + img[:, 0, :, 1]).reshape(n)


GRU_NC = 8  # csrc/gru.cuh kGruNC: CTAs per cluster, each owning H/8 hidden units


def pack_gru_whh(w_hh_f, w_hh_b):
    """[3H, H] x 2 directions -> bf16 K-major UMMA images of the per-CTA slices
    [dir][rank 8][H/8][3U][8], U = H/8; slice row gate*U + u is W_hh row gate*H + rank*U + u."""
    out = []
    for w in (w_hh_f, w_hh_b):
        m, h = w.shape
        u = h // GRU_NC
        x = w.float().view(3, GRU_NC, u, h // 8, 8)          # gate, rank, u, j, e
        out.append(x.permute(1, 3, 0, 2, 4).reshape(GRU_NC, h // 8, 3 * u, 8))
    return to_bf16_bits(torch.stack(out))


def unpack_gru_whh(bits, h):
    """Inverse of pack_gru_whh (CPU emulator): -> fp32 [2, 3H, H]."""
    u = h // GRU_NC
    x = torch.from_numpy(np.ascontiguousarray(bits)).view(torch.bfloat16).float()
    x = x.view(2, GRU_NC, h // 8, 3, u, 8)                    # dir, rank, j, gate, u, e
    return x.permute(0, 3, 1, 4, 2, 5).reshape(2, 3 * h, h).contiguous()


def dft_basis(window, n_fft):
    """Windowed real-DFT basis f32 [n_fft][2 * nbp]: cos | sin, nbp = bins padded to 32."""
    nb = n_fft // 2 + 1
    nbp = (nb + 31) // 32 * 32
    # the mel kernel folds the DFT about n_fft / 2 (csrc/small_ops.cuh): that needs w[n] == w[n_fft - n], which the
    # reference's periodic Hann window satisfies (torchaudio MelSpectrogram default, condition.py:69-90)
    wd = window.double()
    if n_fft % 8 or (wd[1:] - wd[1:].flip(0)).abs().max() > 1e-6 * wd.abs().max():
        raise NotImplementedError("mel front end: the window must be symmetric about n_fft / 2 and n_fft a multiple of 8")
    n = torch.arange(n_fft, dtype=torch.float64)[:, None]
    k = torch.arange(nb, dtype=torch.float64)[None, :]
    ang = 2.0 * math.pi * ((n * k) % n_fft) / n_fft
    basis = torch.zeros(n_fft, 2 * nbp, dtype=torch.float64)
    w = window.double()[:, None]
    basis[:, :nb] = torch.cos(ang) * w
    basis[:, nbp : nbp + nb] = torch.sin(ang) * w
    return basis.float(), nb, nbp


def w4_dft(basis, nb, nbp, n_fft, hop):
    """The framed, windowed real DFT of the mel front end (condition.py:92-108) as a multi-tap GEMM on the padded signal
    stored in rows of hop / 2 samples, phase-split by 2 (row m of phase s = samples [(2m + s) half, (2m + s + 1) half)):
    frame f starts (n_fft - hop) / 2 samples in front of f hop, i.e. at half-hop 2f - (n_fft / hop - 1), and spans
    2 n_fft / hop half-hops; half-hop 2(f + q) + s is source s at row shift q.  Columns: cos bins [0, nb) then, from
    column nbq (bins padded to 64), sin bins.  -> (W4 [2 nbq, 2, n_shift, c_pad], smin, c_pad, nbq); rows / channels of
    the padding are zero.  Run by the split-bf16 tensor-core GEMM of csrc/hp_ops.cuh (1e-5, far below the bf16
    rounding of the mel planes)."""
    half = hop // 2
    assert hop % 2 == 0 and half % 8 == 0 and n_fft % half == 0 and ((n_fft - hop) // 2) % half == 0
    c_pad = (half + 15) // 16 * 16
    nbq = (nb + 63) // 64 * 64
    j0 = -(((n_fft - hop) // 2) // half)
    taps = []
    for j in range(n_fft // half):
        hh = j0 + j
        s = hh % 2
        taps.append((j, s, (hh - s) // 2))
    smin, smax = min(q for _, _, q in taps), max(q for _, _, q in taps)
    w4 = torch.zeros(2 * nbq, 2, smax - smin + 1, c_pad)
    for j, s, q in taps:
        rows = basis[j * half : (j + 1) * half]          # [half, 2 nbp]
        w4[:nb, s, q - smin, :half] = rows[:, :nb].t()
        w4[nbq : nbq + nb, s, q - smin, :half] = rows[:, nbp : nbp + nb].t()
    return w4, smin, c_pad, nbq


class PackedModel:
    def __init__(self, hp=False):
        self.blob = Blob()
        self.layers = {}
        self.meta = {}
        # hp: also keep every dense layer's fp32 image for the fp32 parity tier (precision="fp32": rel-L2 <= 1e-4,
        # csrc/hp_ops.cuh): `w_hp` = f32 [S][Q][C_in][N_total] of the same multi-tap GEMM, `whh_hp` = f32
        # [dir][k < H][gate 3][u < H] of the GRU recurrence.  Doubles the blob, so only on request.
        self.hp = hp

    def add_gemm(self, name, w4, smin, c_out, up=1, bias=None, slope=None, resid=False):
        """resid: the op that uses this layer adds a residual / condition tensor; for C_out <= 128 it goes
        through identity MMAs whose images take N_TILE^2 * 2 bytes of shared memory (csrc/gemm_tc.cuh)."""
        n, s, q, ci = w4.shape
        n_tile = pick_ntile_up(c_out, up) if up > 1 else pick_ntile(c_out)
        kc = pick_kc(ci, n, n_tile, s, q, reserve=n_tile * n_tile if (resid and c_out <= 128) else 0)
        bias_n = bias.float().repeat(n // c_out) if bias is not None else None  # column n -> channel n % C_out
        self.layers[name] = dict(
            w_w=self.blob.add(pack_w4(w4, kc, n_tile // 2, bias_n)),   # per (N tile, half): images, then the bias image
            bias_w=self.blob.add_f32(bias) if bias is not None else -1,
            n_src=s, smin=smin, smax=smin + q - 1, KC=kc, N_TILE=n_tile, N_total=n,
            C_out=c_out, C_in=ci, up=up, slope=slope,
        )
        if self.hp:
            self.layers[name]["w_hp"] = self.blob.add_f32(w4.permute(1, 2, 3, 0).contiguous())
            # tensor-core variant of the fp32 tier (csrc/hp_ops.cuh gemm_hp_tc_kernel): split-bf16 images
            tc_kc = 32 if ci % 32 == 0 else 16
            self.layers[name].update(w_tc=self.blob.add(pack_w4_split(w4, tc_kc, n_tile)), tc_KC=tc_kc, tc_NT=n_tile)

    def add_conv_block(self, sd, prefix, name, rate=None, direction="none", s2d=False):
        """ConvBlock parameters (blocks.py:240-329).  s2d: also pack the three convs for phase-split-by-2 tensors
        (layers `<name>.convN.s2d`; the program builder picks them for the score network's finest level)."""
        for cn in ("conv1", "conv2", "conv3"):
            p = f"{prefix}.{cn}"
            w = eff_weight(sd, p + ".conv")
            w4, smin = w4_plain(w)
            resid = cn == "conv3" or (cn == "conv1" and name.startswith("s.dec"))
            self.add_gemm(f"{name}.{cn}", w4, smin, w4.shape[0], bias=sd[p + ".conv.bias"],
                          slope=float(sd[p + ".prelu.weight"]), resid=resid)
            if s2d:
                w4s, smin_s = w4_same_s2d(w)
                self.add_gemm(f"{name}.{cn}.s2d", w4s, smin_s, w.shape[0], up=2, bias=sd[p + ".conv.bias"],
                              slope=float(sd[p + ".prelu.weight"]), resid=resid)
        if direction in ("down", "up"):
            p = f"{prefix}.rate_change_conv"
            w = eff_weight(sd, p + ".conv")
            lp = sd.get(p + ".low_pass_filter.weights")
            lp = lp.float() if lp is not None else None
            bias = sd[p + ".bias"] if p + ".bias" in sd else sd.get(p + ".conv.bias")
            if direction == "down":
                w4, smin = w4_down(w, lp)
                self.add_gemm(f"{name}.rc", w4, smin, w.shape[0], bias=bias, slope=float(sd[p + ".prelu.weight"]))
            else:
                w4, smin = w4_up(w, lp)
                self.add_gemm(f"{name}.rc", w4, smin, w.shape[1], up=rate, bias=bias,
                              slope=float(sd[p + ".prelu.weight"]), resid=name.startswith("s.dec"))

    def add_gru(self, sd, prefix, name, layer):
        sufs = ("", "_reverse")
        w_ih = torch.cat([sd[f"{prefix}.weight_ih_l{layer}{s}"].float() for s in sufs])  # [6H, C]
        b_ih = torch.cat([sd[f"{prefix}.bias_ih_l{layer}{s}"].float() for s in sufs])
        self.add_gemm(f"{name}.ih", w_ih[:, None, None, :].contiguous(), 0, w_ih.shape[0], bias=b_ih)
        whh = [sd[f"{prefix}.weight_hh_l{layer}{s}"].float() for s in sufs]
        self.layers[f"{name}.hh"] = dict(
            whh_w=self.blob.add(pack_gru_whh(*whh)),
            bhh_w=self.blob.add_f32(torch.cat([sd[f"{prefix}.bias_hh_l{layer}{s}"].float() for s in sufs])),
            H=whh[0].shape[1],
        )
        if self.hp:
            h = whh[0].shape[1]
            self.layers[f"{name}.hh"]["whh_hp"] = self.blob.add_f32(torch.stack([w.view(3, h, h).permute(2, 0, 1) for w in whh]).contiguous())

    def add_conv_in(self, sd, prefix, name):
        w = eff_weight(sd, prefix)  # [C, 1, 3]
        if w.shape[1] != 1 or w.shape[2] != 3:
            raise NotImplementedError(f"{prefix}: the 1 -> C edge conv kernel handles fb_kernel_size == 3, input_channels == 1 (got {tuple(w.shape)})")
        self.layers[name] = dict(w_w=self.blob.add_f32(w[:, 0, :].t().contiguous()),
                                 bias_w=self.blob.add_f32(sd[prefix + ".bias"]), C=w.shape[0])


def pack_model(sd, cfg, hp=False):
    """sd: reference-layout state dict (tensors on CPU); cfg: resolved `model` config node.
    hp: also pack the fp32 weight images of the fp32 parity tier (PackedModel.hp)."""
    sd = {k: v.detach().cpu() for k, v in sd.items()}
    pm = PackedModel(hp=hp)
    scfg, ccfg = cfg["score_model"], cfg["condition_model"]
    edm = cfg.get("edm") is not None
    sp, cp = ("_edm_model" if edm else "score_model"), "condition_model"
    rates = [int(r) for r in scfg["rate_factors"]]
    extra = bool(scfg.get("extra_conv_block", False))
    nlev = len(rates)
    n_blocks = nlev + (1 if extra else 0)
    c0 = int(scfg["n_channels"])
    chans = [c0 * 2**i for i in range(nlev + 1)]
    pm.meta = dict(rates=rates, extra=extra, chans=chans, edm=edm, n_blocks=n_blocks,
                   simple_embed=scfg.get("time_embedding") == "simple",
                   noise_cond_dim=int(scfg.get("noise_cond_dim", 512)))
    if scfg.get("encoder_gru_conv_sandwich"):
        raise NotImplementedError("encoder_gru_conv_sandwich is not used by the supported configs")
    if scfg.get("seq_model", "gru") != "gru" or ccfg.get("seq_model", "gru") != "gru":
        raise NotImplementedError("only seq_model == 'gru' is supported")

    # finest level of the score network stored phase-split by 2 (needs an even first rate factor: that level's block
    # output doubles as the phase-split input of the first down conv, and 2 C columns must fit one N tile)
    s2d0 = S2D and rates[0] == 2 and 2 * c0 <= 128 and c0 % 16 == 0
    pm.meta["s2d0"] = s2d0

    # ---------------- score network (score.py:214-298)
    pm.add_conv_in(sd, sp + ".input_conv", "s.in")
    film_w, film_b, film_rows = [], [], {}
    row = 0

    def add_film(prefix, key):
        nonlocal row
        w = eff_weight(sd, prefix)
        film_w.append(w)
        film_b.append(sd[prefix + ".bias"].float())
        film_rows[key] = row
        row += w.shape[0]

    for i in range(n_blocks):
        down = i < nlev
        pm.add_conv_block(sd, f"{sp}.encoder.ds_modules.{i}", f"s.enc{i}", rates[i] if down else None,
                          "down" if down else "none", s2d=s2d0 and i == 0)
        add_film(f"{sp}.encoder.cond_proj.{i}", f"enc{i}")
    pm.add_gru(sd, sp + ".encoder.gru", "s.gru", 0)
    up_rates = ([None] if extra else []) + rates[::-1]
    for i, r in enumerate(up_rates):
        pm.add_conv_block(sd, f"{sp}.decoder.up_modules.{i}", f"s.dec{i}", r, "up" if r else "none",
                          s2d=s2d0 and i == len(up_rates) - 1)
        add_film(f"{sp}.decoder.noise_cond_proj.{i}", f"dec{i}")
        w4, smin = w4_plain(eff_weight(sd, f"{sp}.decoder.signal_cond_proj.{i}"))
        pm.add_gemm(f"s.dec{i}.sc", w4, smin, w4.shape[0], bias=sd[f"{sp}.decoder.signal_cond_proj.{i}.bias"])
    pm.layers["s.film"] = dict(w_w=pm.blob.add_f32(torch.cat(film_w)), bias_w=pm.blob.add_f32(torch.cat(film_b)),
                               rows=film_rows, R=row)
    if pm.meta["simple_embed"]:
        emb = torch.stack([sd[sp + ".sigma_block.weight"].reshape(()), sd[sp + ".sigma_block.bias"].reshape(())])
        pm.layers["s.embed"] = dict(w_w=pm.blob.add_f32(emb), n_rff=0)
    else:
        parts = [sd[sp + ".sigma_block.freq"].float().reshape(-1)]
        for i in (1, 2, 3):
            parts += [sd[f"{sp}.sigma_block.layer{i}.prelu.weight"].float().reshape(-1),
                      sd[f"{sp}.sigma_block.layer{i}.lin.weight"].float().reshape(-1),
                      sd[f"{sp}.sigma_block.layer{i}.lin.bias"].float().reshape(-1)]
        pm.layers["s.embed"] = dict(w_w=pm.blob.add_f32(torch.cat(parts)), n_rff=int(sd[sp + ".sigma_block.freq"].numel()))
    wo = eff_weight(sd, sp + ".output_conv.conv")  # [1, C, 3]
    if wo.shape[0] != 1 or wo.shape[2] != 3:
        raise NotImplementedError(f"output_conv: the C -> 1 edge conv kernel handles fb_kernel_size == 3, output_channels == 1 (got {tuple(wo.shape)})")
    pm.layers["s.out"] = dict(w_w=pm.blob.add_f32(wo[0].t().contiguous()), bias_w=pm.blob.add_f32(sd[sp + ".output_conv.conv.bias"]),
                              slope_a=float(sd[sp + ".prelu.weight"]), slope_b=float(sd[sp + ".output_conv.prelu.weight"]))

    # ---------------- signal_decoupling_layer (universe_gan.py:119-128): weights of the AUX op (csrc/aux_ops.cuh)
    ap = "signal_decoupling_layer"
    if ap + ".conv.bias" in sd:
        wa = eff_weight(sd, ap + ".conv")   # [1, C, 3]
        Ca = wa.shape[1]
        if ap + ".prelu.act.act.alpha" in sd:
            kup = sd[ap + ".prelu.act.upsample.kernel"].float()[:, 0, :]     # [2][KU]: the two output phases
            kdn = sd[ap + ".prelu.act.downsample.kernel"].float()[0, 0, :]   # [KD]
            head, mode = sd[ap + ".prelu.act.act.alpha"].float().reshape(-1), 0
        elif ap + ".prelu.weight" in sd:
            kup, kdn = torch.zeros(2, 0), torch.zeros(0)
            head, mode = sd[ap + ".prelu.weight"].float().reshape(-1).expand(Ca).contiguous(), 1
        else:
            head = None
        if head is not None and kup.shape[0] == 2 and wa.shape[2] == 3:
            blob = torch.cat([head, kup.reshape(-1), kdn, wa[0].reshape(-1), sd[ap + ".conv.bias"].float().reshape(-1)])
            pm.layers["s.aux"] = dict(w_w=pm.blob.add_f32(blob), mode=mode, KU=int(kup.shape[1]), KD=int(kdn.numel()), C=int(Ca))

    # ---------------- conditioner (condition.py:273-377)
    pm.add_conv_in(sd, cp + ".input_conv", "c.in")
    tot_ds = math.prod(rates)
    n_fft = int(ccfg.get("n_mel_oversample", 4)) * tot_ds
    basis, nb, nbp = dft_basis(sd[cp + ".input_mel.mel_spec.spectrogram.window"], n_fft)
    fb = sd[cp + ".input_mel.mel_spec.mel_scale.fb"].float()
    pm.layers["c.mel"] = dict(basis_w=pm.blob.add_f32(basis), fb_w=pm.blob.add_f32(fb), n_fft=n_fft, hop=tot_ds,
                              n_bins=nb, nbp=nbp, n_mels=fb.shape[1])
    # the same DFT as a tensor-core GEMM (csrc/hp_ops.cuh gemm_hp_tc_kernel; program.py MEL_TC): always with its fp32 /
    # split-bf16 images, whatever the tier of the rest of the model
    if tot_ds % 16 == 0:
        w4d, smin_d, c_pad, nbq = w4_dft(basis, nb, nbp, n_fft, tot_ds)
        was_hp, pm.hp = pm.hp, True
        pm.add_gemm("c.mel.dft", w4d, smin_d, w4d.shape[0])
        pm.hp = was_hp
        # bins [lo, hi) that feed each mel band (triangular filters: a band sees a few dozen bins, not all n_fft / 2 + 1)
        nzr = []
        for m in range(fb.shape[1]):
            nz = torch.nonzero(fb[:, m]).reshape(-1)
            nzr.append((int(nz.min()), int(nz.max()) + 1) if nz.numel() else (0, 0))
        rng_w = pm.blob.add(np.asarray(nzr, dtype=np.int32))
        pm.layers["c.mel"].update(dft_cpad=c_pad, dft_nbq=nbq, half=tot_ds // 2, fb_rng_w=rng_w)
    w4, smin = w4_plain(eff_weight(sd, cp + ".input_mel.conv"))
    pm.add_gemm("c.mel.conv", w4, smin, w4.shape[0], bias=sd[cp + ".input_mel.conv.bias"])
    pm.add_conv_block(sd, cp + ".input_mel.conv_block", "c.melblk")
    for i in range(n_blocks):
        down = i < nlev
        pm.add_conv_block(sd, f"{cp}.encoder.ds_modules.{i}", f"c.enc{i}", rates[i] if down else None,
                          "down" if down else "none")
        if i < nlev - 1:
            p = f"{cp}.encoder.st_convs.{i}"
            w = eff_weight(sd, p + ".conv")
            w4, smin = w4_down(w)
            pm.add_gemm(f"c.st{i}", w4, smin, w.shape[0], bias=sd[p + ".conv.bias"], slope=float(sd[p + ".prelu.weight"]))
    pm.add_conv_block(sd, cp + ".encoder.conv_block1", "c.cb1")
    pm.add_gru(sd, cp + ".encoder.gru", "c.gru0", 0)
    pm.add_gru(sd, cp + ".encoder.gru", "c.gru1", 1)
    pm.add_conv_block(sd, cp + ".encoder.conv_block2", "c.cb2")
    pm.add_conv_block(sd, cp + ".decoder.input_conv_block", "c.dec_in")
    for i, r in enumerate(up_rates):
        pm.add_conv_block(sd, f"{cp}.decoder.up_modules.{i}", f"c.dec{i}", r, "up" if r else "none")
    pm.meta["gru_residual"] = bool(ccfg.get("encoder_gru_residual", False))
    return pm
