"""Weight folding / packing algebra (open_universe_b200/weights.py) against torch.nn.functional."""
import math

import torch
import torch.nn.functional as F

from open_universe_b200 import weights
from open_universe_b200.networks.universe.layers import LowPass


def _apply_w4(x, w4, smin, up=1):
    """Reference evaluation of the multi-tap GEMM on x [B, C, L] with sources = phases of x."""
    n, s, q, ci = w4.shape
    B, C, L = x.shape
    M = L // s
    xs = x.view(B, C, M, s)  # phase t = l % s
    out = torch.zeros(B, n, M)
    for t in range(s):
        xp = F.pad(xs[..., t], (8, 8))
        for j in range(q):
            sh = smin + j
            out += torch.einsum("nc,bcm->bnm", w4[:, t, j, :], xp[:, :, 8 + sh : 8 + sh + M])
    if up > 1:
        co = n // up
        out = out.view(B, up, co, M).permute(0, 2, 3, 1).reshape(B, co, M * up)
    return out


def test_plain_conv():
    torch.manual_seed(0)
    w, x = torch.randn(24, 16, 5), torch.randn(2, 16, 50)
    w4, smin = weights.w4_plain(w)
    assert torch.allclose(_apply_w4(x, w4, smin), F.conv1d(x, w, padding="same"), atol=1e-4)


def test_down_conv_with_and_without_antialias():
    torch.manual_seed(1)
    for r in (2, 4, 5):
        w, x = torch.randn(32, 16, r), torch.randn(2, 16, 20 * r)
        w4, smin = weights.w4_down(w)
        assert torch.allclose(_apply_w4(x, w4, smin), F.conv1d(x, w, stride=r), atol=1e-4)
        lp = LowPass(2 * r + 1).weights
        ref = F.conv1d(F.conv1d(x, lp[None, None].expand(16, 1, -1), padding="same", groups=16), w, stride=r)
        w4, smin = weights.w4_down(w, lp)
        assert torch.allclose(_apply_w4(x, w4, smin), ref, atol=1e-4)


def test_up_conv_with_and_without_antialias():
    torch.manual_seed(2)
    for r in (2, 3, 5, 8):
        w, x = torch.randn(32, 16, r), torch.randn(2, 32, 20)
        w4, smin = weights.w4_up(w)
        assert torch.allclose(_apply_w4(x, w4, smin, up=r), F.conv_transpose1d(x, w, stride=r), atol=1e-4)
        lp = LowPass(2 * r + 1).weights
        y = F.conv_transpose1d(x, w, stride=r)
        ref = F.conv1d(y, lp[None, None].expand(16, 1, -1), padding="same", groups=16)
        w4, smin = weights.w4_up(w, lp)
        assert torch.allclose(_apply_w4(x, w4, smin, up=r), ref, atol=1e-4)


def test_pack_unpack_roundtrip():
    torch.manual_seed(3)
    w4 = torch.randn(512, 2, 3, 128).to(torch.bfloat16).float()
    bits = weights.pack_w4(w4, 64, 128)
    assert torch.equal(weights.unpack_w4(bits.reshape(-1), 512, 2, 3, 128, 64, 128), w4)
    # with the bias image (hi + lo bf16 split of the fp32 bias, one image per N tile)
    bias = torch.randn(512)
    bits = weights.pack_w4(w4, 64, 128, bias)
    assert bits.size == 512 * 2 * 3 * 128 + 512 * 16
    w_back, b_back = weights.unpack_w4(bits.reshape(-1), 512, 2, 3, 128, 64, 128, has_bias=True)
    assert torch.equal(w_back, w4)
    assert torch.allclose(b_back, bias, rtol=2.0**-15, atol=1e-30)


def test_tile_choices():
    for c in (32, 48, 64, 80, 96, 128, 192, 256, 384, 512, 768):
        kc = weights.pick_kc(c)
        assert c % kc == 0 and kc % 16 == 0
        # single-N-tile k=5 layer: the resident-weights choice must still be a legal K group
        kc = weights.pick_kc(c, min(c, 128), weights.pick_ntile(c), 1, 5)
        assert c % kc == 0 and kc % 16 == 0
    for c in (32, 48, 64, 96, 128, 192, 256, 384, 512, 768, 1536, 2304):
        n = weights.pick_ntile(c)
        assert c % n == 0 and n % 16 == 0 and n <= 256


def test_binomial_taps_closed_form():
    # SURVEY.md appendix A
    assert torch.allclose(LowPass(5).weights, torch.tensor([0.267261, 1.069045, 1.603567, 1.069045, 0.267261]), atol=1e-5)
    assert abs(LowPass(17).weights[8].item() - 2.164396) < 1e-5


def test_dft_basis_matches_rfft():
    torch.manual_seed(4)
    win = torch.hann_window(64, periodic=True)
    basis, nb, nbp = weights.dft_basis(win, 64)
    x = torch.randn(5, 64)
    ri = x @ basis
    ref = torch.fft.rfft(x * win)
    assert torch.allclose(ri[:, :nb], ref.real, atol=1e-4) and torch.allclose(ri[:, nbp : nbp + nb], -ref.imag, atol=1e-4)


def test_unsupported_edge_conv_kernel_size_is_rejected():
    """conv_in / conv_out kernels are 3-tap: any other fb_kernel_size must fail at pack time, not run wrong."""
    import pytest
    from conftest import build_model
    from open_universe_b200 import weights

    model, cfg = build_model("universepp_16k")
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    key = next(k for k in sd if k.endswith("condition_model.input_conv.weight_v") or k.endswith("condition_model.input_conv.weight"))
    sd[key] = torch.zeros(sd[key].shape[0], 1, 5)
    with pytest.raises(NotImplementedError):
        weights.pack_model(sd, model.model_config())


def test_split_bf16_images_of_the_fp32_tier():
    """pack_w4_split: item (N tile, source, K group, shift) = [hi | lo][kc/8][n_tile][8]; hi + lo reproduces the fp32
    weight to 2^-16 relative (the "bf16x3" operands of csrc/hp_ops.cuh gemm_hp_tc_kernel)."""
    import numpy as np
    torch.manual_seed(0)
    n, s, q, ci, kc, nt = 96, 2, 3, 64, 32, 48
    w4 = torch.randn(n, s, q, ci)
    bits = weights.pack_w4_split(w4, kc, nt)
    x = torch.from_numpy(np.ascontiguousarray(bits)).view(torch.bfloat16).float()
    x = x.view(n // nt, s, ci // kc, q, 2, kc // 8, nt, 8)
    rec = (x[:, :, :, :, 0] + x[:, :, :, :, 1]).permute(0, 5, 1, 3, 2, 4, 6).reshape(n, s, q, ci)   # nt, n, s, q, g, j, e
    assert (rec - w4).abs().max() <= 2.0 ** -16 * w4.abs().max()
    assert torch.equal(x[:, :, :, :, 0].permute(0, 5, 1, 3, 2, 4, 6).reshape(n, s, q, ci), w4.to(torch.bfloat16).float())


def test_dft_as_multi_tap_gemm_matches_framed_dft():
    """weights.w4_dft: frames @ windowed basis == the multi-tap GEMM on the signal in half-hop rows, phase-split by 2
    (reference condition.py:92-108: (n_fft - hop) / 2 zero padding in front, hop-spaced frames)."""
    torch.manual_seed(1)
    for hop, ov in ((160, 4), (240, 4)):
        n_fft = ov * hop
        basis, nb, nbp = weights.dft_basis(torch.hann_window(n_fft, periodic=True), n_fft)
        w4, smin, c_pad, nbq = weights.w4_dft(basis, nb, nbp, n_fft, hop)
        F = 9
        sig = torch.randn(F * hop)
        padl = (n_fft - hop) // 2
        xp = torch.zeros((F - 1) * hop + n_fft)
        xp[padl : padl + F * hop] = sig[: xp.numel() - padl]
        ref = xp.unfold(0, n_fft, hop) @ basis                      # [F, 2 nbp]
        half = hop // 2
        rows = torch.zeros(2, F + 8, c_pad)                          # phase planes with 4 zero rows on either side
        hh = sig.view(2 * F, half)
        rows[0, 4 : 4 + F, :half] = hh[0::2]
        rows[1, 4 : 4 + F, :half] = hh[1::2]
        acc = torch.zeros(F, w4.shape[0])
        for s in range(2):
            for qi in range(w4.shape[2]):
                q = smin + qi
                acc += rows[s, 4 + q : 4 + q + F] @ w4[:, s, qi, :].t()
        assert torch.allclose(acc[:, :nb], ref[:, :nb], atol=2e-4) and torch.allclose(acc[:, nbq : nbq + nb], ref[:, nbp : nbp + nb], atol=2e-4)
        assert acc[:, nb:nbq].abs().max() == 0
