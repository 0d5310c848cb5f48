"""TEST INFRASTRUCTURE ONLY (CPU) -- never imported by the product path.

Executes a compiled op program (open_universe_b200/program.py) on the CPU with
the exact memory layout, packed weights and rounding points (bf16 stores) of
the CUDA kernels, one op at a time.  It serves two purposes:
  * on the CPU box: proves the compiler's bookkeeping (layouts, weight folding
    and packing, epilogue flags, buffer aliasing) against the oracle without a GPU;
  * on the GPU box: gives the expected arena contents after every op, so a
    kernel bug is localised to one launch.
"""
import math

import numpy as np
import torch

from open_universe_b200 import _abi
from open_universe_b200.weights import unpack_gru_whh, unpack_w4


def _dt(t):
    return torch.bfloat16 if t.dtype == _abi.DT_BF16 else torch.float32


def op_outputs(op):
    """[(label, arena offset, nbytes, torch dtype)] of everything `op` writes into the arena (per-op GPU-vs-emulator
    comparisons: tests/test_gpu_modules.py, tools/gpu_debug.py)."""
    outs = []
    t = op.type

    def tens(label, u):
        es = 2 if u.dtype == _abi.DT_BF16 else 4
        outs.append((label, u.off, u.split * u.nch * u.P * 8 * es, torch.bfloat16 if es == 2 else torch.float32))

    if t in (_abi.OP_GEMM, _abi.OP_GRU, _abi.OP_CONV_IN, _abi.OP_MEL_FIN):
        for i in range(op.n_out):
            tens(f"out{i}", op.out[i].t)
        if t == _abi.OP_GEMM and op.rm_out >= 0:
            if op.H > 0:   # GRU gate-input layout [t][dir][rank 8][b < Bp][3U + 4]
                outs.append(("rm", op.rm_out, op.M_valid * 16 * op.i1 * (3 * (op.H // 8) + 4) * 4, torch.float32))
            else:
                outs.append(("rm", op.rm_out, op.m_rows * op.rm_ld * 4, torch.float32))
    elif t == _abi.OP_CONVBLOCK:
        for i in range(op.n_out):
            tens(f"out{i}", op.out[i].t)
        if op.i2:
            tens("cond", op.out[2].t)
    elif t == _abi.OP_PREP:
        outs += [("sig", op.p0, op.B * op.Lp * 4, torch.float32), ("rms", op.p1, op.B * 4, torch.float32)]
    elif t == _abi.OP_SIGP:
        tens("out0", op.out[0].t)
    elif t in (_abi.OP_MEL, _abi.OP_MEL_POW):
        outs += [("mel", op.p1, op.B * op.F * op.C * 4, torch.float32),
                 ("energy", op.p2, op.B * ((op.F + op.T - 1) // op.T) * 4, torch.float32)]
    elif t in (_abi.OP_INIT_X, _abi.OP_CONV_OUT, _abi.OP_AUX):
        outs.append(("x", op.p0, op.B * op.Lp * 4, torch.float32))
    elif t == _abi.OP_FILM:
        outs.append(("film", op.p1, op.i0 * op.i2 * 4, torch.float32))
    return outs


class Emulator:
    def __init__(self, pm, pg):
        self.pm, self.pg = pm, pg
        self.blob = np.frombuffer(pm.blob.bytes(), dtype=np.uint8)
        self.arena = torch.zeros(pg.arena, dtype=torch.uint8)
        init = np.frombuffer(pg.init_blob, dtype=np.uint8)
        self.arena[pg.init_off : pg.init_off + len(init)] = torch.from_numpy(init.copy())
        self._wcache = {}

    # ---------------------------------------------------------------- views
    def w_f32(self, off, n):
        return torch.from_numpy(self.blob[off : off + 4 * n].copy()).view(torch.float32)

    def a_f32(self, off, n):
        return self.arena[off : off + 4 * n].view(torch.float32)

    def view(self, t):
        """[split, nch, P, 8] view of a planes tensor."""
        es = 2 if t.dtype == _abi.DT_BF16 else 4
        n = t.split * t.nch * t.P * 8
        return self.arena[t.off : t.off + n * es].view(_dt(t)).view(t.split, t.nch, t.P, 8)

    def read_tensor(self, t, B, L):
        """-> [B, C, L] fp32 logical tensor."""
        v = self.view(t).float()
        rows = (L + t.split - 1) // t.split
        out = torch.zeros(B, t.nch * 8, rows * t.split)
        for b in range(B):
            blk = v[:, :, t.G + b * t.Lh : t.G + b * t.Lh + rows, :]  # [split, nch, rows, 8]
            out[b] = blk.permute(1, 3, 2, 0).reshape(t.nch * 8, rows * t.split)
        return out[:, :, :L]

    def _scatter(self, t, b_idx, l_idx, vals):
        """vals [n, C] -> tensor t at (b, l)."""
        v = self.view(t)
        ph = l_idx % t.split
        row = t.G + b_idx * t.Lh + l_idx // t.split
        C = t.nch * 8
        v[ph, :, row, :] = vals.to(v.dtype).view(-1, C // 8, 8)

    def _gather(self, t, b_idx, l_idx):
        v = self.view(t)
        ph = l_idx % t.split
        row = t.G + b_idx * t.Lh + l_idx // t.split
        return v[ph, :, row, :].float().reshape(len(b_idx), -1)

    @staticmethod
    def _act(v, o):
        if o.act >= 1:
            v = torch.where(v >= 0, v, o.a1 * v)
        if o.act == 2:
            v = torch.where(v >= 0, v, o.a2 * v)
        return v

    def _write_outs(self, op, b_idx, l_idx, v):
        for i in range(op.n_out):
            o = op.out[i]
            self._scatter(o.t, b_idx, l_idx, self._act(v, o))

    # ---------------------------------------------------------------- ops
    def op_gemm(self, op):
        a = self.view(op.a).float()  # [n_src, nch, P, 8]
        hp = op.p5 >= 0   # fp32 tier (csrc/hp_ops.cuh gemm_hp_kernel): f32 planes, f32 weight image [S][Q][C_in][N], no rounding
        if op.a_act:
            # raw A tensor: the kernel's transform warps apply PReLU in fp32 and round to bf16 once (csrc/gemm_tc.cuh)
            a = torch.where(a >= 0, a, np.float32(op.a_slope) * a)
            if not hp:
                a = a.to(torch.bfloat16).float()
        n_shift = op.smax - op.smin + 1
        C_in = op.a.nch * 8
        has_bias = op.bias_w >= 0
        key = (op.w_w, op.N_total, op.n_src, n_shift, C_in, has_bias, hp)
        if key not in self._wcache and hp:
            w = self.w_f32(op.p5, op.n_src * n_shift * C_in * op.N_total).view(op.n_src, n_shift, C_in, op.N_total)
            bias = self.w_f32(op.bias_w, op.C_out).repeat(op.N_total // op.C_out) if has_bias else None
            self._wcache[key] = (w.permute(3, 0, 1, 2).contiguous(), bias)
        elif key not in self._wcache:
            nb = op.N_total * op.n_src * n_shift * C_in * 2 + (op.N_total * 32 if has_bias else 0)
            bits = self.blob[op.w_w : op.w_w + nb].view(np.int16)
            res = unpack_w4(bits, op.N_total, op.n_src, n_shift, C_in, op.KC, op.N_TILE // 2, has_bias)   # packed per half N tile
            self._wcache[key] = res if has_bias else (res, None)
        w4, bias_n = self._wcache[key]
        M = op.m_rows
        acc = torch.zeros(M, op.N_total)
        for s in range(op.n_src):
            for q in range(n_shift):
                r0 = op.a.G + op.smin + q
                x = a[s, :, r0 : r0 + M, :].permute(1, 0, 2).reshape(M, C_in)
                acc += x @ w4[:, s, q, :].t()
        g = torch.arange(M)
        b_idx, m = g // op.Lh_A, g % op.Lh_A
        valid = m < op.M_valid
        b_idx, m, acc = b_idx[valid], m[valid], acc[valid]
        g_rows = g[valid]
        if op.rm_out >= 0 and op.H == 0:
            rm_plain = self.a_f32(op.rm_out, op.m_rows * op.rm_ld).view(op.m_rows, op.rm_ld)   # row-major [b Lh + t][N]
        elif op.rm_out >= 0:
            # GRU gate-input layout [t][dir][rank 8][b < Bp][3U + 4] (csrc/gru.cuh)
            H, Bp = op.H, op.i1
            U, S = H // 8, 3 * (H // 8) + 4
            rm = self.a_f32(op.rm_out, op.M_valid * 16 * Bp * S).view(op.M_valid, 2, 8, Bp, S)
        # the kernel adds the bias on the tensor core from its packed hi + lo bf16 image
        bias = bias_n[: op.C_out] if bias_n is not None else None
        film = self.a_f32(op.film, 2 * op.C_out) if op.film >= 0 else None
        for ph in range(op.N_total // op.C_out):
            v = acc[:, ph * op.C_out : (ph + 1) * op.C_out].clone()
            l = op.up * m + ph
            if bias is not None:
                v += bias
            if op.add1.off >= 0:
                v += self._gather(op.add1, b_idx, l)
            if op.add2.off >= 0:
                v += self._gather(op.add2, b_idx, l)
            v *= np.float32(op.scale)
            if film is not None:
                v = film[: op.C_out] * v + film[op.C_out :]
            if op.rm_out >= 0 and op.H == 0:
                rm_plain[g_rows, ph * op.C_out : (ph + 1) * op.C_out] = v
            elif op.rm_out >= 0:
                assert op.up == 1 and op.C_out == 6 * H
                vv = v.view(-1, 2, 3, 8, U).permute(0, 1, 3, 2, 4).reshape(-1, 2, 8, 3 * U)  # [row, dir, rank, gate*U + u]
                rm[m, :, :, b_idx, : 3 * U] = vv
            self._write_outs(op, b_idx, l, v)

    def op_prep(self, op, io):
        B, T, Lp = op.B, op.T, op.Lp
        mix = io["mix"].float()[op.i1 : op.i1 + B]   # i1 = first utterance of this lane
        x = torch.zeros(B, Lp)
        x[:, op.i0 : op.i0 + T] = mix
        rms = mix.square().mean(dim=1).sqrt()
        mean = x.mean(dim=1, keepdim=True)
        x = x - mean
        gain = op.f0 / x.std(dim=1, keepdim=True).clamp(min=1e-5)
        self.a_f32(op.p0, B * Lp).copy_((x * gain).reshape(-1))
        self.a_f32(op.p1, B).copy_(rms)

    def op_mel(self, op):
        B, Lp, F_, n_mels = op.B, op.Lp, op.F, op.C
        n_fft, hop, nb, padl, nbp = op.i0, op.i1, op.i2, op.i3, op.H
        sig = self.a_f32(op.p0, B * Lp).view(B, Lp)
        xp = torch.zeros(B, (F_ - 1) * hop + n_fft)
        n = min(Lp, xp.shape[1] - padl)
        xp[:, padl : padl + n] = sig[:, :n]
        frames = xp.unfold(-1, n_fft, hop)  # [B, F, n_fft]
        basis = self.w_f32(op.p3, n_fft * 2 * nbp).view(n_fft, 2 * nbp)
        fb = self.w_f32(op.p4, nb * n_mels).view(nb, n_mels)
        ri = frames @ basis
        power = ri[..., :nb] ** 2 + ri[..., nbp : nbp + nb] ** 2
        mel = power @ fb  # [B, F, n_mels]
        self.a_f32(op.p1, B * F_ * n_mels).copy_(mel.reshape(-1))
        ft = op.T   # frames per partial energy sum (= per CTA of the kernel)
        nft = (F_ + ft - 1) // ft
        e = torch.zeros(B, nft * ft)
        e[:, :F_] = (mel**2).sum(-1)
        self.a_f32(op.p2, B * nft).copy_(e.view(B, nft, ft).sum(-1).reshape(-1))

    def op_sigp(self, op):
        """padded signal [B][Lp] -> f32 planes in rows of C = hop / 2 samples, phase-split by 2 (csrc sig_planes_kernel)."""
        B, Lp, F_, half = op.B, op.Lp, op.F, op.C
        sig = self.a_f32(op.p0, B * Lp).view(B, Lp)
        t = op.out[0].t
        vals = torch.zeros(B * 2 * F_, t.nch * 8)
        vals[:, :half] = sig[:, : 2 * F_ * half].reshape(B * 2 * F_, half)
        b_idx = torch.arange(B).repeat_interleave(2 * F_)
        l_idx = torch.arange(2 * F_).repeat(B)
        self._scatter(t, b_idx, l_idx, vals)

    def op_mel_pow(self, op):
        """DFT rows [b Lh + f][cos | sin] -> power spectrum -> mel filterbank (csrc mel_pow_kernel); outputs as op_mel."""
        B, F_, n_mels, Lh = op.B, op.F, op.C, op.Lh_A
        nb, nbq, ld = op.i0, op.i1, op.i2
        d = self.a_f32(op.p0, B * Lh * ld).view(B, Lh, ld)[:, :F_]
        power = d[..., :nb] ** 2 + d[..., nbq : nbq + nb] ** 2
        fb = self.w_f32(op.p4, nb * n_mels).view(nb, n_mels)
        mel = power @ fb
        self.a_f32(op.p1, B * F_ * n_mels).copy_(mel.reshape(-1))
        ft = op.T
        nft = (F_ + ft - 1) // ft
        e = torch.zeros(B, nft * ft)
        e[:, :F_] = (mel**2).sum(-1)
        self.a_f32(op.p2, B * nft).copy_(e.view(B, nft, ft).sum(-1).reshape(-1))

    def op_mel_fin(self, op):
        B, F_, n_mels, nft = op.B, op.F, op.C, op.i0
        mel = self.a_f32(op.p0, B * F_ * n_mels).view(B, F_, n_mels)
        energy = self.a_f32(op.p1, B * nft).view(B, nft).sum(-1)
        norm = (energy / F_).sqrt().clamp(min=1e-5)
        v = mel / norm[:, None, None]
        b_idx = torch.arange(B).repeat_interleave(F_)
        l_idx = torch.arange(F_).repeat(B)
        self._write_outs(op, b_idx, l_idx, v.reshape(B * F_, n_mels))

    def op_conv_in(self, op):
        B, Lp, C = op.B, op.Lp, op.C
        x = self.a_f32(op.p0, B * Lp).view(B, Lp) * np.float32(op.f0)
        w = self.w_f32(op.p3, 3 * C).view(3, C)
        bias = self.w_f32(op.p4, C)
        xp = torch.nn.functional.pad(x, (1, 1))
        v = xp[:, :-2, None] * w[0] + xp[:, 1:-1, None] * w[1] + xp[:, 2:, None] * w[2] + bias  # [B, Lp, C]
        b_idx = torch.arange(B).repeat_interleave(Lp)
        l_idx = torch.arange(Lp).repeat(B)
        self._write_outs(op, b_idx, l_idx, v.reshape(B * Lp, C))

    def op_init_x(self, op, io):
        z = io["noise"][op.i0].float().reshape(-1, op.Lp)[op.i1 : op.i1 + op.B]
        self.a_f32(op.p0, op.B * op.Lp).copy_((z * np.float32(op.f0)).reshape(-1))

    def op_aux(self, op, io):
        """aux_to_wav (universe_gan.py:147-151) + x = sig [+ sigma_n z_n] (universe.py:317-331), in the packed
        weights' terms: alpha | up kernel [2][KU] | down kernel [KD] | conv [C][3] | bias."""
        from oracle import universe_oracle as uo

        B, Lp, C, KU, KD = op.B, op.Lp, op.C, op.H, op.F
        y = self.read_tensor(op.a, B, Lp)  # [B, C, Lp]
        w = self.w_f32(op.p3, C + 2 * KU + KD + 3 * C + 1)
        head, kup, kdn = w[:C], w[C : C + 2 * KU].view(2, 1, KU), w[C + 2 * KU : C + 2 * KU + KD].view(1, 1, KD)
        cw, bias = w[C + 2 * KU + KD : C + 2 * KU + KD + 3 * C].view(1, C, 3), w[-1:]
        if op.i3 == 0:
            y = uo.resample(y, kup, 1, 2)
            a = torch.exp(head)[None, :, None]
            y = y + (1.0 / (a + 1e-9)) * torch.sin(y * a) ** 2
            y = uo.resample(y, kdn, 2, 1)
        else:
            y = torch.where(y >= 0, y, head[None, :, None] * y)
        sig = torch.nn.functional.conv1d(y, cw, bias, padding="same")[:, 0, :]
        if op.f0 != 0.0:
            z = io["noise"][op.i0].float().reshape(-1, Lp)[op.i1 : op.i1 + B]
            sig = sig + np.float32(op.f0) * z
        self.a_f32(op.p0, B * Lp).copy_(sig.reshape(-1))

    def op_conv_out(self, op, io):
        B, Lp, C = op.B, op.Lp, op.C
        a = self.read_tensor(op.a, B, Lp)  # [B, C, Lp]
        w = self.w_f32(op.p3, 3 * C).view(3, C)
        bias = self.w_f32(op.p4, 1)
        ap = torch.nn.functional.pad(a, (1, 1))
        net = bias + sum((ap[:, :, k : k + Lp] * w[k][None, :, None]).sum(1) for k in range(3))  # [B, Lp]
        ca, cb, cc = [np.float32(v) for v in io["coef"][op.i0]]
        x = self.a_f32(op.p0, B * Lp).view(B, Lp)
        xn = ca * x + cb * net
        if cc != 0:
            xn = xn + cc * io["noise"][op.i0 + 1].float().reshape(-1, Lp)[op.i1 : op.i1 + B]
        x.copy_(xn)
        if io.get("dbg") is not None:
            B_total = op.i2 if op.i2 > 0 else B
            io["dbg"].setdefault(op.i0, torch.zeros(B_total, Lp))[op.i1 : op.i1 + B] = net

    def op_post(self, op, io):
        B, T, Lp = op.B, op.T, op.Lp
        x = self.a_f32(op.p0, B * Lp).view(B, Lp)[:, op.i0 : op.i0 + T].clone()
        if io.get("keep_rms"):
            rms = self.a_f32(op.p1, B)
            x_rms = x.square().mean(dim=1).sqrt().clamp(min=1e-5)
            x = x * (rms / x_rms)[:, None]
        scale = x.abs().max(dim=1, keepdim=True).values
        B_total = op.i2 if op.i2 > 0 else B
        if io.get("out") is None or io["out"].shape[0] != B_total:
            io["out"] = torch.zeros(B_total, T)
        io["out"][op.i1 : op.i1 + B] = torch.where(scale > 1.0, x / scale, x)

    def op_film(self, op):
        n_steps, kind, R, D, n_rff = op.i0, op.i1, op.i2, op.C, op.H
        ls = self.a_f32(op.p0, n_steps)
        if kind == 0:
            wb = self.w_f32(op.p3, 2)
            t = torch.arange(D // 2)
            f = 0.5 * torch.sigmoid(wb[0] * ls[:, None] + wb[1])
            p = 2.0 * math.pi * f * t
            g = torch.cat([torch.sin(p), torch.cos(p)], dim=-1)
        else:
            sizes = [(2 * n_rff, 4 * n_rff), (4 * n_rff, 8 * n_rff), (8 * n_rff, D)]
            tot = n_rff + sum(1 + i * o + o for i, o in sizes)
            buf = self.w_f32(op.p3, tot)
            freq, pos = buf[:n_rff], n_rff
            p = 2.0 * math.pi * freq[None, :] * ls[:, None]
            g = torch.cat([torch.sin(p), torch.cos(p)], dim=-1)
            for i, o in sizes:
                a = buf[pos]
                W = buf[pos + 1 : pos + 1 + i * o].view(o, i)
                bb = buf[pos + 1 + i * o : pos + 1 + i * o + o]
                pos += 1 + i * o + o
                g = g @ W.t() + bb
                g = torch.where(g >= 0, g, a * g)
        W = self.w_f32(op.p4, R * D).view(R, D)
        b = self.w_f32(op.p5, R)
        self.a_f32(op.p1, n_steps * R).copy_((g @ W.t() + b).reshape(-1))

    def op_gru(self, op, round_h=True):
        B, T, H, Lh = op.B, op.T, op.H, op.Lh_A
        Bp, U = op.i1, H // 8
        S = 3 * U + 4
        if op.p5 >= 0:
            # fp32 tier (csrc/hp_ops.cuh gru_hp_kernel): gate inputs row-major [b Lh + t][6H], f32 weights [dir][k][gate][u]
            round_h = False
            gi_bt = self.a_f32(op.p0, B * Lh * 6 * H).view(B, Lh, 6 * H)[:, :T]
            whh = self.w_f32(op.p5, 2 * H * 3 * H).view(2, H, 3, H).permute(0, 2, 3, 1).reshape(2, 3 * H, H)
        else:
            gi_blk = self.a_f32(op.p0, T * 16 * Bp * S).view(T, 2, 8, Bp, S)[:, :, :, :B, : 3 * U]  # [t, dir, rank, b, gate*U+u]
            # -> [b, t, dir*3H + gate*H + rank*U + u]
            gi_bt = gi_blk.reshape(T, 2, 8, B, 3, U).permute(3, 0, 1, 4, 2, 5).reshape(B, T, 6 * H)
            bits = self.blob[op.p3 : op.p3 + 2 * 3 * H * H * 2].view(np.int16)
            whh = unpack_gru_whh(bits.copy(), H)
        bhh = self.w_f32(op.p4, 6 * H).view(2, 3 * H)
        out = torch.zeros(B, T, 2 * H)
        for d in range(2):
            h = torch.zeros(B, H)
            g_d = gi_bt[:, :, d * 3 * H : (d + 1) * 3 * H]
            for t in (range(T) if d == 0 else range(T - 1, -1, -1)):
                hb = h.to(torch.bfloat16).float() if round_h else h
                gh = hb @ whh[d].t() + bhh[d]
                git = g_d[:, t]
                r = torch.sigmoid(git[:, :H] + gh[:, :H])
                z = torch.sigmoid(git[:, H : 2 * H] + gh[:, H : 2 * H])
                n = torch.tanh(git[:, 2 * H :] + r * gh[:, 2 * H :])
                h = (1 - z) * n + z * h
                out[:, t, d * H : (d + 1) * H] = h
        b_idx = torch.arange(B).repeat_interleave(T)
        l_idx = torch.arange(T).repeat(B)
        v = out.reshape(B * T, 2 * H)
        if op.add1.off >= 0:
            v = v + self._gather(op.add1, b_idx, l_idx)
        v = v * np.float32(op.scale)
        self._write_outs(op, b_idx, l_idx, v)

    # ---------------------------------------------------------------- driver
    def run_op(self, i, io):
        op = self.pg.ops[i]
        t = op.type
        if t == _abi.OP_GEMM:
            self.op_gemm(op)
        elif t == _abi.OP_PREP:
            self.op_prep(op, io)
        elif t == _abi.OP_MEL:
            self.op_mel(op)
        elif t == _abi.OP_MEL_FIN:
            self.op_mel_fin(op)
        elif t == _abi.OP_SIGP:
            self.op_sigp(op)
        elif t == _abi.OP_MEL_POW:
            self.op_mel_pow(op)
        elif t == _abi.OP_CONV_IN:
            self.op_conv_in(op)
        elif t == _abi.OP_INIT_X:
            self.op_init_x(op, io)
        elif t == _abi.OP_CONV_OUT:
            self.op_conv_out(op, io)
        elif t == _abi.OP_POST:
            self.op_post(op, io)
        elif t == _abi.OP_FILM:
            self.op_film(op)
        elif t == _abi.OP_GRU:
            self.op_gru(op)
        elif t == _abi.OP_AUX:
            self.op_aux(op, io)
        elif t == _abi.OP_CONVBLOCK:
            for part in op.parts:   # fused on the GPU; here the three convs run one after the other
                self.op_gemm(part)
        else:
            raise ValueError(t)

    def run(self, mix, noises, coefs, keep_rms=False, dbg=False, first=0, last=None):
        io = dict(mix=mix, noise=noises, coef=coefs, keep_rms=keep_rms, dbg={} if dbg else None)
        last = len(self.pg.ops) if last is None else last
        for i in range(first, last):
            self.run_op(i, io)
        return io
