"""Op-program builder: turns a packed model + a shape (B, T, n_steps) into the flat
list of uv_op records (one per kernel launch) that the native executor captures
into a CUDA graph.  The topology mirrors the reference forward passes:

  Universe.enhance            networks/universe/universe.py:231-375
  ConditionerNetwork.forward  networks/universe/condition.py:346-377 (+189-220, 264-270)
  ScoreNetwork.forward        networks/universe/score.py:278-298 (+105-128, 197-211)
  ConvBlock.forward           networks/universe/blocks.py:331-416

Every tensor consumed by a tensor-core GEMM is stored already PReLU-activated
with the consumer's slope (the MMA reads shared memory directly, so the
activation is applied by the producer's epilogue); raw copies are kept only
where a residual / skip / condition needs them.
"""
import copy
import math
import os

import numpy as np
import torch

from . import _abi
from ._abi import DT_BF16, DT_F32, UvTens, new_op

G = 8  # guard rows (csrc/uv_common.cuh kGuard)
SQRT1_2 = 1.0 / math.sqrt(2.0)


def _ru(x, m):
    return (x + m - 1) // m * m


class Level:
    """Row geometry of one time resolution."""

    def __init__(self, B, L):
        self.L = L
        self.Lh = _ru(L + 2, 8)
        self.rows = B * self.Lh
        self.P = G + _ru(self.rows, 512) + G   # GEMM tiles are 256 or 512 rows (csrc/gemm_tc.cuh kMaxTileRows)


MEL_FT = 32   # frames per CTA of the mel kernel (csrc/small_ops.cuh kMelFT)
# The framed DFT of the mel front end as a tensor-core GEMM (split-bf16, csrc/hp_ops.cuh) between two small kernels
# (signal -> f32 planes in half-hop rows; power + mel filterbank) instead of the one CUDA-core kernel (UV_MEL_TC=0)
MEL_TC = os.environ.get("UV_MEL_TC", "1") != "0"
# A tensor that is consumed both raw (residual / skip) and through a PReLU (the next conv) is written ONCE, raw; the
# consuming GEMM activates it in shared memory (csrc/gemm_tc.cuh, a_act).  UV_RAW_A=0 restores the two-copy scheme.
RAW_A = os.environ.get("UV_RAW_A", "1") != "0"
ALTERNATE = os.environ.get("UV_ALTERNATE", "1") != "0"   # alternate the tile direction along each chain of kernels
RAW_A_MAXC = int(os.environ.get("UV_RAW_A_MAXC", "512"))   # channel limit of the scheme (measured on B200: it pays at every level, 46.2 ms per pass against 46.7 with 128 and 48.5 with 32)
# ... and of its use in front of the encoder's strided down convs (they re-read and re-activate all r phases of A for
# every N tile, so the transform warps pace the wide levels)
RAW_A_RC_MAXC = int(os.environ.get("UV_RAW_A_RC_MAXC", "128"))
# one launch per ConvBlock of the 64-column levels (csrc/convblock_pair.cuh): C = 64, or C = 32 in space-to-depth form
FUSE_CONVBLOCKS = os.environ.get("UV_FUSE_CONVBLOCKS", "1") != "0"
ROW_TABLES_ON = os.environ.get("UV_ROW_TABLES", "1") != "0"   # debug: 0 builds programs without per-utterance tables (no enhance(lengths=))
ROW_TABLES = 8   # distinct row spaces (samples, the U-Net levels, frames) a program may have
RECYCLE = os.environ.get("UV_RECYCLE", "1") != "0"   # liveness-based reuse of the activation slots of the arena (finalize_arena)
VBASE = 1 << 44   # planes tensors carry virtual offsets >= VBASE until finalize_arena() places them


def raw_ok(C):
    """One raw copy + in-kernel activation for a C-channel tensor?  Pays where the layers are HBM bound."""
    return RAW_A and C <= RAW_A_MAXC


class _PartSink:
    """Builds GEMM records exactly like Program.gemm() without emitting them (the parts of a fused op)."""

    def __init__(self, pg):
        self.pg = pg

    def gemm(self, *args, **kw):
        n = len(self.pg.ops)
        op = self.pg.gemm(*args, **kw)
        del self.pg.ops[n:], self.pg.names[n:]
        return op


class Program:
    def __init__(self, pm, B, T, n_steps, sigmas, log10_sigmas, in_scales, arena_base=0, lane=0, hp=False):
        self.pm = pm
        self.hp = hp             # fp32 parity tier: every planes tensor is f32, dense ops carry the fp32 weight images
        self.ops = []
        self.names = []
        self.arena = arena_base
        self.varena = VBASE      # virtual address space of the planes tensors (placed by finalize_arena)
        self.lane = lane
        self.rev = 0
        self.tensors = {}
        self.B, self.T, self.n_steps = B, T, n_steps
        self.sigmas, self.log10_sigmas, self.in_scales = sigmas, log10_sigmas, in_scales

    # ---- arena
    def alloc(self, nbytes, align=1024):
        off = _ru(self.arena, align)
        self.arena = off + nbytes
        return off

    def tens(self, name, lev, C, split=1, dtype=DT_BF16):
        """Allocates (or returns) the planes tensor `name` whose row space is `lev`."""
        if self.hp:
            dtype = DT_F32
        if name in self.tensors:
            t = self.tensors[name]
            assert (t.nch, t.split, t.Lh, t.P, t.dtype) == (C // 8, split, lev.Lh, lev.P, dtype), name
            return t
        assert C % 8 == 0
        es = 2 if dtype == DT_BF16 else 4
        off = _ru(self.varena, 1024)
        self.varena = off + split * (C // 8) * lev.P * 8 * es
        t = UvTens(off, dtype, C // 8, split, lev.Lh, lev.P, G)
        self.tensors[name] = t
        return t

    # ---- per-utterance tables of variable-length calls (row_tables() below fills them; csrc uv_set_row_tables)
    def rows_geo(self):
        """uv_row_geo records of this lane's utterances."""
        return self.rowtab + self.b_off * 16

    def rows_mv(self, M_valid):
        """int32 table of this lane's utterances: valid rows in a row space of M_valid rows per full-length utterance."""
        assert M_valid > 0 and self.Lp % M_valid == 0, (self.Lp, M_valid)
        ratio = self.Lp // M_valid
        if ratio not in self.row_ratios:
            self.row_ratios.append(ratio)
            assert len(self.row_ratios) <= ROW_TABLES
        k = self.row_ratios.index(ratio)
        return self.rowtab + self.B_total * 16 + (k * self.B_total + self.b_off) * 4

    def emit(self, op, name):
        op.lane = self.lane
        if getattr(self, "rowtab", None) is not None and ROW_TABLES_ON:
            if op.type in (_abi.OP_GEMM, _abi.OP_CONVBLOCK):
                op.rows = self.rows_mv(op.M_valid)
            elif op.type != _abi.OP_FILM:
                op.rows = self.rows_geo()
        if ALTERNATE and op.step != -2 and op.type in (_abi.OP_GEMM, _abi.OP_CONV_IN, _abi.OP_CONV_OUT, _abi.OP_CONVBLOCK):
            # the kernels of a chain walk their tensors in alternating directions: a consumer starts at the end its
            # producer just wrote, which is the part of the tensor the L2 still holds (csrc/gemm_tc.cuh, `reverse`)
            op.i3 = self.rev
            self.rev ^= 1
        self.ops.append(op)
        self.names.append(name if self.lane == 0 else f"L{self.lane}.{name}")
        return op

    # ---- op helpers
    def gemm(self, name, layer, a, lev_rows, M_valid, outs, step=-1, add1=None, add2=None, scale=1.0,
             film=-1, rm_out=-1, bias=True, a_slope=None):
        ly = self.pm.layers[layer]
        op = new_op(_abi.OP_GEMM, step)
        op.a = a
        assert a.nch * 8 == ly["C_in"] and a.split == ly["n_src"], (name, a.nch * 8, ly["C_in"], a.split, ly["n_src"])
        assert (a.Lh, a.P) == (lev_rows.Lh, lev_rows.P), name
        op.n_src, op.smin, op.smax, op.KC = ly["n_src"], ly["smin"], ly["smax"], ly["KC"]
        op.w_w, op.N_TILE, op.N_total = ly["w_w"], ly["N_TILE"], ly["N_total"]
        op.m_rows, op.Lh_A, op.M_valid = lev_rows.rows, lev_rows.Lh, M_valid
        op.up, op.C_out = ly["up"], ly["C_out"]
        op.bias_w = ly["bias_w"] if bias else -1
        op.film = film
        if add1 is not None:
            op.add1 = add1
        if add2 is not None:
            op.add2 = add2
        op.scale = scale
        op.n_out = len(outs)
        for i, (t, act, a1, a2) in enumerate(outs):
            op.out[i].t, op.out[i].act, op.out[i].a1, op.out[i].a2 = t, act, a1, a2
        op.rm_out = rm_out
        op.rm_ld = ly["N_total"]
        if a_slope is not None:   # `a` is a raw tensor: PReLU(a_slope) is applied inside the kernel
            op.a_act, op.a_slope = 1, a_slope
        if a.dtype == DT_F32:     # fp32 operands: f32 [S][Q][C_in][N_total] image of the same GEMM (csrc/hp_ops.cuh)
            op.p5 = ly["w_hp"]
            op.p4, op.i0, op.i2 = ly["w_tc"], ly["tc_KC"], ly["tc_NT"]   # split-bf16 images, K group, N tile (tensor-core variant)
        return self.emit(op, name)


def _op_tensors(op):
    """Every uv_tens record of an op."""
    yield op.a
    yield op.add1
    yield op.add2
    for k in range(3):
        yield op.out[k].t
    for part in getattr(op, "parts", ()):   # the GEMM records of a fused ConvBlock (run by the CPU emulator)
        yield from _op_tensors(part)


def _tens_bytes(t):
    return t.split * t.nch * t.P * 8 * (2 if t.dtype == DT_BF16 else 4)


def finalize_arena(pg, recycle=True, persistent=()):
    """Places the planes tensors of a lane program (virtual offsets >= VBASE) behind its raw buffers.

    recycle=False: one slot per tensor, in allocation order.
    recycle=True: a slot is shared by tensors whose live ranges do not overlap.  Time = op index over the front end /
    conditioner and the FIRST sampler step; the later steps name exactly the same tensors in the same order, so the
    placement of step 0 holds for all of them.  A tensor touched both before and inside the sampler loop (the hoisted
    condition projections), and every tensor listed in `persistent`, lives to the end.  Slots are only shared inside
    a class (utterance pitch, plane pitch, element size): the rows between utterances and the guard rows of every
    plane are never written and serve as the zero padding of the convolutions (csrc/uv_common.cuh), which stays true
    only if a plane always holds tensors of one row geometry.  The result is checked op by op on the CPU by the
    emulator tests (same program, same offsets).
    Sets pg.arena, pg.live {name: (first, last)} and patches every record."""
    tens = {t.off: (name, t) for name, t in pg.tensors.items() if isinstance(t, UvTens)}
    assert all(off >= VBASE for off in tens)
    n_ops = len(pg.ops)
    first_step = [i for i, op in enumerate(pg.ops) if op.step == 0]
    t_end = (first_step[-1] + 1) if first_step else n_ops
    live = {}
    in_loop, pre_loop = set(), set()
    for i, op in enumerate(pg.ops):
        if op.step > 0:
            continue   # same tensors as step 0 (asserted below)
        for u in _op_tensors(op):
            if u.off < 0:
                continue
            assert u.off in tens, ("op references an unknown tensor", pg.names[i])
            lo, hi = live.get(u.off, (i, i))
            live[u.off] = (min(lo, i), max(hi, i))
            (in_loop if op.step == 0 else pre_loop).add(u.off)
    if first_step:
        n0 = len(first_step)
        for n in range(1, pg.n_steps):
            for k, i0 in enumerate(first_step):
                a, b = pg.ops[i0], pg.ops[pg.step_first_op[n] + k]
                assert [u.off for u in _op_tensors(a)] == [u.off for u in _op_tensors(b)], (pg.names[i0], "step", n)
        assert pg.step_first_op[1] - pg.step_first_op[0] == n0 if pg.n_steps > 1 else True
    keep = {t.off for t in persistent if t is not None}
    for off in tens:
        lo, hi = live.get(off, (0, 0))
        if off in keep or (off in in_loop and off in pre_loop):
            hi = 1 << 30
        live[off] = (lo, hi)
    base = _ru(pg.arena, 1024)
    pg.planes_lo = base
    place = {}
    if not recycle:
        for off in sorted(tens):
            place[off] = base
            base = _ru(base + _tens_bytes(tens[off][1]), 1024)
    else:
        classes = {}
        for off, (_, t) in tens.items():
            classes.setdefault((t.Lh, t.P, t.dtype), []).append(off)
        for key in sorted(classes):
            _, P, dtype = key
            plane = P * 8 * (2 if dtype == DT_BF16 else 4)   # a multiple of 128 bytes (P is a multiple of 8)
            # interval allocation in units of planes: tensors in order of their first use, first fit among the
            # planes whose occupant's life has ended
            busy = []    # (first plane, n planes, last use)
            n_planes = 0
            for off in sorted(classes[key], key=lambda o: (live[o][0], -live[o][1], o)):
                lo, hi = live[off]
                need = tens[off][1].split * tens[off][1].nch
                busy = [b for b in busy if b[2] >= lo]   # an output may not share planes with an input of the same op
                busy.sort()
                pos = 0
                for b0, bn, _ in busy:
                    if pos + need <= b0:
                        break
                    pos = max(pos, b0 + bn)
                busy.append((pos, need, hi))
                n_planes = max(n_planes, pos + need)
                place[off] = base + pos * plane
            base = _ru(base + n_planes * plane, 1024)
    for op in pg.ops:
        for u in _op_tensors(op):
            if u.off >= VBASE:
                u.off = place[u.off]
    pg.live = {}
    for off, (name, t) in tens.items():
        pg.live[name] = live[off]
        t.off = place[off]
    pg.arena = base
    pg.planes_hi = base
    return pg


def row_tables(pg, lengths=None):
    """Host block of the per-utterance tables of a call (uv_set_row_tables) -> (bytes, arena offset, clear offset,
    clear bytes).  lengths[b] = valid samples of row b of the caller's [B, T] tensor (None: every row is T long).
    Each row gets the geometry the reference gives a file of that length on its own: pad = tot_ds - T_b % tot_ds
    (a full block when T_b is a multiple: universe.py:219-223), Lp_b = T_b + pad, F_b = Lp_b / tot_ds frames, and
    Lp_b / ratio valid rows in every row space of the program.  The cleared range covers the activation planes of
    all lanes (and the raw buffers of the later lanes in between, which every call rewrites)."""
    lanes = getattr(pg, "lanes", None) or [pg]
    p0 = lanes[0]
    Bt, T, td = p0.B_total, p0.T, p0.tot_ds
    if lengths is None:
        lengths = [T] * Bt
    lengths = [int(v) for v in lengths]
    if len(lengths) != Bt or min(lengths) < 1 or max(lengths) > T:
        raise ValueError(f"lengths: need {Bt} values in [1, {T}]")
    geo = np.zeros((Bt, 4), np.int32)
    for b, tb in enumerate(lengths):
        pad = td - tb % td
        geo[b] = (tb, pad // 2, tb + pad, (tb + pad) // td)
    mv = np.zeros((ROW_TABLES, Bt), np.int32)
    for k, ratio in enumerate(p0.row_ratios):
        assert (geo[:, 2] % ratio == 0).all()
        mv[k] = geo[:, 2] // ratio
    lo, hi = min(l.planes_lo for l in lanes), max(l.planes_hi for l in lanes)
    assert p0.rowtab + geo.nbytes + mv.nbytes <= lo
    return geo.tobytes() + mv.tobytes(), p0.rowtab, lo, hi - lo


def gru_geometry(H, B):
    """(NB, Bp, S): utterances per GRU cluster, batch pitch and row pitch (floats) of the gate-input
    buffer [t][dir][rank 8][b < Bp][S = 3U + 4].  Mirrors gru_smem_layout() in csrc/gru.cuh: NB is limited
    by the gate-thread mapping (U/8 * NB <= 256), the 227 KB of shared memory and, by choice, 32: the
    recurrence is latency bound and more, smaller clusters shorten every phase of a step."""
    U = H // 8

    def smem(nb):
        gh = _ru(3 * U * (nb + 1) * 4, 128)
        step = gh + nb * (3 * U + 4) * 4 + 2 * (U // 8) * nb * 144       # gh, gi, two output staging buffers
        return 2 * nb * H * 2 + max(3 * U * H * 2 + 2048, step) + 256   # W staging is overlaid by the step buffers

    nb = int(os.environ.get("UV_GRU_NB", "32"))   # utterances per cluster
    while nb > 16 and ((U // 8) * nb > 256 or smem(nb) > 227 * 1024):
        nb -= 16
    nb = min(nb, _ru(B, 16))
    return nb, _ru(B, nb), 3 * U + 4


def _act(t, slope):
    return (t, 1, slope, 1.0)


def _raw(t):
    return (t, 0, 1.0, 1.0)


def sampler_tables(cfg, n_steps):
    """sigma schedule in fp32 exactly as the reference builds it (universe.py:308-311,380-386)
    plus per-step network-input scale and embedding input (universe.py:175-209)."""
    d = cfg["diffusion"]
    s_min, s_max = float(d["sigma_min"]), float(d["sigma_max"])
    time = torch.linspace(0, 1, n_steps).flip(dims=[0])
    sigma = s_min * (s_max / s_min) ** time  # fp32 tensor
    edm = cfg.get("edm")
    if edm is not None:
        level_db = float(edm.get("data_level_db", cfg.get("normalization_kwargs", {}).get("level_db", 0.0)))
        sigma_data = 10.0 ** (level_db / 20.0)
        sigma_norm = (sigma**2 + sigma_data**2) ** 0.5
        w_in = 1.0 / sigma_norm
        emb_in = float(edm["noise"]) * sigma
    else:
        w_in = torch.ones_like(sigma)
        emb_in = sigma
    return sigma, torch.log10(emb_in), w_in


def step_coefs(cfg, n_steps, epsilon):
    """Per-step (a, b, c) with x <- a x + b net + c z  (universe.py:301-305, 334-343, 197-209).
    EDM: score = (w_skip x + w_out net - x) / s^2, so x + s^2 eta score = (1 - eta + eta w_skip) x + eta w_out net."""
    d = cfg["diffusion"]
    s_min, s_max = float(d["sigma_min"]), float(d["sigma_max"])
    sigma, _, _ = sampler_tables(cfg, n_steps)
    sig = [float(s) for s in sigma]
    delta_t = 1.0 / (n_steps - 1)
    gamma = (s_max / s_min) ** -delta_t
    eta = 1 - gamma**epsilon
    beta = math.sqrt(1 - gamma ** (2 * (epsilon - 1.0)))
    edm = cfg.get("edm")
    out = []
    for n in range(n_steps):
        s = sig[n]
        last = n == n_steps - 1
        e = 1.0 if last else eta
        c = 0.0 if last else beta * sig[n + 1]
        if edm is not None:
            level_db = float(edm.get("data_level_db", cfg.get("normalization_kwargs", {}).get("level_db", 0.0)))
            sd2 = (10.0 ** (level_db / 20.0)) ** 2
            w_skip = sd2 / (s * s + sd2)
            w_out = s * math.sqrt(sd2) / math.sqrt(s * s + sd2)
            out.append((1.0 - e + e * w_skip, e * w_out, c))
        else:
            out.append((1.0, s * s * e, c))
    return out


LANE_MIN_B = 16   # smallest batch that is split into two utterance lanes


def default_lanes(B, aux=False):
    """Two lanes whenever the batch is large enough: the lanes' kernels overlap each other's GRU recurrence
    (16 of 148 SMs busy for hundreds of dependent steps) and launch gaps.  UV_LANES=1 forces one lane."""
    if aux or os.environ.get("UV_LANES", "") == "1":
        return 1
    return 2 if B >= LANE_MIN_B else 1


PRECISIONS = ("bf16", "fp32")


def build_program(pm, cfg, B, T, n_steps, aux=False, lanes=None, recycle=None, precision="bf16", mel_tc=None):
    """Program of one enhance() call.  precision = "fp32": the parity tier of north_star (rel-L2 <= 1e-4): same
    topology, every activation stored in f32 planes, dense layers and recurrences run by the fp32 kernels of
    csrc/hp_ops.cuh from the fp32 weight images (pm packed with hp=True), no fused ConvBlocks.  lanes = 2 splits the utterances into two independent halves whose op
    lists are concatenated (lane 0 first); each lane owns its tensors, the FiLM table is shared.  The native
    plan captures lane 1 on a second stream (csrc/capi.cu capture_program).  One cross-lane edge staggers the
    lanes: lane 1 starts when lane 0 has finished the op named by UV_LANE_STAGGER (default: the conditioner
    encoder's second rate-change conv, about a third of a score step of lane 0's work), so that their
    recurrences do not coincide.  The result carries lane 0's attributes plus `lanes` (the per-lane programs)."""
    if precision not in PRECISIONS:
        raise ValueError(f"precision must be one of {PRECISIONS}")
    hp = precision == "fp32"
    if hp and not pm.hp:
        raise ValueError("precision='fp32' needs a model packed with pack_model(..., hp=True)")
    if lanes is None:
        lanes = default_lanes(B, aux)
    if recycle is None:
        recycle = RECYCLE
    if lanes == 1 or B < 2:
        return _finish_init(_build_lane(pm, cfg, B, T, n_steps, aux=aux, recycle=recycle, hp=hp, mel_tc=mel_tc))
    assert not aux, "aux-signal programs are single-lane (run_ops ranges address one op list)"
    B0 = (B + 1) // 2
    pg0 = _build_lane(pm, cfg, B0, T, n_steps, b_off=0, B_total=B, recycle=recycle, hp=hp, mel_tc=mel_tc)
    pg1 = _build_lane(pm, cfg, B - B0, T, n_steps, b_off=B0, B_total=B, arena_base=pg0.arena, lane=1, shared=pg0, recycle=recycle, hp=hp,
                      mel_tc=mel_tc)
    stagger = os.environ.get("UV_LANE_STAGGER", "c.enc3.rc")
    if stagger in pg0.names:
        pg0.ops[pg0.names.index(stagger)].sync |= 1
        first = next(i for i, op in enumerate(pg1.ops) if op.step != -2)
        pg1.ops[first].sync |= 2
    pg = copy.copy(pg0)
    pg.lanes = [pg0, pg1]
    pg.lane_ops = [len(pg0.ops), len(pg1.ops)]
    pg.ops = pg0.ops + pg1.ops
    pg.names = pg0.names + pg1.names
    pg.arena = pg1.arena
    pg.B = B
    return _finish_init(pg)


def _finish_init(pg):
    """Init blob of the plan: the row tables of a full-length batch, then (at its own offset) log10 sigma."""
    p0 = (getattr(pg, "lanes", None) or [pg])[0]
    tables = row_tables(pg)[0]
    pg.init_blob = tables + bytes(p0.log10s - p0.rowtab - len(tables)) + p0.log10s_blob
    pg.init_off = p0.rowtab
    return pg


def _build_lane(pm, cfg, B, T, n_steps, aux=False, b_off=0, B_total=None, arena_base=0, lane=0, shared=None, recycle=True, hp=False,
                mel_tc=None):
    """aux=True also evaluates the tail of the conditioner decoder's last block (y_hat, the input of aux_to_wav:
    reference condition.py:264-270,362; universe.py:313-331), which the default path never uses.
    b_off / B_total: this lane's first utterance in / the size of the caller's batch (io tensors cover all lanes)."""
    if B_total is None:
        B_total = B
    meta = pm.meta
    rates, chans, extra = meta["rates"], meta["chans"], meta["extra"]
    nlev = len(rates)
    tot_ds = math.prod(rates)
    pad = tot_ds - T % tot_ds
    Lp = T + pad
    sigma, log10s, w_in = sampler_tables(cfg, n_steps)
    pg = Program(pm, B, T, n_steps, sigma, log10s, w_in, arena_base=arena_base, lane=lane, hp=hp)
    pg.lanes = [pg]
    pg.b_off, pg.B_total = b_off, B_total
    pg.Lp, pg.pad_left, pg.tot_ds = Lp, pad // 2, tot_ds
    lv = [Level(B, Lp // math.prod(rates[:i])) for i in range(nlev + 1)]
    pg.levels = lv
    Fr = lv[nlev].L  # bottleneck frames
    Cb = chans[nlev]
    L = pm.layers

    # ---- fixed buffers
    pg.sig = pg.alloc(B * Lp * 4)        # normalised, padded mixture (f32 [B][Lp])
    pg.x = pg.alloc(B * Lp * 4)          # sampler state (f32 [B][Lp])
    pg.mix_rms = pg.alloc(B * 4)
    pg.coef = pg.alloc(n_steps * 16 + 16)  # uv_step_coef[n_steps] | keep_rms flag
    R = L["s.film"]["R"]
    if shared is None:
        # per-utterance geometry of the whole call: uv_row_geo [B_total] | int32 [ROW_TABLES][B_total] (row_tables());
        # the plan is created with the tables of a full-length batch (init blob = tables | log10 sigma, _finish_init)
        pg.rowtab = pg.alloc(B_total * 16 + ROW_TABLES * B_total * 4, align=1024)
        pg.row_ratios = []
        pg.log10s = pg.alloc(n_steps * 4)
        pg.film = pg.alloc(n_steps * R * 4)
        pg.init_off = pg.rowtab
        pg.log10s_blob = np.asarray(log10s.numpy(), dtype=np.float32).tobytes()
        pg.init_blob = b""

        # ---- plan-time setup: FiLM table (sigma_block.py:50-57,73-78; score.py:109,208)
        op = new_op(_abi.OP_FILM, -2)
        op.i0, op.i1, op.i2, op.C, op.H = n_steps, 0 if meta["simple_embed"] else 1, R, meta["noise_cond_dim"], L["s.embed"]["n_rff"]
        op.p0, op.p1, op.p3, op.p4, op.p5 = pg.log10s, pg.film, L["s.embed"]["w_w"], L["s.film"]["w_w"], L["s.film"]["bias_w"]
        pg.emit(op, "film_table")
    else:
        pg.log10s, pg.film, pg.init_off, pg.init_blob = shared.log10s, shared.film, shared.init_off, shared.init_blob
        pg.rowtab, pg.row_ratios = shared.rowtab, shared.row_ratios

    # ---- PREP (universe.py:219-223,272; utils/norm.py:47-87)
    level = 10.0 ** (float(cfg.get("normalization_kwargs", {}).get("level_db", 0.0)) / 20.0)
    mel = L["c.mel"]
    nft = (Fr + MEL_FT - 1) // MEL_FT
    pg.mel_energy = pg.alloc(B * nft * 4)
    pg.melraw = pg.alloc(B * Fr * mel["n_mels"] * 4)
    op = new_op(_abi.OP_PREP)
    op.B, op.T, op.Lp, op.i0, op.f0, op.p0, op.p1 = B, T, Lp, pg.pad_left, level, pg.sig, pg.mix_rms
    op.i1, op.i2 = b_off, B_total
    pg.emit(op, "prep")

    # =========================================================== conditioner
    lb = lv[nlev]
    # mel front-end (condition.py:92-114)
    use_tc = (MEL_TC if mel_tc is None else mel_tc) and "c.mel.dft" in L
    if use_tc:
        # signal -> f32 planes in rows of hop / 2 samples (phase-split by 2) -> framed DFT as ONE multi-tap GEMM on the
        # tensor cores (weights.w4_dft) -> power spectrum + mel filterbank
        dl = L["c.mel.dft"]
        sigp = pg.tens("c.sigp", lb, mel["dft_cpad"], split=2, dtype=DT_F32)
        op = new_op(_abi.OP_SIGP)
        op.B, op.Lp, op.F, op.C, op.p0 = B, Lp, Fr, mel["half"], pg.sig
        op.n_out = 1
        op.out[0].t = sigp
        pg.emit(op, "c.sigp")
        pg.dft = pg.alloc(lb.rows * dl["N_total"] * 4)   # f32 row-major [b Lh + f][cos bins | sin bins]
        pg.gemm("c.mel.dft", "c.mel.dft", sigp, lb, lb.L, [], rm_out=pg.dft, bias=False)
        op = new_op(_abi.OP_MEL_POW)
        op.B, op.F, op.C, op.T, op.Lh_A = B, Fr, mel["n_mels"], MEL_FT, lb.Lh
        op.i0, op.i1, op.i2 = mel["n_bins"], mel["dft_nbq"], dl["N_total"]
        op.p0, op.p1, op.p2, op.p4, op.p5 = pg.dft, pg.melraw, pg.mel_energy, mel["fb_w"], mel["fb_rng_w"]
        pg.emit(op, "c.mel.pow")
    else:
        op = new_op(_abi.OP_MEL)
        op.B, op.Lp, op.F, op.C, op.T = B, Lp, Fr, mel["n_mels"], MEL_FT
        op.i0, op.i1, op.i2, op.i3 = mel["n_fft"], mel["hop"], mel["n_bins"], (mel["n_fft"] - mel["hop"]) // 2
        op.H = mel["nbp"]
        op.p0, op.p1, op.p2, op.p3, op.p4 = pg.sig, pg.melraw, pg.mel_energy, mel["basis_w"], mel["fb_w"]
        pg.emit(op, "c.mel")
    t_mel = pg.tens("c.mel", lb, mel["n_mels"])
    op = new_op(_abi.OP_MEL_FIN)
    op.B, op.F, op.C, op.i0, op.p0, op.p1 = B, Fr, mel["n_mels"], nft, pg.melraw, pg.mel_energy
    op.n_out = 1
    op.out[0].t = t_mel
    pg.emit(op, "c.mel_fin")

    def conv_block(pfx, name, lev, C, h_raw, h_act, v_outs, step=-1, film=-1, cond=None, cond_out=None,
                   v_add2=None, v_scale=SQRT1_2, only_conv1=False, s2d=False):
        """conv1 -> [+cond]/sqrt2 -> [FiLM] -> conv2 -> conv3 -> (h + c)/sqrt2  (blocks.py:387-403).
        h_act = PReLU_conv1(h), h_raw = h (h_act = None: only h exists, conv1 activates it in the kernel).
        v_outs: outputs of the residual sum."""
        sfx = ""
        if s2d:
            # every tensor of this block is stored phase-split by 2 in the next level's row space; the convs are the
            # ".s2d" packings (2 sources x 3 shifts, both output phases of a row in one tile; weights.w4_same_s2d)
            assert h_act is None and cond_out is None and not only_conv1 and v_add2 is None
            sfx = ".s2d"
            c1 = pg.tens(f"{name}.c1", lev, C, split=2)
            c2 = pg.tens(f"{name}.c2", lev, C, split=2)
        else:
            # (one name per block: finalize_arena() shares the slots of tensors whose lives do not overlap)
            c1 = pg.tens(f"{name}.c1", lev, C)
            c2 = pg.tens(f"{name}.c2", lev, C)
        s2, s3 = L[name + ".conv2"]["slope"], L[name + ".conv3"]["slope"]
        outs = [] if only_conv1 else [_act(c1, s2)]
        if cond_out is not None:
            outs.append(_raw(cond_out))
        lys = [L[name + f".conv{k}" + sfx] for k in (1, 2, 3)]
        fuse = (
            FUSE_CONVBLOCKS and h_act is None and not only_conv1 and v_add2 is None and C * (2 if s2d else 1) == 64
            and all(ly["N_TILE"] == 64 and ly["N_total"] == 64 and ly["KC"] * ly["n_src"] == 64 and ly["smin"] == -ly["smax"] for ly in lys)
            and len(v_outs) + (1 if cond_out is not None else 0) <= 3 and (cond_out is None or len(v_outs) <= 2)
            and h_raw.dtype == DT_BF16 and (cond is None or (cond.dtype == DT_BF16 and (cond.nch, cond.split, cond.Lh, cond.P) == (h_raw.nch, h_raw.split, h_raw.Lh, h_raw.P)))
            and all(t.dtype == DT_BF16 and (t.nch == 8 or (s2d and t.nch == 4 and t.split == 2)) for t, *_ in v_outs)
            and (cond_out is None or (cond_out.dtype == DT_BF16 and cond_out.nch == 8))
            # the fused epilogues compute PReLU(v) as max(v, a v)
            and s2 <= 1.0 and s3 <= 1.0 and all((act < 1 or a1 <= 1.0) and (act < 2 or a2 <= 1.0) for _, act, a1, a2 in v_outs)
        )
        sink = _PartSink(pg) if fuse else pg
        g1 = sink.gemm(name + ".conv1", name + ".conv1" + sfx, h_act if h_act is not None else h_raw, lev, lev.L, outs, step, add1=cond,
                       scale=SQRT1_2 if cond is not None else 1.0, film=film,
                       a_slope=None if h_act is not None else L[name + ".conv1"]["slope"])
        if only_conv1:
            return
        g2 = sink.gemm(name + ".conv2", name + ".conv2" + sfx, c1, lev, lev.L, [_act(c2, s3)], step)
        g3 = sink.gemm(name + ".conv3", name + ".conv3" + sfx, c2, lev, lev.L, v_outs, step, add1=h_raw, add2=v_add2, scale=v_scale)
        if not fuse:
            return
        # one launch for the whole block (csrc/convblock_pair.cuh); the three GEMM records stay attached for the CPU
        # emulator, which runs them one after the other (same arithmetic, same bf16 intermediates c1 / c2)
        op = new_op(_abi.OP_CONVBLOCK, step)
        op.a = h_raw
        op.a_act, op.a_slope = 1, L[name + ".conv1"]["slope"]
        op.w_w, op.p0, op.p1 = (ly["w_w"] for ly in lys)
        op.i0 = sum((1 << k) for k, g in enumerate((g1, g2, g3)) if g.bias_w >= 0)
        op.i1 = sum(((ly["n_src"] | ((ly["smax"] - ly["smin"] + 1) << 4)) << (8 * k)) for k, ly in enumerate(lys))
        op.m_rows, op.Lh_A, op.M_valid = lev.rows, lev.Lh, lev.L
        if cond is not None:
            op.add1 = cond
        op.f0 = g1.scale
        op.film, op.C = film, C
        op.f1, op.f2 = s2, s3
        op.scale = v_scale
        op.n_out = len(v_outs)
        for i, (t, act, a1, a2) in enumerate(v_outs):
            op.out[i].t, op.out[i].act, op.out[i].a1, op.out[i].a2 = t, act, a1, a2
        if cond_out is not None:
            op.i2 = 1
            op.out[2].t = cond_out
        op.parts = [g1, g2, g3]
        pg.emit(op, name + ".block")

    def gru_layer(name, x_raw, out_list, step=-1, add1=None, scale=1.0):
        """in-proj GEMM -> recurrence (score.py:117 / condition.py:212)."""
        H = L[name + ".hh"]["H"]
        nb, bp, srow = gru_geometry(H, B)
        gi = pg.tensors.get("gi")
        if gi is None:
            # fp32 tier: plain row-major gate inputs [b Lh + t][6H] (csrc/hp_ops.cuh gru_hp_kernel)
            gi = pg.alloc(lb.rows * 6 * H * 4 if hp else Fr * 16 * bp * srow * 4)
            pg.tensors["gi"] = gi
        op = pg.gemm(name + ".ih", name + ".ih", x_raw, lb, lb.L, [], step, rm_out=gi)
        if not hp:
            op.H, op.i1 = H, bp           # the epilogue scatters into the GRU's [t][dir][rank][b][3U+4] layout
        op = new_op(_abi.OP_GRU, step)
        op.B, op.T, op.H, op.Lh_A = B, Fr, H, lb.Lh
        op.i0, op.i1 = nb, bp
        op.p0, op.p3, op.p4 = gi, L[name + ".hh"]["whh_w"], L[name + ".hh"]["bhh_w"]
        if hp:
            op.p5 = L[name + ".hh"]["whh_hp"]   # f32 [dir][k][gate][u]: selects the fp32 recurrence kernel
        if add1 is not None:
            op.add1 = add1
        op.scale = scale
        op.n_out = len(out_list)
        for i, (t, act, a1, a2) in enumerate(out_list):
            op.out[i].t, op.out[i].act, op.out[i].a1, op.out[i].a2 = t, act, a1, a2
        pg.emit(op, name + ".hh")

    # mel conv + ConvBlock (condition.py:112-113)
    mb_raw, mb_act = pg.tens("c.melblk.h", lb, Cb), pg.tens("c.melblk.ha", lb, Cb)
    pg.gemm("c.mel.conv", "c.mel.conv", t_mel, lb, lb.L, [_raw(mb_raw), _act(mb_act, L["c.melblk.conv1"]["slope"])])
    x_mel = pg.tens("c.x_mel", lb, Cb, dtype=DT_F32)
    conv_block("c.", "c.melblk", lb, Cb, mb_raw, mb_act, [_raw(x_mel)])

    # input conv (condition.py:360) and encoder (condition.py:189-206)
    raw0 = raw_ok(chans[0])
    # finest level of the score network phase-split by 2 ("space to depth", weights.w4_same_s2d); UV_S2D=0 disables
    s2d0 = bool(meta.get("s2d0")) and raw0 and os.environ.get("UV_S2D", "1") != "0"
    pg.s2d0 = s2d0
    h_raw, h_act = pg.tens("c.h0", lv[0], chans[0]), None if raw0 else pg.tens("c.h0a", lv[0], chans[0])
    op = new_op(_abi.OP_CONV_IN)
    op.B, op.Lp, op.C, op.p0, op.p3, op.p4, op.f0 = B, Lp, chans[0], pg.sig, L["c.in"]["w_w"], L["c.in"]["bias_w"], 1.0
    op.n_out = 1 if raw0 else 2
    op.out[0].t = h_raw
    if not raw0:
        op.out[1].t, op.out[1].act, op.out[1].a1, op.out[1].a2 = h_act, 1, L["c.enc0.conv1"]["slope"], 1.0
    pg.emit(op, "c.in")
    st_inputs = []
    for i in range(nlev):
        C = chans[i]
        vds = pg.tens(f"c.enc{i}.vds", lv[i + 1], C, split=rates[i])
        outs = [_act(vds, L[f"c.enc{i}.rc"]["slope"])]
        if i < nlev - 1:
            S = math.prod(rates[i:])
            vst = pg.tens(f"c.enc{i}.vst", lb, C, split=S)
            outs.append(_act(vst, L[f"c.st{i}"]["slope"]))
            st_inputs.append((i, vst))
        conv_block("c.", f"c.enc{i}", lv[i], C, h_raw, h_act, outs)
        nxt = f"c.enc{i+1}.conv1" if (i + 1 < nlev or extra) else "c.cb1.conv1"
        h_raw = pg.tens(f"c.h{i+1}", lv[i + 1], chans[i + 1])
        if raw_ok(chans[i + 1]) and (i + 1 < nlev or extra):
            h_act = None
            pg.gemm(f"c.enc{i}.rc", f"c.enc{i}.rc", vds, lv[i + 1], lv[i + 1].L, [_raw(h_raw)])
        else:
            h_act = pg.tens(f"c.h{i+1}a", lv[i + 1], chans[i + 1])
            pg.gemm(f"c.enc{i}.rc", f"c.enc{i}.rc", vds, lv[i + 1], lv[i + 1].L, [_raw(h_raw), _act(h_act, L[nxt]["slope"])])
    # extra 512-channel block (condition.py:159-166): x = its v_out
    if extra:
        x_last = pg.tens("c.x_last", lb, Cb, dtype=DT_F32)
        conv_block("c.", f"c.enc{nlev}", lb, Cb, h_raw, h_act, [_raw(x_last)])
    else:
        x_last = h_raw
    # out = (x_mel + sum_i st_conv_i(res_i) + x) / sqrt(n)   (condition.py:197-206): the st GEMMs
    # chain through add1, the last one also adds x and applies the 1/sqrt(n) scale.
    assert st_inputs, "at least two rate factors are required"
    acc = x_mel
    s_raw, s_act = pg.tens("c.sum", lb, Cb), pg.tens("c.suma", lb, Cb)
    n_sum = len(st_inputs) + 2
    for k, (i, vst) in enumerate(st_inputs):
        if k < len(st_inputs) - 1:
            nacc = pg.tens(f"c.stacc{k % 2}", lb, Cb, dtype=DT_F32)
            pg.gemm(f"c.st{i}", f"c.st{i}", vst, lb, lb.L, [_raw(nacc)], add1=acc)
            acc = nacc
        else:
            pg.gemm(f"c.st{i}", f"c.st{i}", vst, lb, lb.L, [_raw(s_raw), _act(s_act, L["c.cb1.conv1"]["slope"])],
                    add1=acc, add2=x_last, scale=1.0 / math.sqrt(n_sum))
    # conv_block1 -> 2-layer bi-GRU (+residual) -> conv_block2   (condition.py:208-216)
    g_in = pg.tens("c.gru_in", lb, Cb)
    conv_block("c.", "c.cb1", lb, Cb, s_raw, s_act, [_raw(g_in)])
    g_mid = pg.tens("c.gru_mid", lb, Cb)
    gru_layer("c.gru0", g_in, [_raw(g_mid)])
    b2_raw, b2_act = pg.tens("c.cb2.h", lb, Cb), pg.tens("c.cb2.ha", lb, Cb)
    if meta["gru_residual"]:
        gru_layer("c.gru1", g_mid, [_raw(b2_raw), _act(b2_act, L["c.cb2.conv1"]["slope"])], add1=g_in, scale=SQRT1_2)
    else:
        gru_layer("c.gru1", g_mid, [_raw(b2_raw), _act(b2_act, L["c.cb2.conv1"]["slope"])])
    # conv_block2 -> decoder.input_conv_block (condition.py:216, 266)
    d_raw, d_act = pg.tens("c.dec_in.h", lb, Cb), pg.tens("c.dec_in.ha", lb, Cb)
    conv_block("c.", "c.cb2", lb, Cb, b2_raw, b2_act, [_raw(d_raw), _act(d_act, L["c.dec_in.conv1"]["slope"])])
    # decoder (condition.py:264-270): conditions = conv1 outputs of every up module
    up_rates = ([None] if extra else []) + rates[::-1]
    n_dec = len(up_rates)
    first = "c.dec0.conv1" if extra else "c.dec0.rc"
    y_raw, y_act = pg.tens("c.dec.y4", lb, Cb), pg.tens("c.dec.y4a", lb, Cb)
    conv_block("c.", "c.dec_in", lb, Cb, d_raw, d_act, [_raw(y_raw), _act(y_act, L[first]["slope"])])
    conds = []
    lev_i = nlev
    for i, r in enumerate(up_rates):
        name = f"c.dec{i}"
        last = i == n_dec - 1
        if r is not None:
            lev_i -= 1
            lev, C = lv[lev_i], chans[lev_i]
            if raw_ok(C) and not (last and not aux):
                u_raw, u_act = pg.tens(f"{name}.u", lev, C), None
                outs = [_raw(u_raw)]
            else:
                u_raw, u_act = pg.tens(f"{name}.u", lev, C), pg.tens(f"{name}.ua", lev, C)
                outs = [_act(u_act, L[name + ".conv1"]["slope"])] + ([] if (last and not aux) else [_raw(u_raw)])
            pg.gemm(name + ".rc", name + ".rc", y_act, lv[lev_i + 1], lv[lev_i + 1].L, outs)
            y_raw, y_act = u_raw, u_act
        lev, C = lv[lev_i], chans[lev_i]
        cond = pg.tens(f"cond{i}", lev, C)
        conds.append(cond)
        if last and aux:
            pg.y_hat = pg.tens("c.y_hat", lev, C, dtype=DT_F32)
            conv_block("c.", name, lev, C, y_raw, y_act, [_raw(pg.y_hat)], cond_out=cond)
        elif last:
            # y_hat is unused by enhance() (universe.py:314-316): only conv1 (the condition) is needed
            conv_block("c.", name, lev, C, y_raw, y_act, [], cond_out=cond, only_conv1=True)
        else:
            nxt = f"c.dec{i+1}.rc"
            v_act = pg.tens(f"{name}.va", lev, C)
            conv_block("c.", name, lev, C, y_raw, y_act, [_act(v_act, L[nxt]["slope"])], cond_out=cond)
            y_raw, y_act = None, v_act
    # hoisted step-invariant 1x1 condition projections (score.py:168,192,209)
    scs = []
    lev_i = nlev
    for i, r in enumerate(up_rates):
        if r is not None:
            lev_i -= 1
        lev, C = lv[lev_i], chans[lev_i]
        sc = pg.tens(f"sc{i}", lv[1], C, split=2) if (s2d0 and lev_i == 0) else pg.tens(f"sc{i}", lev, C)
        pg.gemm(f"s.dec{i}.sc", f"s.dec{i}.sc", conds[i], lev, lev.L, [_raw(sc)])
        scs.append(sc)

    # =========================================================== sampler / score network
    op = new_op(_abi.OP_INIT_X)
    op.B, op.Lp, op.p0, op.f0, op.i0 = B, Lp, pg.x, float(sigma[0]), 0
    op.i1, op.i2 = b_off, B_total
    pg.idx_init_x = len(pg.ops)   # ops [0, idx_init_x) = front end + conditioner
    pg.emit(op, "init_x")
    pg.step_first_op = []
    frows = L["s.film"]["rows"]
    n_blocks = meta["n_blocks"]
    for n in range(n_steps):
        fbase = pg.film + n * R * 4
        if s2d0:
            h_raw, h_act = pg.tens("s.h0", lv[1], chans[0], split=2), None
        else:
            h_raw, h_act = pg.tens("s.h0", lv[0], chans[0]), None if raw0 else pg.tens("s.h0a", lv[0], chans[0])
        pg.step_first_op.append(len(pg.ops))
        op = new_op(_abi.OP_CONV_IN, n)
        op.B, op.Lp, op.C, op.p0, op.p3, op.p4 = B, Lp, chans[0], pg.x, L["s.in"]["w_w"], L["s.in"]["bias_w"]
        op.f0 = float(w_in[n])
        op.n_out = 1 if raw0 else 2
        op.out[0].t = h_raw
        if not raw0:
            op.out[1].t, op.out[1].act, op.out[1].a1, op.out[1].a2 = h_act, 1, L["s.enc0.conv1"]["slope"], 1.0
        pg.emit(op, f"s{n}.in")
        skips = []
        for i in range(n_blocks):
            lev, C = lv[min(i, nlev)], chans[min(i, nlev)]
            if i < nlev:
                # consumed only as the residual of the decoder's up conv: stored phase-split in the
                # coarser row space so that conv's GEMM can stream it like an A operand (identity MMA)
                skip = pg.tens(f"s.skip{i}", lv[i + 1], C, split=rates[i])
            else:
                skip = pg.tens(f"s.skip{i}", lev, C)
            skips.append(skip)
            outs = [_raw(skip)]
            rc_slope = None
            if i < nlev:
                if raw_ok(C) and C <= RAW_A_RC_MAXC:   # the down conv reads the (phase-split) skip tensor itself and activates it in the kernel
                    vds, rc_slope = skip, L[f"s.enc{i}.rc"]["slope"]
                else:
                    vds = pg.tens(f"s.enc{i}.vds", lv[i + 1], C, split=rates[i])
                    outs.append(_act(vds, L[f"s.enc{i}.rc"]["slope"]))
            if s2d0 and i == 0:
                conv_block("s.", f"s.enc{i}", lv[1], C, h_raw, None, outs, step=n, film=fbase + frows[f"enc{i}"] * 4, s2d=True)
            else:
                conv_block("s.", f"s.enc{i}", lev, C, h_raw, h_act, outs, step=n, film=fbase + frows[f"enc{i}"] * 4)
            if i < nlev:
                h_raw = pg.tens(f"s.h{i+1}", lv[i + 1], chans[i + 1])
                outs = [_raw(h_raw)]
                h_act = None
                if i + 1 < n_blocks and not raw_ok(chans[i + 1]):
                    h_act = pg.tens(f"s.h{i+1}a", lv[i + 1], chans[i + 1])
                    outs.append(_act(h_act, L[f"s.enc{i+1}.conv1"]["slope"]))
                pg.gemm(f"s.enc{i}.rc", f"s.enc{i}.rc", vds, lv[i + 1], lv[i + 1].L, outs, n, a_slope=rc_slope)
        gru_in = skips[-1] if extra else h_raw
        # decoder (score.py:197-211)
        lev_i = nlev
        y_act = None
        for i, r in enumerate(up_rates):
            name = f"s.dec{i}"
            res = skips[n_blocks - 1 - i]
            if i == 0:
                # GRU (score.py:117) fused with the first decoder block's (h + res)/sqrt2 when it has no up conv
                if r is None:
                    u_raw, u_act = pg.tens("s.dec0.u", lb, Cb), None if raw_ok(Cb) else pg.tens("s.dec0.ua", lb, Cb)
                    gru_layer("s.gru", gru_in, [_raw(u_raw)] + ([] if raw_ok(Cb) else [_act(u_act, L[name + ".conv1"]["slope"])]), step=n,
                              add1=res, scale=SQRT1_2)
                else:
                    y_act = pg.tens("s.gru_out", lb, Cb)
                    gru_layer("s.gru", gru_in, [_act(y_act, L[name + ".rc"]["slope"])], step=n)
            if r is not None:
                lev_i -= 1
                lev, C = lv[lev_i], chans[lev_i]
                if s2d0 and lev_i == 0:
                    u_raw, u_act = pg.tens(f"{name}.u", lv[1], C, split=2), None   # the up conv writes both phases of a row
                else:
                    u_raw, u_act = pg.tens(f"{name}.u", lev, C), None if raw_ok(C) else pg.tens(f"{name}.ua", lev, C)
                pg.gemm(name + ".rc", name + ".rc", y_act, lv[lev_i + 1], lv[lev_i + 1].L,
                        [_raw(u_raw)] + ([] if raw_ok(C) else [_act(u_act, L[name + ".conv1"]["slope"])]), n, add1=res, scale=SQRT1_2)
            lev, C = lv[lev_i], chans[lev_i]
            blk_s2d = s2d0 and lev_i == 0
            v_act = pg.tens(f"{name}.va", lv[1], C, split=2) if blk_s2d else pg.tens(f"{name}.va", lev, C)
            if i < n_dec - 1:
                vo = _act(v_act, L[f"s.dec{i+1}.rc"]["slope"])
            else:
                vo = (v_act, 2, L["s.out"]["slope_a"], L["s.out"]["slope_b"])  # score.py:290 (two PReLUs)
            conv_block("s.", name, lv[1] if blk_s2d else lev, C, u_raw, u_act, [vo], step=n, film=fbase + frows[f"dec{i}"] * 4,
                       cond=scs[i], s2d=blk_s2d)
            y_act = v_act
        op = new_op(_abi.OP_CONV_OUT, n)
        op.a = y_act
        op.B, op.Lp, op.C, op.p0, op.p3, op.p4, op.i0 = B, Lp, chans[0], pg.x, L["s.out"]["w_w"], L["s.out"]["bias_w"], n
        op.p1 = pg.coef
        op.i1, op.i2 = b_off, B_total
        pg.emit(op, f"s{n}.out")
    pg.idx_post = len(pg.ops)
    op = new_op(_abi.OP_POST)
    op.B, op.T, op.Lp, op.i0, op.p0, op.p1, op.p2 = B, T, Lp, pg.pad_left, pg.x, pg.mix_rms, pg.coef + n_steps * 16
    op.i1, op.i2 = b_off, B_total
    pg.emit(op, "post")
    # use_aux_signal / warm_start (universe.py:313-331): one AUX record per way of starting, run one at a time by
    # Universe._enhance_aux through uv_run_ops: [idx_aux] x = aux_to_wav(y_hat), [idx_aux + 1 + n] x = aux_to_wav(y_hat)
    # + sigma_n z_n (the sampler then resumes at step n).  A model without the decoupling layer keeps the host path.
    if aux and "s.aux" in L and getattr(pg, "y_hat", None) is not None and L["s.aux"]["C"] == pg.y_hat.nch * 8:
        pg.idx_aux = len(pg.ops)
        for k in range(n_steps + 1):
            op = new_op(_abi.OP_AUX, -3)
            op.a = pg.y_hat
            op.B, op.Lp, op.C, op.p0, op.p3 = B, Lp, L["s.aux"]["C"], pg.x, L["s.aux"]["w_w"]
            op.i3, op.H, op.F = L["s.aux"]["mode"], L["s.aux"]["KU"], L["s.aux"]["KD"]
            op.i0, op.f0 = (k - 1, float(sigma[k - 1])) if k > 0 else (0, 0.0)
            op.i1, op.i2 = b_off, B_total
            pg.emit(op, "aux" if k == 0 else f"aux.warm{k - 1}")
    # the aux-signal paths read y_hat after the conditioner has run (universe.py _enhance_aux)
    finalize_arena(pg, recycle=recycle, persistent=[getattr(pg, "y_hat", None)])
    pg.arena = _ru(pg.arena, 1024)
    return pg


def algorithmic_flops(cfg, Lp, n_steps):
    """Dense-layer work of one utterance in the REFERENCE's formulation (2*Co*Ci*k*L_out per conv,
    no credit for the wider anti-alias-folded kernels).  Returns a dict with the
    conditioner total, the per-step score-network total, and the share executed by the tensor-core
    GEMM kernel (everything except the GRU recurrence and the 1<->C edge convs)."""
    sc, cc = cfg["score_model"], cfg["condition_model"]
    rates = [int(r) for r in sc["rate_factors"]]
    c0 = int(sc["n_channels"])
    nlev = len(rates)
    extra = bool(sc.get("extra_conv_block", False))
    chans = [c0 * 2**i for i in range(nlev + 1)]
    lens = [Lp // math.prod(rates[:i]) for i in range(nlev + 1)]
    Cb, Fr, H = chans[nlev], lens[nlev], chans[nlev] // 2

    def block(C, L):
        return 2 * C * C * (5 + 3 + 3) * L

    def rc(C, L):  # rate change between (C @ L) and (2C @ L/r)
        return 4 * C * C * L

    gru_ih = 2 * Cb * 3 * H * 2 * Fr
    gru_hh = 2 * H * 3 * H * 2 * Fr
    edge = 2 * 3 * c0 * Lp  # 1 -> C or C -> 1, k = 3
    # score network, one evaluation
    s_gemm = 0
    sc = 0   # signal_cond_proj 1x1 convs: step invariant, the program evaluates them once (hoisted)
    for i in range(nlev):
        s_gemm += 2 * (block(chans[i], lens[i]) + rc(chans[i], lens[i]))  # encoder + decoder block at this level
        sc += 2 * chans[i] * chans[i] * lens[i]
    if extra:
        s_gemm += 2 * block(Cb, Fr)
        sc += 2 * Cb * Cb * Fr
    s_gemm += sc
    s_gemm += gru_ih
    score = s_gemm + gru_hh + 2 * edge
    # conditioner
    n_mels = int(cc.get("n_mels", 80))
    c_gemm = 2 * n_mels * Cb * 3 * Fr + block(Cb, Fr)                    # mel conv + mel ConvBlock
    for i in range(nlev):
        c_gemm += 2 * (block(chans[i], lens[i]) + rc(chans[i], lens[i]))
        if i < nlev - 1:
            c_gemm += 2 * chans[i] * Cb * lens[i]                        # st_conv
    n_bottleneck = 3 + (2 if extra else 0)                               # conv_block1/2, decoder input (+ extra enc/dec)
    c_gemm += n_bottleneck * block(Cb, Fr)
    c_gemm += 2 * gru_ih
    cond = c_gemm + 2 * gru_hh + edge
    # the hoisted 1x1 projections run once, not per step; algorithmic count keeps the reference's per-step figure
    # gemm_executed / total_executed: the same count without the n_steps - 1 repeats of the hoisted projections
    return dict(cond=cond, score=score, total=cond + n_steps * score, gemm=c_gemm + n_steps * s_gemm,
                gru_hh=(2 + n_steps) * gru_hh, gemm_executed=c_gemm + n_steps * s_gemm - (n_steps - 1) * sc,
                total_executed=cond + n_steps * score - (n_steps - 1) * sc)
