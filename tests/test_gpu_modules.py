"""Module-level GPU parity through the C ABI's per-op entry points (uv_run_ops + uv_arena_read / uv_arena_write):
a failing module is named instead of showing up as a drift of the whole enhance() output.
  * every launch of a 2 x 4 s program against the CPU emulator of the op program (same layouts, same rounding points),
  * the conditioner network's five conditioning tensors against the oracle's ConditionerNetwork
    (reference condition.py:346-377),
  * the mel front end against the oracle's MelAdapter.compute_mel_spec (condition.py:92-108),
  * one bidirectional GRU layer alone at 401 steps against torch.nn.GRU (reference score.py:84-90,117)."""
import math

import pytest
import torch

from conftest import build_model
from gpu_util import read_planes, write_planes
from open_universe_b200 import _abi, _native, program, weights
from oracle import universe_oracle as uo
from oracle.emulator import Emulator, op_outputs

pytestmark = pytest.mark.gpu


def _plan(name, B, T, n_steps=8, graph=False, lanes=None, recycle=None):
    model, cfg = build_model(name)
    sd = {k: v.detach().float() for k, v in model.state_dict().items()}
    mcfg = model.model_config()
    pm = weights.pack_model(sd, mcfg)
    pg = program.build_program(pm, mcfg, B, T, n_steps, lanes=lanes, recycle=recycle)
    nm = _native.NativeModel(pm.blob.bytes(), 0)
    plan = _native.NativePlan(nm, pg, use_graph=graph)
    coefs = program.step_coefs(mcfg, n_steps, float(cfg.model.diffusion.epsilon))
    return model, cfg, sd, pm, pg, plan, coefs


class _Io:
    def __init__(self, pg, B, T, n_steps, seed):
        g = torch.Generator().manual_seed(seed)
        self.mix = 0.1 * torch.randn(B, T, generator=g)
        self.noise = torch.stack([torch.randn(B, pg.Lp, generator=g) for _ in range(n_steps)])
        self.dmix, self.dnoise = self.mix.cuda(), self.noise.cuda()
        self.dout = torch.zeros(B, T, device="cuda")
        self.ddbg = torch.zeros(n_steps, B, pg.Lp, device="cuda")

    def ptrs(self):
        return self.dmix.data_ptr(), self.dnoise.data_ptr(), self.dout.data_ptr(), self.ddbg.data_ptr()


def test_every_op_matches_emulator_at_4s():
    B, T = 2, 64000
    model, cfg, sd, pm, pg, plan, coefs = _plan("universepp_16k", B, T)
    io = _Io(pg, B, T, 8, 1235)
    em = Emulator(pm, pg)
    eio = dict(mix=io.mix, noise=io.noise, coef=coefs, keep_rms=False, dbg={})
    worst, bad = {}, []
    for i, (op, name) in enumerate(zip(pg.ops, pg.names)):
        em.run_op(i, eio)
        plan.run_ops(i, i + 1, *io.ptrs(), coefs, False, None)
        for label, off, nbytes, dt in op_outputs(op):
            gt = torch.frombuffer(bytearray(plan.arena_read(off, nbytes)), dtype=dt).float()
            et = em.arena[off : off + nbytes].view(dt).float()
            err = (gt - et).norm().item() / max(et.norm().item(), 1e-12)
            kind = _abi.OP_NAMES[op.type]
            worst[kind] = max(worst.get(kind, 0.0), err)
            if not (err < 2e-2) or not torch.isfinite(gt).all():
                bad.append((i, name, label, err))
    e_out = uo.rel_l2(io.dout.cpu(), eio["out"])
    print("worst rel-L2 per op kind vs emulator:", {k: f"{v:.1e}" for k, v in worst.items()}, f"final out {e_out:.2e}")
    assert not bad, bad[:10]
    assert e_out < 1e-2


def test_conditioner_tensors_match_oracle_at_4s():
    B, T = 2, 64000
    # one slot per tensor: the conditioning tensors are read back after the whole conditioner has run (with slot
    # recycling, the default, a tensor's planes are reused once its last consumer is done)
    model, cfg, sd, pm, pg, plan, coefs = _plan("universepp_16k", B, T, recycle=False)
    io = _Io(pg, B, T, 8, 77)
    plan.run_ops(0, pg.idx_init_x, *io.ptrs(), coefs, False, None)
    # oracle: pad + normalise exactly like enhance() (universe.py:219-223), then ConditionerNetwork.forward
    tot = model.tot_ds
    pad = tot - T % tot
    x = torch.nn.functional.pad(io.mix[:, None, :], (pad // 2, pad - pad // 2))
    x = uo.normalize(x, cfg.model.normalization_kwargs["level_db"])
    with torch.no_grad():
        conds = uo.conditioner(sd, "condition_model", cfg.model.condition_model, x)
    errs = []
    for i, c in enumerate(conds):
        got = read_planes(plan, pg.tensors[f"cond{i}"], B, c.shape[-1])
        errs.append(uo.rel_l2(got, c))
    print("conditioner: rel-L2 of the conditioning tensors (coarse -> fine):", [f"{e:.1e}" for e in errs])
    assert max(errs) < 1e-2


def test_mel_front_end_matches_oracle_at_4s():
    B, T = 3, 64000
    model, cfg, sd, pm, pg, plan, coefs = _plan("universepp_16k", B, T)
    io = _Io(pg, B, T, 8, 5)
    last = pg.names.index("c.mel_fin") + 1
    plan.run_ops(0, last, *io.ptrs(), coefs, False, None)
    tot = model.tot_ds
    pad = tot - T % tot
    x = torch.nn.functional.pad(io.mix[:, None, :], (pad // 2, pad - pad // 2))
    x = uo.normalize(x, cfg.model.normalization_kwargs["level_db"])
    ref = uo.mel_spec(sd, "condition_model.input_mel", x, tot, cfg.model.condition_model.get("n_mel_oversample", 4))
    got = read_planes(plan, pg.tensors["c.mel"], B, ref.shape[-1])
    err = uo.rel_l2(got, ref)
    print(f"mel front end ({ref.shape[1]} bands x {ref.shape[2]} frames, bf16 store): rel-L2 {err:.2e}")
    assert err < 5e-3


@pytest.mark.parametrize("B", [8, 40])
def test_gru_layer_alone_matches_torch_gru_401_steps(B):
    """c.gru0 (input projection GEMM + cluster recurrence) on a random bf16 input, T = 401 dependent steps."""
    T = 64000
    model, cfg, sd, pm, pg, plan, coefs = _plan("universepp_16k", B, T, lanes=1)
    io = _Io(pg, B, T, 8, 3)
    Fr = pg.levels[-1].L
    Cb = pg.tensors["c.gru_in"].nch * 8
    g = torch.Generator().manual_seed(9)
    x = (0.5 * torch.randn(B, Cb, Fr, generator=g)).to(torch.bfloat16).float()
    write_planes(plan, pg.tensors["c.gru_in"], x)
    i_ih, i_hh = pg.names.index("c.gru0.ih"), pg.names.index("c.gru0.hh")
    assert i_hh == i_ih + 1
    plan.run_ops(i_ih, i_hh + 1, *io.ptrs(), coefs, False, None)
    got = read_planes(plan, pg.tensors["c.gru_mid"], B, Fr)
    gru = torch.nn.GRU(Cb, Cb // 2, num_layers=1, batch_first=True, bidirectional=True)
    pfx = "condition_model.encoder.gru."
    with torch.no_grad():
        for n, p in gru.named_parameters():
            p.copy_(sd[pfx + n])
        ref = gru(x.transpose(1, 2))[0].transpose(1, 2)
    err = uo.rel_l2(got, ref)
    worst_t = ((got - ref).norm(dim=1) / ref.norm(dim=1).clamp(min=1e-6)).max().item()
    print(f"GRU layer alone, B={B}, {Fr} steps: rel-L2 {err:.2e}, worst (b, t) column {worst_t:.2e}")
    assert math.isfinite(err) and err < 1e-2
