"""uv_mel_spec_static: the allocation-free entry point of the C ABI (include/universe_b200.h; no counterpart in the
reference).  Called through the C ABI on two different problems in a row
-- the static scratch array is shared and keeps the previous call's contents -- and compared with the existing entry
point (uv_run_ops over the plan's mel ops, read back with uv_arena_read) within the tolerance the mel test uses
(bf16 store of the planes tensor: rel-L2 <= 5e-3).  Collected last on purpose."""
import pytest
import torch

from conftest import build_model
from gpu_util import read_planes
from open_universe_b200 import _abi, _native, program, weights
from oracle import universe_oracle as uo

pytestmark = pytest.mark.gpu


def _reference_and_signal(pm, nm, mcfg, B, T, seed):
    """Existing entry point: PREP + MEL + MEL_FIN of a plan through uv_run_ops.  Returns the plan (keeps the arena that
    holds the padded, normalised signal alive), a device view of that signal and the mel tensor [B, n_mels, F]."""
    # (mel_tc=False: the plan's CUDA-core mel op, the same kernel the static entry point launches; the default plan runs
    # the framed DFT as a tensor-core GEMM, which agrees to ~1e-5 but not bit for bit)
    pg = program.build_program(pm, mcfg, B, T, 8, lanes=1, mel_tc=False)
    plan = _native.NativePlan(nm, pg, use_graph=False)
    g = torch.Generator().manual_seed(seed)
    mix = (0.1 * torch.randn(B, T, generator=g)).cuda()
    noise = torch.zeros(8, B, pg.Lp, device="cuda")
    out = torch.zeros(B, T, device="cuda")
    coefs = program.step_coefs(mcfg, 8, 1.3)
    last = pg.names.index("c.mel_fin") + 1
    plan.run_ops(0, last, mix.data_ptr(), noise.data_ptr(), out.data_ptr(), None, coefs, False, None)
    F = pg.Lp // pm.layers["c.mel"]["hop"]
    ref = read_planes(plan, pg.tensors["c.mel"], B, F)
    sig = plan.arena_view(pg.sig, B * pg.Lp, torch.device("cuda", 0)).view(B, pg.Lp)
    return plan, pg, sig, ref


def test_static_scratch_entry_point_matches_existing_one_twice_in_a_row():
    model, cfg = build_model("universepp_16k")
    mcfg = model.model_config()
    pm = weights.pack_model({k: v.detach().float() for k, v in model.state_dict().items()}, mcfg)
    nm = _native.NativeModel(pm.blob.bytes(), 0)
    mel = pm.layers["c.mel"]
    cases = [_reference_and_signal(pm, nm, mcfg, 3, 16000, 11), _reference_and_signal(pm, nm, mcfg, 2, 9000, 12)]
    outs = [torch.full(tuple(ref.shape), float("nan"), device="cuda") for _, _, _, ref in cases]
    torch.cuda.synchronize()
    for k, ((plan, pg, sig, ref), o) in enumerate(zip(cases, outs)):
        B = sig.shape[0]
        if k == 1:
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        rc = nm.mel_spec_static(mel, sig.data_ptr(), B, pg.Lp, o.data_ptr(), None)
        assert rc == 0, _native.lib().uv_last_error()
        torch.cuda.synchronize()
        if k == 1:
            assert torch.cuda.mem_get_info()[0] == free_before   # the call allocated nothing on the device
        err = uo.rel_l2(o.cpu(), ref)
        print(f"uv_mel_spec_static call {k}: B={B} Lp={pg.Lp}, rel-L2 vs uv_run_ops path {err:.2e}")
        assert torch.isfinite(o).all() and err < 5e-3
        # same arithmetic, so after the same bf16 rounding the two entry points agree bit for bit
        assert torch.equal(o.cpu().to(torch.bfloat16).float(), ref)
    # first problem again after the second one used the shared scratch: same bits as before
    plan, pg, sig, ref = cases[0]
    again = torch.empty_like(outs[0])
    assert nm.mel_spec_static(mel, sig.data_ptr(), sig.shape[0], pg.Lp, again.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert torch.equal(again, outs[0])


def test_static_scratch_entry_point_rejects_oversize_and_bad_arguments():
    model, cfg = build_model("universepp_16k")
    mcfg = model.model_config()
    pm = weights.pack_model({k: v.detach().float() for k, v in model.state_dict().items()}, mcfg)
    nm = _native.NativeModel(pm.blob.bytes(), 0)
    mel = pm.layers["c.mel"]
    x = torch.zeros(16, device="cuda")
    Lp = 64160
    F = Lp // mel["hop"]
    per_clip = (F * mel["n_mels"] + (F + 31) // 32) * 4
    B_max = _abi.UV_STATIC_SCRATCH_BYTES // per_clip
    # one clip more than the static array holds: rejected before anything is launched (the pointers are never read)
    rc = nm.mel_spec_static(mel, x.data_ptr(), B_max + 1, Lp, x.data_ptr(), None)
    assert rc == -1 and b"static array holds" in _native.lib().uv_last_error()
    assert nm.mel_spec_static(mel, x.data_ptr(), 1, Lp + 1, x.data_ptr(), None) == -1      # not a multiple of hop
    assert nm.mel_spec_static(mel, None, 1, Lp, x.data_ptr(), None) == -1
    torch.cuda.synchronize()
