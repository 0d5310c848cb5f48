"""Deterministic synthetic weights / inputs (no checkpoint or dataset is reachable offline).
Used by bench.py (random-init model of the named architecture) and by the golden-vector
tests (the same seeded weights are rebuilt on any box with this torch build)."""
import math

import torch


def seed_weights(model, seed=0):
    """Fills every parameter of `model` deterministically (CPU generator, state-dict order) with
    well-scaled values; PReLU slopes, weight-norm gains and the time embedding are perturbed away
    from their defaults so that a slope / gain mixed up between layers changes the output."""
    g = torch.Generator().manual_seed(seed)
    named = dict(model.named_parameters())
    with torch.no_grad():
        for name, p in named.items():
            shape = tuple(p.shape)
            if name.endswith("prelu.weight"):
                v = 0.05 + 0.4 * torch.rand(shape, generator=g)
            elif name.endswith("sigma_block.weight"):
                v = torch.full(shape, 0.7)
            elif name.endswith("sigma_block.bias"):
                v = torch.full(shape, -0.3)
            elif name.endswith("act.alpha"):
                v = torch.zeros(shape)
            elif name.endswith("weight_g"):
                continue  # set from weight_v below
            elif p.ndim >= 2:
                fan_in = p[0].numel()
                b = 1.0 / math.sqrt(fan_in)
                v = (2 * torch.rand(shape, generator=g) - 1) * b
            else:
                v = 0.02 * torch.randn(shape, generator=g)
            p.copy_(v.to(p.device, p.dtype))
        for name, p in named.items():
            if name.endswith("weight_g"):
                v = named[name[: -len("weight_g")] + "weight_v"].detach().float().cpu()
                dims = tuple(range(1, v.ndim))
                gain = v.pow(2).sum(dim=dims, keepdim=True).sqrt() * (0.8 + 0.4 * torch.rand(tuple(p.shape), generator=g))
                p.copy_(gain.to(p.device, p.dtype))
        for name, b in model.named_buffers():
            if name.endswith("sigma_block.freq"):
                b.copy_((16 * torch.randn(tuple(b.shape), generator=g)).to(b.device, b.dtype))
    # keep the EMA shadow (what eval() swaps in, reference universe.py:841-862) in sync
    if getattr(model, "ema", None) is not None:
        model.ema.shadow_params = [p.detach().clone() for p in model.model_parameters()]
        model.ema.collected_params = None
    if hasattr(model, "_invalidate"):
        model._invalidate()
    return model


def perturb_snake_alpha(model):
    """seed_weights() leaves the snake's log-alpha of the signal decoupling layer at 0 (its default); the
    aux-signal fixtures spread it so that a channel mix-up or a missing exp() shows."""
    sdl = getattr(model, "signal_decoupling_layer", None)
    if sdl is not None and hasattr(sdl.prelu, "act"):
        a = sdl.prelu.act.act.alpha
        with torch.no_grad():
            a.copy_(torch.linspace(-0.7, 0.9, a.numel()).to(a))
        if getattr(model, "ema", None) is not None:
            model.ema.shadow_params = [p.detach().clone() for p in model.model_parameters()]
        if hasattr(model, "_invalidate"):
            model._invalidate()
    return model


def synthetic_noisy_speech(B, T, seed=1234, device="cpu"):
    """Broadband 'noisy speech' stand-in: one-pole low-passed noise + white noise."""
    g = torch.Generator().manual_seed(seed)
    x = 0.1 * torch.randn(B, T, generator=g)
    y = torch.empty_like(x)
    acc = torch.zeros(B)
    # cheap one-pole via cumulative filter in blocks (exact recursion is not needed for a benchmark signal)
    k = torch.exp(-torch.arange(64, dtype=torch.float32) / 8.0)
    y = torch.nn.functional.conv1d(torch.nn.functional.pad(x[:, None], (63, 0)), k.flip(0)[None, None])[:, 0]
    y = y / y.std(dim=1, keepdim=True) * 0.1
    return (y + 0.05 * torch.randn(B, T, generator=g)).to(device)
