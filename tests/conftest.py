import json
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# The CPU checks are thousands of small matmuls: on a container whose CPU quota is below its visible core count,
# 8 intra-op threads oversubscribe and each call takes ~100x longer (measured here: 3.6 s vs 0.03 s per 200 calls).
torch.set_num_threads(max(1, min(4, os.cpu_count() or 1)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA sm_100 (B200) device")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_cfg(model_name):
    from open_universe_b200 import config

    with open(os.path.join(GOLDEN, f"config_{model_name}.json")) as f:
        return config.to_attr(json.load(f))


def build_model(model_name, seed=7):
    """Shell model of the named architecture with the deterministic golden weights."""
    from open_universe_b200 import config, networks, testing  # noqa: F401

    cfg = load_cfg(model_name)
    model = config.instantiate(cfg.model)
    testing.seed_weights(model, seed=seed)
    return model, cfg


def golden_inputs(fix, tot_ds):
    """Rebuilds the inputs of oracle/gen_golden.py::make_inputs."""
    B, T = fix["B"], fix["T"]
    Be = B * (fix["kwargs"].get("ensemble") or 1)   # the noise covers every ensemble member
    n_steps = fix["kwargs"].get("n_steps", 8)
    seed = fix["seed_inputs"]
    g = torch.Generator().manual_seed(seed)
    mix = 0.1 * torch.randn(Be, T, generator=g) + 0.01
    mix[0] *= 30.0
    Lp = T + (tot_ds - T % tot_ds)
    gn = torch.Generator().manual_seed(seed + 1)
    noise = torch.stack([torch.randn(Be, Lp, generator=gn) for _ in range(n_steps)])
    return mix[:B], noise


GOLDEN_CASES = ["pp16k", "pp16k_opts", "orig16k", "pp24k"]


def load_fixture(case):
    return torch.load(os.path.join(GOLDEN, f"enhance_{case}.pt"), weights_only=False)
