"""`load_model` with the reference's contract (open_universe/inference_utils/model_loader.py:33-140):
checkpoint path -> config lookup (dir/config.yaml, ../../.hydra/config.yaml, parents[3]/.hydra/config.yaml)
-> model instantiation -> state dict (+ EMA overlay) -> eval mode.  A non-existing path is treated as a
Hugging Face model id `repo[:revision]` (files weights.ckpt + config.yaml)."""
from pathlib import Path

import torch

from .. import networks  # noqa: F401  (registers the _target_ classes)
from ..config import instantiate, load_config


def ckpt_to_config_path(ckpt_path):
    ckpt_path = Path(ckpt_path)
    candidates = [ckpt_path.parent / "config.yaml"]
    parents = ckpt_path.parents
    if len(parents) > 1:
        candidates.append(parents[1] / ".hydra/config.yaml")
    if len(parents) > 3:
        candidates.append(parents[3] / ".hydra/config.yaml")
    for c in candidates:
        if c.exists():
            return c
    raise ValueError(f"Could not find the configuration file for model {ckpt_path}.")


def load_model(ckpt_path, device=None, strict=True, return_config=False, hf_token=None):
    if not Path(ckpt_path).exists():
        try:
            from huggingface_hub import hf_hub_download

            ckpt_path = str(ckpt_path)
            repo_id, _, revision = ckpt_path.partition(":")
            revision = revision or None
            ckpt_path = hf_hub_download(repo_id=repo_id, filename="weights.ckpt", revision=revision, token=hf_token)
            config_path = hf_hub_download(repo_id=repo_id, filename="config.yaml", revision=revision, token=hf_token)
        except Exception as e:
            print(f"{ckpt_path} is not a local file and download from HF hub failed.")
            raise e
    else:
        ckpt_path = Path(ckpt_path)
        config_path = ckpt_to_config_path(ckpt_path)

    config = load_config(config_path)
    model = instantiate(config.model)
    model = model.to(device)
    data = torch.load(ckpt_path, map_location=device, weights_only=False)

    if getattr(model, "ema", None) is not None and "ema" in data:
        model.ema.load_state_dict(data["ema"])
        model.load_state_dict(data["state_dict"], strict=False)
    elif getattr(model, "ema", None) is not None:
        model.load_state_dict(data["state_dict"], strict=strict)
        model.ema.shadow_params = [p.clone().detach() for p in model.model_parameters()]
    else:
        model.load_state_dict(data["state_dict"], strict=strict)
    model.eval()
    if return_config:
        return model, config
    return model
