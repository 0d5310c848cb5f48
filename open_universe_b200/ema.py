"""Container for the EMA shadow parameters stored in reference checkpoints under the key
"ema" (torch_ema.ExponentialMovingAverage state; reference universe.py:119-128,832-869).
Inference only: the shadow list is positional, in `model_parameters()` order."""
import torch


class ExponentialMovingAverage:
    def __init__(self, parameters, decay):
        self.decay = decay
        self.num_updates = 0
        self.shadow_params = [p.clone().detach() for p in parameters]
        self.collected_params = None

    def store(self, parameters):
        self.collected_params = [p.clone().detach() for p in parameters]

    def copy_to(self, parameters):
        for s, p in zip(self.shadow_params, parameters):
            p.data.copy_(s.data)

    def restore(self, parameters):
        if self.collected_params is None:
            raise RuntimeError("no parameters stored")
        for c, p in zip(self.collected_params, parameters):
            p.data.copy_(c.data)

    def to(self, *args, **kwargs):
        self.shadow_params = [p.to(*args, **kwargs) if p.is_floating_point() else p.to(*args, **{k: v for k, v in kwargs.items() if k != "dtype"}) for p in self.shadow_params]
        if self.collected_params is not None:
            self.collected_params = [p.to(*args, **kwargs) for p in self.collected_params]

    def state_dict(self):
        return dict(decay=self.decay, num_updates=self.num_updates, shadow_params=self.shadow_params,
                    collected_params=self.collected_params)

    def load_state_dict(self, sd):
        shadow = sd["shadow_params"]
        if len(shadow) != len(self.shadow_params):
            raise ValueError(f"EMA state has {len(shadow)} tensors, model has {len(self.shadow_params)} parameters")
        for new, old in zip(shadow, self.shadow_params):
            if tuple(new.shape) != tuple(old.shape):
                raise ValueError("EMA shadow parameter shape mismatch")
        self.decay = sd.get("decay", self.decay)
        self.num_updates = sd.get("num_updates", 0)
        self.shadow_params = [n.detach().clone().to(o.device, o.dtype) for n, o in zip(shadow, self.shadow_params)]
        self.collected_params = None
