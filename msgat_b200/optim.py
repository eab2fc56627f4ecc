"""Adam of the reference's trainer (``src/engine.py:106``: ``optim.Adam(model.parameters(), lr=1e-3,
weight_decay=5e-4)``) as ONE kernel launch over flat buffers (``csrc/loss.cu:k_adam``).

``FlatAdam`` re-homes every trainable parameter in one contiguous fp32 buffer (each ``p.data`` becomes a view
of it, names and shapes unchanged, so ``state_dict`` / ``load_state_dict`` still work) next to the flat gradient
buffer of ``distributed.FlatGradients``.  The step counter lives on the device, so a captured CUDA graph replays
the right bias correction; the 1 / world_size of the gradient average is folded into the gradient read.
"""
from __future__ import annotations

import torch

from . import _lib
from .distributed import FlatGradients


class FlatAdam:
    def __init__(self, grads: FlatGradients, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, grad_scale: float = 1.0):
        self.grads = grads
        self.lr, self.betas, self.eps, self.weight_decay, self.grad_scale = lr, betas, eps, weight_decay, grad_scale
        flat_g = grads.flat
        self.flat_p = torch.zeros_like(flat_g)   # same (256-byte aligned, zero padded) layout as the gradients
        for p, o in zip(grads.params, grads.offsets):
            n = p.numel()
            self.flat_p[o:o + n].copy_(p.data.reshape(-1))
            p.data = self.flat_p[o:o + n].view_as(p)
        self.m = torch.zeros_like(flat_g)
        self.v = torch.zeros_like(flat_g)
        self.step_count = torch.zeros(1, dtype=torch.int32, device=flat_g.device)
        self._scratch = torch.zeros(4, dtype=torch.int32, device=flat_g.device)

    def zero_grad(self, set_to_none: bool = False):
        self.grads.zero()

    def step(self):
        g = self.grads.flat
        _lib.check(_lib.get().msgat_adam_step(_lib.ptr(self.flat_p), _lib.ptr(g), _lib.ptr(self.m), _lib.ptr(self.v),
                                              g.numel(), self.lr, self.betas[0], self.betas[1], self.eps,
                                              self.weight_decay, self.grad_scale, _lib.ptr(self.step_count),
                                              _lib.ptr(self._scratch), _lib.stream_of(g)), "msgat_adam_step")

    def state_dict(self):
        return {"m": self.m, "v": self.v, "step": self.step_count, "lr": self.lr, "betas": self.betas, "eps": self.eps,
                "weight_decay": self.weight_decay}

    def load_state_dict(self, sd):
        self.m.copy_(sd["m"]); self.v.copy_(sd["v"]); self.step_count.copy_(sd["step"])
        self.lr, self.betas, self.eps, self.weight_decay = sd["lr"], tuple(sd["betas"]), sd["eps"], sd["weight_decay"]
