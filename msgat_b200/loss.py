"""Device-side loss and metrics of the training step.

``HuberLoss`` / ``huber_loss`` keep the reference's signatures (``src/loss.py:12-52``) and ``Metrics`` its
interface (``src/metrics.py:11-38``: ``update``, ``MAE`` / ``MAPE`` / ``RMSE``, ``todict``), but one kernel
(``csrc/loss.cu``) reads ``pred`` and ``target`` once and produces the loss, its gradient and the three
running metric sums with no host synchronisation; the reference reads four scalars back per step
(``engine.py:66``, ``metrics.py:24,30,34``).  ``Metrics`` values are fetched from the device only when asked.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import _lib


def _scratch(device) -> torch.Tensor:
    n = int(_lib.get().msgat_loss_scratch_bytes())
    return torch.zeros((n + 7) // 8, dtype=torch.float64, device=device)


class Metrics:
    """Streaming MAE / MAPE / RMSE (metrics.py:11-38) whose running sums live on the device."""

    def __init__(self, mask_value: float = 0.0, device=None):
        self.mask_value = float(mask_value)
        self.state: Optional[torch.Tensor] = None   # double[4]: sum |r|, 100 sum |r / y| (y > mask), sum r^2, n
        self._device = device
        self._scratch = None
        self._loss = None

    def _ensure(self, device):
        if self.state is None or self.state.device != device:
            self.state = torch.zeros(4, dtype=torch.float64, device=device)
            self._scratch = _scratch(device)
            self._loss = torch.empty(1, dtype=torch.float32, device=device)

    def update(self, y_pred: torch.Tensor, y_true: torch.Tensor):
        """metrics.py:20-35, without the three ``.item()`` calls."""
        pred = y_pred.detach().float().contiguous()
        true = y_true.detach().float().contiguous()
        self._ensure(pred.device)
        _lib.check(_lib.get().msgat_loss_forward(_lib.ptr(pred), _lib.ptr(true), pred.numel(), 1.0, self.mask_value,
                                                 _lib.ptr(self._loss), None, _lib.ptr(self.state),
                                                 _lib.ptr(self._scratch), _lib.stream_of(pred)), "msgat_loss_forward")

    def _sums(self):
        if self.state is None:
            return 0.0, 0.0, 0.0, 0.0
        return tuple(float(v) for v in self.state.cpu())   # the one device -> host read

    @property
    def n(self):
        return int(self._sums()[3])

    @property
    def MAE(self):
        ae, _, _, n = self._sums()
        return ae / n if n else 0.0

    @property
    def MAPE(self):
        _, ape, _, n = self._sums()
        return ape / n if n else 0.0

    @property
    def RMSE(self):
        _, _, se, n = self._sums()
        return (se / n) ** 0.5 if n else 0.0

    def todict(self):
        ae, ape, se, n = self._sums()
        if not n:
            return {"MAE": 0.0, "MAPE": 0.0, "RMSE": 0.0}
        return {"MAE": ae / n, "MAPE": ape / n, "RMSE": (se / n) ** 0.5}


class _HuberFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, output, target, delta, metrics, scratch):
        lib = _lib.get()
        pred = output.float().contiguous()
        true = target.detach().float().contiguous()
        if pred.shape != true.shape:
            raise ValueError(f"huber_loss: output {tuple(pred.shape)} and target {tuple(true.shape)} differ")
        loss = torch.empty(1, dtype=torch.float32, device=pred.device)
        dpred = torch.empty_like(pred) if ctx.needs_input_grad[0] else None
        state = None
        if metrics is not None:
            metrics._ensure(pred.device)
            state = metrics.state
        _lib.check(lib.msgat_loss_forward(_lib.ptr(pred), _lib.ptr(true), pred.numel(), float(delta),
                                          0.0 if metrics is None else metrics.mask_value, _lib.ptr(loss),
                                          _lib.ptr(dpred), _lib.ptr(state), _lib.ptr(scratch), _lib.stream_of(pred)),
                   "msgat_loss_forward")
        ctx.dpred = dpred
        return loss.view(())

    @staticmethod
    def backward(ctx, gout):
        dpred = ctx.dpred
        if dpred is None:
            return None, None, None, None, None
        g = gout.detach().float().contiguous().view(1)
        out = torch.empty_like(dpred)
        _lib.check(_lib.get().msgat_loss_backward(_lib.ptr(dpred), _lib.ptr(g), dpred.numel(), _lib.ptr(out),
                                                  _lib.stream_of(dpred)), "msgat_loss_backward")
        return out, None, None, None, None


_SCRATCH = {}


def huber_loss(output: torch.Tensor, target: torch.Tensor, delta: float = 1.0,
               metrics: Optional[Metrics] = None) -> torch.Tensor:
    """``huber_loss`` of loss.py:28-52; with ``metrics`` the same pass also does ``metrics.update(output, target)``."""
    key = (str(output.device), torch.cuda.current_stream(output.device).cuda_stream if output.is_cuda else 0)
    sc = _SCRATCH.get(key)
    if sc is None:
        sc = _SCRATCH[key] = _scratch(output.device)
    with torch.autocast(device_type=output.device.type, enabled=False):
        return _HuberFunction.apply(output, target, delta, metrics, sc)


class HuberLoss(nn.Module):
    """Drop-in for ``loss.HuberLoss`` (loss.py:12-25)."""

    def __init__(self, delta: float = 1.0):
        super().__init__()
        self.delta = delta

    def forward(self, output: torch.Tensor, target: torch.Tensor, metrics: Optional[Metrics] = None) -> torch.Tensor:
        return huber_loss(output, target, self.delta, metrics)
