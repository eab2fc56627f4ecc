"""One training step of the reference's engine (``src/engine.py:52-63``: forward, loss, backward,
optimizer step) captured in a CUDA graph.  The step launches ~135 kernels of this library plus the
optimizer's; replaying them as one graph removes the launch gaps between them (B200: ~0.7 ms of a
16 ms step).  The arithmetic is the eager step's, launch for launch."""
from __future__ import annotations

from typing import Callable, Sequence

import torch

from .distributed import FlatGradients


class GraphedStep:
    """``step = GraphedStep(model, loss_fn, optimizer, grads, example_batch)``; then
    ``loss = step(X, H, D, Y)`` copies the batch into the captured input buffers, replays the graph
    and returns the (device) loss of that step.  The optimizer must be capturable
    (``msgat_b200.optim.FlatAdam``, or ``torch.optim.Adam(..., fused=True, capturable=True)``), gradients live in
    ``grads.flat``."""

    def __init__(self, model: torch.nn.Module, loss_fn: Callable, optimizer: torch.optim.Optimizer, grads: FlatGradients,
                 example: Sequence[torch.Tensor], warmup: int = 3):
        self.model, self.loss_fn, self.opt, self.grads = model, loss_fn, optimizer, grads
        self.inputs = [t.clone() for t in example]
        # warm-up on a side stream: allocator pools, the adjacency's CSR/CSC cache, optimizer state
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self.eager(*self.inputs)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self.eager(*self.inputs)

    def eager(self, X, H, D, Y):
        out = self.model(X, H, D)
        loss = self.loss_fn(out, Y)
        self.grads.zero()
        loss.backward()
        if getattr(self.opt, "grad_scale", None) is not None:
            self.grads.all_reduce_sum()      # optim.FlatAdam divides by the world size while it reads the gradients
        else:
            self.grads.all_reduce_mean()
        self.opt.step()
        return loss.detach()

    def __call__(self, X, H, D, Y):
        for dst, src in zip(self.inputs, (X, H, D, Y)):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.loss


class DeviceFeeder:
    """Double-buffered host -> device feed: ``stage(batch)`` starts the copy of the NEXT step's pinned host
    tensors on a side stream, ``take()`` makes the caller's stream wait for it and returns the device
    tensors.  The 8 MB of a PEMSD7 batch then move while the previous step computes."""

    def __init__(self, example: Sequence[torch.Tensor], device: torch.device):
        self.stream = torch.cuda.Stream(device=device)
        self.bufs = [[torch.empty_like(t, device=device) for t in example] for _ in range(2)]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.free = [torch.cuda.Event(), torch.cuda.Event()]
        self.head = 0       # next buffer to fill
        self.staged = []    # buffers filled and not yet taken

    def stage(self, host_batch: Sequence[torch.Tensor]):
        i = self.head
        self.head ^= 1
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.free[i])          # the step that read this buffer has finished
            for dst, src in zip(self.bufs[i], host_batch):
                dst.copy_(src, non_blocking=True)
            self.ready[i].record(self.stream)
        self.staged.append(i)

    def take(self):
        i = self.staged.pop(0)
        torch.cuda.current_stream().wait_event(self.ready[i])
        return i, self.bufs[i]

    def release(self, i: int):
        """Call after the step that consumed buffer ``i`` has been enqueued on the current stream."""
        self.free[i].record(torch.cuda.current_stream())
