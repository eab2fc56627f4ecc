"""Drop-in ``GraphAttention`` / ``TemporalAttention`` / ``ChannelAttention``.

Same constructor arguments, ``forward`` signatures and parameter names as the reference
(``src/models/attention.py:26,32,55,62,85,91``); the arithmetic runs in the sm_100a kernels
behind ``msgat_b200.ops.meam_op``.
"""
import torch
from torch import nn

from .. import ops


def _precision(module) -> int:
    return int(getattr(module, "precision", 0))


class GraphAttention(nn.Module):
    """softmax_j(q Wg q^T) * adjacency, aggregated over nodes (attention.py:32-36)."""

    def __init__(self, n_channels: int, n_timesteps: int):
        super().__init__()
        self.n_channels, self.n_timesteps = n_channels, n_timesteps
        self.Wg = nn.Parameter(torch.empty(n_timesteps, n_timesteps))
        self.alpha = nn.Parameter(torch.empty(n_channels))

    def forward(self, signals: torch.Tensor, adjacency: torch.Tensor) -> torch.Tensor:
        C = signals.shape[1]
        cfg = ops.MeamConfig(0, 0, C, 0, precision=_precision(self))
        return ops.meam_op(signals, cfg, ops.sparse_adjacency(adjacency), {"sa_Wg": self.Wg, "sa_alpha": self.alpha})

    def extra_repr(self) -> str:
        return f"n_channels={self.n_channels}, n_timesteps={self.n_timesteps}"


class TemporalAttention(nn.Module):
    """softmax_i((q Wt1^T)(q Wt2^T)^T) over the 12 steps (attention.py:62-66)."""

    RANK = 10  # hard-coded in the reference (attention.py:58-59)

    def __init__(self, n_channels: int, n_nodes: int):
        super().__init__()
        self.n_channels, self.n_nodes = n_channels, n_nodes
        self.Wt1 = nn.Parameter(torch.empty(self.RANK, n_nodes))
        self.Wt2 = nn.Parameter(torch.empty(self.RANK, n_nodes))
        self.alpha = nn.Parameter(torch.empty(n_channels))

    def forward(self, signals: torch.Tensor) -> torch.Tensor:
        C = signals.shape[1]
        cfg = ops.MeamConfig(0, C, 0, 0)
        return ops.meam_op(signals, cfg, None, {"ta_Wt1": self.Wt1, "ta_Wt2": self.Wt2, "ta_alpha": self.alpha})

    def extra_repr(self) -> str:
        return f"n_channels={self.n_channels}, n_nodes={self.n_nodes}"


class ChannelAttention(nn.Module):
    """softmax_i(q Wc q^T) over channels (attention.py:91-94)."""

    def __init__(self, n_nodes: int, n_timesteps: int):
        super().__init__()
        self.n_nodes, self.n_timesteps = n_nodes, n_timesteps
        self.Wc = nn.Parameter(torch.empty(n_timesteps, n_timesteps))
        self.alpha = nn.Parameter(torch.empty(n_nodes))

    def forward(self, signals: torch.Tensor) -> torch.Tensor:
        C = signals.shape[1]
        cfg = ops.MeamConfig(C, 0, 0, 0)
        return ops.meam_op(signals, cfg, None, {"ca_Wc": self.Wc, "ca_alpha": self.alpha})

    def extra_repr(self) -> str:
        return f"n_nodes={self.n_nodes}, n_timesteps={self.n_timesteps}"
