"""``TimeEmbedding`` (reference ``src/models/embeddings.py:12-39``): a gather and an add,
<0.01 % of the step; it stays stock PyTorch.  Its output, the
gate, is consumed inside the fused TPC head kernel."""
import torch
from torch import nn


class TimeEmbedding(nn.Module):
    def __init__(self, n_components: int, n_nodes: int, n_timesteps: int):
        super().__init__()
        self.n_components, self.n_nodes, self.n_timesteps = n_components, n_nodes, n_timesteps
        width = n_components * n_nodes * n_timesteps
        self.d_ebd = nn.Embedding(7, width)
        self.h_ebd = nn.Embedding(24, width)

    def forward(self, H: torch.Tensor, D: torch.Tensor) -> torch.Tensor:
        gate = self.h_ebd(H) + self.d_ebd(D)
        return gate.view(gate.shape[0], self.n_components, self.n_nodes, self.n_timesteps)

    def extra_repr(self) -> str:
        return f"n_components={self.n_components}, n_nodes={self.n_nodes}, n_timesteps={self.n_timesteps}"
