"""CPU oracle for the MS-GAT attention hot path.  TEST INFRASTRUCTURE ONLY.

This file is a functional restatement (torch CPU ops, any float dtype, autograd
for the backward) of the algorithm in the reference's ``src/models/attention.py``
and ``src/models/msgat.py``.  It is the checker, never the product: only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it.  Nothing under ``msgat_b200/`` does.

Parity status: PINNED.  ``tests/golden/*.npz`` were produced by importing the
unmodified reference modules from ``/root/reference/src`` in the build container
(``tests/golden/make_golden.py``, committed); ``tests/test_oracle.py`` checks every
function here against those fixtures (fp64 outputs, input grads and every
parameter grad).  The reference itself ships no tests or golden vectors
(SURVEY.md section 4), so the pin is "outputs of the reference itself run here".

All parameters are read from a flat ``dict`` keyed exactly like the reference's
``state_dict()`` (``tpcs.0.tgacns.1.gacn.gatt.Wg`` ...), so a reference checkpoint
can be fed in directly.

Shapes: signals ``[B, C, N, T]`` (T innermost), adjacency ``[N, N]``.
"""
from __future__ import annotations

import re
from typing import Dict, List, Sequence

import torch

# 0.0 = the reference's ReLU (msgat.py:131).  The parity tests move the switching point by the
# fp32 noise of a pre-activation (+-tau) to measure how much the reference's OWN gradient changes
# when a near-zero pre-activation changes sign: that sensitivity bounds what any fp32 path can match.
RELU_THRESHOLD = 0.0

Params = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------
# attention primitives (reference: src/models/attention.py)
# ----------------------------------------------------------------------------
def graph_attention(h: torch.Tensor, adj: torch.Tensor, Wg: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    """Spatial attention, reference ``GraphAttention.forward`` attention.py:32-36.

    q[b,n,t] = sum_c alpha[c] h[b,c,n,t]           (:33, k and q are the same tensor)
    P        = softmax_j( q Wg q^T )  over ALL N   (:34)
    A        = P * adj   (mask AFTER the softmax)  (:36)
    out[b,c,n,t] = sum_i A[b,n,i] h[b,c,i,t]       (:36)
    """
    B, C, N, T = h.shape
    q = (h * alpha.view(1, C, 1, 1)).sum(1)  # [B,N,T]
    scores = torch.matmul(torch.matmul(q, Wg), q.transpose(1, 2))  # [B,N,N]
    att = torch.softmax(scores, dim=-1) * adj
    flat = h.permute(0, 2, 1, 3).reshape(B, N, C * T)  # node-major rows
    out = torch.matmul(att, flat)  # [B,N,C*T]
    return out.reshape(B, N, C, T).permute(0, 2, 1, 3)


def temporal_attention(h: torch.Tensor, Wt1: torch.Tensor, Wt2: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    """Temporal attention, reference ``TemporalAttention.forward`` attention.py:62-66.

    q[b,t,n] = sum_c alpha[c] h[b,c,n,t]; K1 = q Wt1^T, K2 = q Wt2^T  ([B,T,10]);
    P = softmax_i(K1 K2^T) [B,T,T]; out[b,c,n,t] = sum_i P[b,t,i] h[b,c,n,i].
    """
    B, C, N, T = h.shape
    q = (h * alpha.view(1, C, 1, 1)).sum(1).transpose(1, 2)  # [B,T,N]
    k1 = torch.matmul(q, Wt1.t())
    k2 = torch.matmul(q, Wt2.t())
    att = torch.softmax(torch.matmul(k1, k2.transpose(1, 2)), dim=-1)  # [B,T,T]
    return torch.matmul(h, att.transpose(1, 2).unsqueeze(1))  # h[b,c,n,:] @ P^T


def channel_attention(h: torch.Tensor, Wc: torch.Tensor, alpha: torch.Tensor) -> torch.Tensor:
    """Channel attention, reference ``ChannelAttention.forward`` attention.py:91-94.

    q[b,c,t] = sum_n alpha[n] h[b,c,n,t]; P = softmax_i(q Wc q^T) [B,C,C];
    out[b,c,n,t] = sum_i P[b,c,i] h[b,i,n,t].
    """
    B, C, N, T = h.shape
    q = (h * alpha.view(1, 1, N, 1)).sum(2)  # [B,C,T]
    att = torch.softmax(torch.matmul(torch.matmul(q, Wc), q.transpose(1, 2)), dim=-1)  # [B,C,C]
    return torch.matmul(att, h.reshape(B, C, N * T)).reshape(B, C, N, T)


# ----------------------------------------------------------------------------
# branch networks (reference: src/models/msgat.py)
# ----------------------------------------------------------------------------
def channel_mix(x: torch.Tensor, W: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """1x1 channel mixing y[b,o,n,t] = sum_c W[o,c] x[b,c,n,t] (+ bias[o]).

    Equivalent of ``nn.Conv2d(C, O, 1)`` (msgat.py:94,116) and of
    ``x.transpose(1,-1) @ W.T`` (msgat.py:27)."""
    B, C, N, T = x.shape
    y = torch.matmul(W.reshape(W.shape[0], C), x.reshape(B, C, N * T)).reshape(B, -1, N, T)
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y


def causal_conv(x: torch.Tensor, W: torch.Tensor, bias: torch.Tensor, dilation: int) -> torch.Tensor:
    """``Conv2d(k=[1,2], dilation=[1,d], padding=[0,d])`` followed by ``Chomp(d)``
    (msgat.py:68-74, 50-51) written as the causal two-tap filter it is:
        y[:, :, :, t] = W[..., 1] x[t] + W[..., 0] x[t-d] + bias,  x[t<0] = 0.
    ``W`` has the Conv2d layout [O, C, 1, 2]."""
    T = x.shape[-1]
    now = channel_mix(x, W[:, :, 0, 1])
    past = channel_mix(x, W[:, :, 0, 0])
    shifted = torch.zeros_like(past)
    if dilation < T:
        shifted[..., dilation:] = past[..., : T - dilation]
    return now + shifted + bias.view(1, -1, 1, 1)


def layer_norm_t(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """``nn.LayerNorm([T])`` (msgat.py:115,152): normalise the last axis only,
    biased variance, eps 1e-5, affine over T."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w + b


def gacn(h: torch.Tensor, adj: torch.Tensor, p: Params, pre: str) -> torch.Tensor:
    """``GACN.forward`` msgat.py:25-28: graph attention then ``W`` (no bias)."""
    return channel_mix(graph_attention(h, adj, p[pre + "gatt.Wg"], p[pre + "gatt.alpha"]), p[pre + "W"])


def tacn_dilations(p: Params, pre: str) -> List[int]:
    """Indices of the conv layers inside ``TACN.seq`` (msgat.py:68-74): attention
    at seq.0, convs at the odd slots, Chomp modules (no params) at the even ones."""
    idx = sorted({int(m.group(1)) for k in p for m in [re.match(re.escape(pre) + r"seq\.(\d+)\.weight$", k)] if m})
    return idx


def tacn(h: torch.Tensor, p: Params, pre: str, dilations: Sequence[int]) -> torch.Tensor:
    """``TACN.forward`` msgat.py:76-77: temporal attention then the conv chain,
    no non-linearity in between."""
    y = temporal_attention(h, p[pre + "seq.0.Wt1"], p[pre + "seq.0.Wt2"], p[pre + "seq.0.alpha"])
    for i, d in enumerate(dilations):
        slot = 1 + 2 * i
        y = causal_conv(y, p[f"{pre}seq.{slot}.weight"], p[f"{pre}seq.{slot}.bias"], d)
    return y


def cacn(h: torch.Tensor, p: Params, pre: str) -> torch.Tensor:
    """``CACN.forward`` msgat.py:96-97: channel attention then 1x1 conv with bias."""
    y = channel_attention(h, p[pre + "seq.0.Wc"], p[pre + "seq.0.alpha"])
    return channel_mix(y, p[pre + "seq.1.weight"], p[pre + "seq.1.bias"])


def meam(x: torch.Tensor, adj: torch.Tensor, p: Params, pre: str, dilations: Sequence[int]) -> torch.Tensor:
    """``MEAM.forward`` msgat.py:121-131: LN over T, the three branches
    concatenated in the order [cacn, tacn, gacn], residual 1x1 conv of the
    UN-normalised input, ReLU."""
    h = layer_norm_t(x, p[pre + "ln.weight"], p[pre + "ln.bias"])
    cat = torch.cat([cacn(h, p, pre + "cacn."), tacn(h, p, pre + "tacn.", dilations), gacn(h, adj, p, pre + "gacn.")], 1)
    z = cat + channel_mix(x, p[pre + "res.weight"], p[pre + "res.bias"])
    if RELU_THRESHOLD != 0.0:  # conditioning probe of the parity tests only (tests/parity.py)
        return z * (z > RELU_THRESHOLD).to(z.dtype)
    return torch.relu(z)


def tpc(x: torch.Tensor, adj: torch.Tensor, p: Params, pre: str, dilations: Sequence[Sequence[int]]) -> torch.Tensor:
    """``TPC.forward`` msgat.py:155-160: MEAM stack, LN over T, then the
    ``Conv2d(T -> Tout, [1, C])`` head = per-node linear over (c, t)."""
    for j, d in enumerate(dilations):
        x = meam(x, adj, p, f"{pre}tgacns.{j}.", d)
    y = layer_norm_t(x, p[pre + "ln.weight"], p[pre + "ln.bias"])
    fcw = p[pre + "fc.weight"]  # [Tout, T, 1, C]
    B, C, N, T = y.shape
    out = torch.einsum("bcnt,otc->bno", y, fcw[:, :, 0, :])
    return out + p[pre + "fc.bias"].view(1, 1, -1)


def time_gate(H: torch.Tensor, D: torch.Tensor, p: Params, M: int, N: int, Tout: int) -> torch.Tensor:
    """``TimeEmbedding.forward`` embeddings.py:33-36 -> [B, M, N, Tout]."""
    g = p["te.h_ebd.weight"][H] + p["te.d_ebd.weight"][D]
    return g.view(len(H), M, N, Tout)


def msgat(X: torch.Tensor, H: torch.Tensor, D: torch.Tensor, p: Params, dilations: Sequence[Sequence[int]]) -> torch.Tensor:
    """``MSGAT.forward`` msgat.py:202-204 (use_te=True): sum over components of
    tpc_m(X[:, m], adj) * gate_m."""
    adj = p["adj"]
    M, N = X.shape[1], adj.shape[0]
    Tout = p["tpcs.0.fc.bias"].shape[0]
    G = time_gate(H, D, p, M, N, Tout)
    out = 0
    for m in range(M):
        out = out + tpc(X[:, m], adj, p, f"tpcs.{m}.", dilations) * G[:, m]
    return out


def huber(pred: torch.Tensor, target: torch.Tensor, delta: float) -> torch.Tensor:
    """``huber_loss`` loss.py:51-52."""
    r = (pred - target).abs()
    return torch.where(r <= delta, 0.5 * r * r, delta * r - 0.5 * delta * delta).mean()


def metrics_update(state: Dict[str, float], pred: torch.Tensor, true: torch.Tensor, mask_value: float = 0.0):
    """``Metrics.update`` metrics.py:20-35: running sums n, AE, APE, SE and the derived MAE / MAPE / RMSE
    (MAPE divides by ALL elements, not by the number of unmasked ones, as the reference does)."""
    state["n"] = state.get("n", 0) + true.nelement()
    state["AE"] = state.get("AE", 0.0) + float(torch.abs(pred - true).sum())
    mask = true > mask_value
    mp, mt = pred[mask], true[mask]
    state["APE"] = state.get("APE", 0.0) + 100 * float(torch.abs((mp - mt) / mt).sum())
    state["SE"] = state.get("SE", 0.0) + float(torch.square(pred - true).sum())
    state["MAE"] = state["AE"] / state["n"]
    state["MAPE"] = state["APE"] / state["n"]
    state["RMSE"] = (state["SE"] / state["n"]) ** 0.5
    return state


MODEL_DILATIONS = {  # msgat.py:220-229
    "msgat48": [[1, 2], [2, 4]],
    "msgat72": [[1, 2], [2, 4]],
    "msgat96": [[1, 1, 2, 2], [4, 4]],
}
MODEL_WIDTH = {"msgat48": 48, "msgat72": 72, "msgat96": 96}


# ----------------------------------------------------------------------------
# parameter construction with the reference's shapes and init (msgat.py:206-217)
# ----------------------------------------------------------------------------
def init_params(model: str, n_components: int, in_channels: int, adj: torch.Tensor, T: int = 12, Tout: int = 12,
                seed: int = 0, dtype=torch.float32) -> Params:
    """Random-init parameters with the reference's names / shapes / init rule:
    xavier-normal for ndim>=2, U(-s, s), s = size(0)**-0.5 for 1-D (including the
    LayerNorm affine).  The RNG stream is NOT the reference's (tests copy weights
    with ``state_dict`` instead of relying on RNG order)."""
    g = torch.Generator().manual_seed(seed)
    width, dil = MODEL_WIDTH[model], MODEL_DILATIONS[model]
    N = adj.shape[0]
    shapes: Dict[str, tuple] = {}
    shapes["te.d_ebd.weight"] = (7, n_components * N * Tout)
    shapes["te.h_ebd.weight"] = (24, n_components * N * Tout)
    for m in range(n_components):
        chans = [in_channels] + [width] * len(dil)
        for j, d in enumerate(dil):
            C, O = chans[j], chans[j + 1]
            b = O // 3
            pre = f"tpcs.{m}.tgacns.{j}."
            shapes[pre + "ln.weight"] = (T,)
            shapes[pre + "ln.bias"] = (T,)
            shapes[pre + "res.weight"] = (O, C, 1, 1)
            shapes[pre + "res.bias"] = (O,)
            shapes[pre + "cacn.seq.0.Wc"] = (T, T)
            shapes[pre + "cacn.seq.0.alpha"] = (N,)
            shapes[pre + "cacn.seq.1.weight"] = (b, C, 1, 1)
            shapes[pre + "cacn.seq.1.bias"] = (b,)
            shapes[pre + "tacn.seq.0.Wt1"] = (10, N)
            shapes[pre + "tacn.seq.0.Wt2"] = (10, N)
            shapes[pre + "tacn.seq.0.alpha"] = (C,)
            cin = C
            for i, _ in enumerate(d):
                shapes[f"{pre}tacn.seq.{1 + 2 * i}.weight"] = (b, cin, 1, 2)
                shapes[f"{pre}tacn.seq.{1 + 2 * i}.bias"] = (b,)
                cin = b
            shapes[pre + "gacn.W"] = (b, C)
            shapes[pre + "gacn.gatt.Wg"] = (T, T)
            shapes[pre + "gacn.gatt.alpha"] = (C,)
        shapes[f"tpcs.{m}.ln.weight"] = (T,)
        shapes[f"tpcs.{m}.ln.bias"] = (T,)
        shapes[f"tpcs.{m}.fc.weight"] = (Tout, T, 1, width)
        shapes[f"tpcs.{m}.fc.bias"] = (Tout,)
    p: Params = {}
    for k, s in shapes.items():
        if len(s) >= 2:
            rf = 1
            for v in s[2:]:
                rf *= v
            fan_in, fan_out = s[1] * rf, s[0] * rf
            std = (2.0 / (fan_in + fan_out)) ** 0.5
            p[k] = (torch.randn(s, generator=g, dtype=torch.float64) * std).to(dtype)
        else:
            lim = s[0] ** -0.5
            p[k] = ((torch.rand(s, generator=g, dtype=torch.float64) * 2 - 1) * lim).to(dtype)
    p["adj"] = adj.to(dtype)
    return p


def train_step_loss(p: Params, model: str, X, H, D, Y, delta: float = 50.0) -> torch.Tensor:
    """Forward + Huber loss exactly as the train branch of ``Engine.run_epoch``
    (engine.py:54-56) computes it (autocast is a no-op on CPU tensors)."""
    return huber(msgat(X, H, D, p, MODEL_DILATIONS[model]), Y, delta)
