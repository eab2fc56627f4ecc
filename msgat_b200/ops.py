"""torch.autograd.Functions over the C ABI (``include/msgat_sm100.h``).

``meam_op`` is the one MEAM-shaped operator every layer class of the reference maps to
(``MEAM`` msgat.py:121, ``CACN``/``TACN``/``GACN`` msgat.py:96/76/25 and the three bare
attentions attention.py:91/62/32); ``head_op`` is the TPC head fused with the time gate
(msgat.py:158-160,204).  PyTorch only owns memory and streams here: every FLOP of the
path runs in the hand-written kernels of ``csrc/``.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Dict, Optional, Sequence

import torch

from . import _lib

SLOTS = (["ln_w", "ln_b", "res_w", "res_b", "ca_Wc", "ca_alpha", "ca_w", "ca_b", "ta_Wt1", "ta_Wt2", "ta_alpha"]
         + [f"conv_{k}{i}" for i in range(_lib.MAX_CONV) for k in ("w", "b")] + ["sa_W", "sa_Wg", "sa_alpha"])


class SparseAdjacency:
    """CSR + CSC view of the constant normalised adjacency (msgat.py:190), built on the
    device by ``msgat_adj_count`` / ``msgat_adj_fill``."""

    def __init__(self, adj: torch.Tensor):
        lib = _lib.get()
        adj = adj.detach().to(torch.float32).contiguous()
        N = adj.shape[0]
        assert adj.shape == (N, N)
        dev = adj.device
        st = _lib.stream_of(adj)
        rn = torch.empty(N, dtype=torch.int32, device=dev)
        cn = torch.empty(N, dtype=torch.int32, device=dev)
        _lib.check(lib.msgat_adj_count(_lib.ptr(adj), N, _lib.ptr(rn), _lib.ptr(cn), st), "msgat_adj_count")
        zero = torch.zeros(1, dtype=torch.int32, device=dev)
        self.row_ptr = torch.cat([zero, torch.cumsum(rn, 0, dtype=torch.int32)])
        self.csc_ptr = torch.cat([zero, torch.cumsum(cn, 0, dtype=torch.int32)])
        nnz = int(self.row_ptr[-1].item())
        self.N, self.nnz = N, nnz
        self.col_idx = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
        self.val = torch.empty(max(nnz, 1), dtype=torch.float32, device=dev)
        self.csc_row = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
        self.csc_eid = torch.empty(max(nnz, 1), dtype=torch.int32, device=dev)
        _lib.check(lib.msgat_adj_fill(_lib.ptr(adj), N, _lib.ptr(self.row_ptr), _lib.ptr(self.csc_ptr),
                                      _lib.ptr(self.col_idx), _lib.ptr(self.val), _lib.ptr(self.csc_row),
                                      _lib.ptr(self.csc_eid), st), "msgat_adj_fill")
        self.struct = _lib.Adj(N, nnz, *[t.data_ptr() for t in (self.row_ptr, self.col_idx, self.val, self.csc_ptr,
                                                                 self.csc_row, self.csc_eid)])


_ADJ_CACHE: Dict[int, tuple] = {}


def sparse_adjacency(adj: torch.Tensor) -> SparseAdjacency:
    """CSR/CSC of ``adj``, cached per tensor OBJECT: an entry holds a weak reference to the tensor it was built
    from and that tensor's version counter, and is used only while both still match.  (Keying on the storage
    address alone would hand a new model whose adjacency the caching allocator placed at a freed model's
    address the old graph's CSR.)  The adjacency is a frozen parameter (msgat.py:190), so a model builds its
    view once."""
    key = id(adj)
    ent = _ADJ_CACHE.get(key)
    if ent is not None:
        ref, version, ptr, sp = ent
        if ref() is adj and version == adj._version and ptr == adj.data_ptr():
            return sp
    for k in [k for k, e in _ADJ_CACHE.items() if e[0]() is None]:   # drop entries of dead tensors
        del _ADJ_CACHE[k]
    sp = SparseAdjacency(adj)
    _ADJ_CACHE[key] = (weakref.ref(adj), adj._version, adj.data_ptr(), sp)
    return sp


def _sink_of(p):
    """The registered gradient sink of a parameter (see ``distributed.FlatGradients``), if it matches the parameter."""
    sk = getattr(p, "_msgat_grad_sink", None) if p is not None else None
    if sk is None or sk.shape != p.shape or sk.dtype != torch.float32 or sk.device != p.device or not sk.is_contiguous():
        return None
    return sk


class MeamConfig:
    def __init__(self, Oc: int, Ot: int, Os: int, flags: int, dilations: Sequence[int] = (), precision: int = 0,
                 dout_masked: bool = False, x_is_relu: bool = False):
        self.Oc, self.Ot, self.Os, self.flags = Oc, Ot, Os, flags
        self.dilations = tuple(int(d) for d in dilations)
        self.precision, self.dout_masked, self.x_is_relu = precision, dout_masked, x_is_relu
        if len(self.dilations) > _lib.MAX_CONV:
            raise ValueError(f"TACN chains longer than {_lib.MAX_CONV} convs are not supported")


def _desc(cfg: MeamConfig, B: int, Cc: int, N: int, M: int, x_sb: int, x_sc: int, x_sm: int, backward: bool) -> _lib.MeamDesc:
    flags = cfg.flags
    if backward:
        flags |= (_lib.F_DOUT_MASKED if cfg.dout_masked else 0) | (_lib.F_X_IS_RELU if cfg.x_is_relu else 0)
    d = _lib.MeamDesc(B, Cc, N, cfg.Oc, cfg.Ot, cfg.Os, flags, len(cfg.dilations))
    for i, v in enumerate(cfg.dilations):
        d.dil[i] = v
    d.precision = cfg.precision
    d.M = M
    d.x_sb, d.x_sc, d.x_sm = x_sb, x_sc, x_sm
    return d


def _plane_contiguous(x: torch.Tensor) -> torch.Tensor:
    """The (n, t) plane of every channel must be one contiguous, 16-byte aligned run; the leading strides are free."""
    if x.stride(-1) == 1 and x.stride(-2) == x.shape[-1] and x.data_ptr() % 16 == 0 \
            and all(st % 4 == 0 for st in x.stride()[:-2]):
        return x
    return x.contiguous()


def _param_struct(s, tensors: Sequence[Optional[torch.Tensor]]):
    byname = dict(zip(SLOTS, tensors))
    for n in _lib._PARAM_FIELDS + ["sa_W", "sa_Wg", "sa_alpha"]:
        t = byname[n]
        setattr(s, n, None if t is None else t.data_ptr())
    for i in range(_lib.MAX_CONV):
        w, b = byname[f"conv_w{i}"], byname[f"conv_b{i}"]
        s.conv_w[i] = None if w is None else w.data_ptr()
        s.conv_b[i] = None if b is None else b.data_ptr()
    return s


def _param_array(cls, per_component: Sequence[Sequence[Optional[torch.Tensor]]]):
    arr = (cls * len(per_component))()
    for m, tensors in enumerate(per_component):
        _param_struct(arr[m], tensors)
    return arr


def _strides(x: torch.Tensor, layout: str):
    """-> (M, B, C, N, x_sm, x_sb, x_sc) of a grouped input: layout "bm" is [B, M, C, N, 12] (the X of
    MSGAT.forward, msgat.py:202: component m is X[:, m]), "mb" is [M, B, C, N, 12] (a grouped operator's output)."""
    if layout == "bm":
        B, M, Cc, N, T = x.shape
        return M, B, Cc, N, x.stride(1), x.stride(0), x.stride(2)
    M, B, Cc, N, T = x.shape
    return M, B, Cc, N, x.stride(0), x.stride(1), x.stride(2)


class _MeamGroupFunction(torch.autograd.Function):
    """The MEAM-shaped operator on M same-shaped components at once (one launch per kernel for all of them):
    x [B,M,C,N,12] or [M,B,C,N,12] -> out [M,B,O,N,12]; ``params`` = M consecutive groups of ``len(SLOTS)``."""

    @staticmethod
    def forward(ctx, x, layout: str, cfg: MeamConfig, adj: Optional[SparseAdjacency], *params):
        lib = _lib.get()
        if x.dtype != torch.float32:
            x = x.float()  # the path computes in fp32 whatever autocast says (mixed precision)
        if x.shape[-1] != 12:
            raise ValueError("msgat_b200 kernels are specialised for 12 timesteps (main.py:45)")
        x = _plane_contiguous(x)
        M, B, Cc, N, x_sm, x_sb, x_sc = _strides(x, layout)
        if M > _lib.MAX_GROUP:
            raise ValueError(f"at most {_lib.MAX_GROUP} components per grouped call")
        ns = len(SLOTS)
        assert len(params) == M * ns
        ctx.sinks = [_sink_of(p) for p in params]
        params = tuple(None if p is None else p.detach().float().contiguous() for p in params)
        per = [params[m * ns:(m + 1) * ns] for m in range(M)]
        d = _desc(cfg, B, Cc, N, M, x_sb, x_sc, x_sm, backward=False)
        nnz = adj.nnz if adj is not None else 0
        sv, sc = C.c_int64(), C.c_int64()
        _lib.check(lib.msgat_meam_sizes(C.byref(d), nnz, C.byref(sv), C.byref(sc)), "msgat_meam_sizes")
        saved = torch.empty(max(sv.value, 4) * M, dtype=torch.float32, device=x.device)
        # the scratch of the forward is not needed again: the backward allocates its own (nothing large is pinned
        # between the two, and nothing beyond `saved` under no_grad)
        scratch = torch.empty(max(sc.value, 4) * M, dtype=torch.float32, device=x.device)
        out = torch.empty(M, B, cfg.Oc + cfg.Ot + cfg.Os, N, 12, dtype=torch.float32, device=x.device)
        ps = _param_array(_lib.MeamParams, per)
        _lib.check(lib.msgat_meam_forward(C.byref(d), ps, C.byref(adj.struct) if adj is not None else None,
                                          _lib.ptr(x), _lib.ptr(out), _lib.ptr(saved), _lib.ptr(scratch),
                                          _lib.stream_of(x)), "msgat_meam_forward")
        ctx.cfg, ctx.adj, ctx.layout, ctx.scratch_floats = cfg, adj, layout, max(sc.value, 4) * M
        ctx.save_for_backward(x, out, saved, *[p for p in params if p is not None])
        ctx.present = [p is not None for p in params]
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.get()
        x, out, saved, *rest = ctx.saved_tensors
        it = iter(rest)
        params = tuple(next(it) if pr else None for pr in ctx.present)
        cfg, adj, layout = ctx.cfg, ctx.adj, ctx.layout
        M, B, Cc, N, x_sm, x_sb, x_sc = _strides(x, layout)
        ns = len(SLOTS)
        dout = dout.float().contiguous()
        d = _desc(cfg, B, Cc, N, M, x_sb, x_sc, x_sm, backward=True)
        dx = torch.empty(M, B, Cc, N, 12, dtype=torch.float32, device=x.device)
        # a parameter with a registered gradient sink (distributed.FlatGradients(direct=True)) gets its gradient
        # written straight into the flat buffer by the kernels; autograd is handed None for it (no add kernel)
        grads = [None if p is None else (sk if sk is not None else torch.empty_like(p)) for p, sk in zip(params, ctx.sinks)]
        scratch = torch.empty(ctx.scratch_floats, dtype=torch.float32, device=x.device)
        ps = _param_array(_lib.MeamParams, [params[m * ns:(m + 1) * ns] for m in range(M)])
        gs = _param_array(_lib.MeamGrads, [grads[m * ns:(m + 1) * ns] for m in range(M)])
        _lib.check(lib.msgat_meam_backward(C.byref(d), ps, C.byref(adj.struct) if adj is not None else None,
                                           _lib.ptr(x), _lib.ptr(out), _lib.ptr(dout), _lib.ptr(saved),
                                           _lib.ptr(scratch), _lib.ptr(dx), gs, _lib.stream_of(x)),
                   "msgat_meam_backward")
        if layout == "bm":
            dx = dx.permute(1, 0, 2, 3, 4)
        return (dx, None, None, None, *[None if sk is not None else g for g, sk in zip(grads, ctx.sinks)])


def meam_group_op(x: torch.Tensor, layout: str, cfg: MeamConfig, adj: Optional[SparseAdjacency],
                  params: Sequence[Dict[str, Optional[torch.Tensor]]]) -> torch.Tensor:
    """The same layer of M components (the loop of msgat.py:203-204) as one grouped operator.
    ``x`` is [B,M,C,N,12] (``layout="bm"``) or [M,B,C,N,12] (``"mb"``); returns [M,B,O,N,12]."""
    flat = [p.get(n) for p in params for n in SLOTS]
    with torch.autocast(device_type=x.device.type, enabled=False):
        return _MeamGroupFunction.apply(x, layout, cfg, adj, *flat)


def meam_op(x: torch.Tensor, cfg: MeamConfig, adj: Optional[SparseAdjacency],
            params: Dict[str, Optional[torch.Tensor]]) -> torch.Tensor:
    """Run the MEAM-shaped operator on one component; ``params`` maps slot names (``SLOTS``) to tensors."""
    return meam_group_op(x.unsqueeze(0), "mb", cfg, adj, [params])[0]


class _HeadFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ln_w, ln_b, fc_w, fc_b, gate, acc, x_is_relu):
        lib = _lib.get()
        x = x.float().contiguous()
        B, Cc, N, T = x.shape
        if T != 12:
            raise ValueError("msgat_b200 kernels are specialised for 12 timesteps (main.py:45)")
        Tout = fc_w.shape[0]
        ctx.sinks = [_sink_of(t) for t in (ln_w, ln_b, fc_w, fc_b)]
        ln_w, ln_b, fc_w, fc_b = (t.detach().float().contiguous() for t in (ln_w, ln_b, fc_w, fc_b))
        if gate is not None:
            gate = gate.detach().float()
            if gate.stride(2) != 1 or gate.stride(1) != Tout:
                gate = gate.contiguous()
        d = _lib.HeadDesc(B, Cc, N, Tout, 0 if acc is None else 1, 1 if x_is_relu else 0,
                          0 if gate is None else gate.stride(0))
        head_out = torch.empty(B, N, Tout, dtype=torch.float32, device=x.device)
        out = torch.empty_like(head_out) if acc is None else acc.detach().float().contiguous().clone()
        _lib.check(lib.msgat_head_forward(C.byref(d), _lib.ptr(x), _lib.ptr(ln_w), _lib.ptr(ln_b), _lib.ptr(fc_w),
                                          _lib.ptr(fc_b), _lib.ptr(gate), _lib.ptr(head_out), _lib.ptr(out),
                                          _lib.stream_of(x)), "msgat_head_forward")
        ctx.save_for_backward(x, ln_w, ln_b, fc_w, gate if gate is not None else head_out, head_out)
        ctx.has_gate, ctx.has_acc, ctx.x_is_relu = gate is not None, acc is not None, x_is_relu
        ctx.gate_needs_grad = gate is not None and ctx.needs_input_grad[5]
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.get()
        x, ln_w, ln_b, fc_w, gate, head_out = ctx.saved_tensors
        if not ctx.has_gate:
            gate = None
        dout = dout.float().contiguous()
        B, Cc, N, T = x.shape
        Tout = fc_w.shape[0]
        d = _lib.HeadDesc(B, Cc, N, Tout, 0, 1 if ctx.x_is_relu else 0, 0 if gate is None else gate.stride(0))
        dx = torch.empty_like(x)
        sk = ctx.sinks
        g_lnw = sk[0] if sk[0] is not None else torch.empty_like(ln_w)
        g_lnb = sk[1] if sk[1] is not None else torch.empty_like(ln_b)
        g_fcw = sk[2] if sk[2] is not None else torch.empty_like(fc_w)
        g_fcb = sk[3] if sk[3] is not None else torch.empty(Tout, dtype=torch.float32, device=x.device)
        g_gate = torch.empty_like(head_out) if ctx.gate_needs_grad else None
        scratch = torch.empty(max(int(lib.msgat_head_scratch_floats(C.byref(d))), 4), dtype=torch.float32,
                              device=x.device)
        _lib.check(lib.msgat_head_backward(C.byref(d), _lib.ptr(x), _lib.ptr(ln_w), _lib.ptr(ln_b), _lib.ptr(fc_w),
                                           _lib.ptr(gate), _lib.ptr(head_out), _lib.ptr(dout), _lib.ptr(dx),
                                           _lib.ptr(g_lnw), _lib.ptr(g_lnb), _lib.ptr(g_fcw), _lib.ptr(g_fcb),
                                           _lib.ptr(g_gate), _lib.ptr(scratch), _lib.stream_of(x)),
                   "msgat_head_backward")
        ret = [None if k is not None else g for g, k in zip((g_lnw, g_lnb, g_fcw, g_fcb), sk)]
        return (dx, *ret, g_gate, (dout if ctx.has_acc else None), None)


def head_op(x, ln_w, ln_b, fc_w, fc_b, gate=None, acc=None, x_is_relu=False):
    """TPC head + gate (+ running sum over components): returns ``acc + head(x) * gate``."""
    with torch.autocast(device_type=x.device.type, enabled=False):
        return _HeadFunction.apply(x, ln_w, ln_b, fc_w, fc_b, gate, acc, x_is_relu)


def _ptr_array(tensors):
    return (C.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])


class _HeadsFunction(torch.autograd.Function):
    """The heads of all M components and the gated sum over them (msgat.py:158-160, 203-204) as grouped launches:
    x [M,B,C,N,12], gate [B,M,N,Tout] -> [B,N,Tout]; ``params`` = ln_w * M, ln_b * M, fc_w * M, fc_b * M."""

    @staticmethod
    def forward(ctx, x, gate, x_is_relu, *params):
        lib = _lib.get()
        x = x.float().contiguous()
        M, B, Cc, N, T = x.shape
        if T != 12:
            raise ValueError("msgat_b200 kernels are specialised for 12 timesteps (main.py:45)")
        ctx.sinks = [_sink_of(t) for t in params]
        params = [t.detach().float().contiguous() for t in params]
        ln_w, ln_b, fc_w, fc_b = params[:M], params[M:2 * M], params[2 * M:3 * M], params[3 * M:]
        Tout = fc_w[0].shape[0]
        if gate is not None:
            gate = gate.detach().float()
            if gate.stride(3) != 1 or gate.stride(2) != Tout or gate.stride(0) == 0:
                gate = gate.contiguous()
        d = _lib.HeadsDesc(B, Cc, N, Tout, M, 1 if x_is_relu else 0, x.stride(0),
                           0 if gate is None else gate.stride(0), 0 if gate is None else gate.stride(1))
        head_out = torch.empty(M, B, N, Tout, dtype=torch.float32, device=x.device)
        out = torch.empty(B, N, Tout, dtype=torch.float32, device=x.device)
        _lib.check(lib.msgat_heads_forward(C.byref(d), _lib.ptr(x), _ptr_array(ln_w), _ptr_array(ln_b), _ptr_array(fc_w),
                                           _ptr_array(fc_b), _lib.ptr(gate), _lib.ptr(head_out), _lib.ptr(out),
                                           _lib.stream_of(x)), "msgat_heads_forward")
        ctx.save_for_backward(x, head_out, *([gate] if gate is not None else []), *ln_w, *ln_b, *fc_w)
        ctx.has_gate, ctx.x_is_relu, ctx.M = gate is not None, x_is_relu, M
        ctx.gate_needs_grad = gate is not None and ctx.needs_input_grad[1]
        return out

    @staticmethod
    def backward(ctx, dout):
        lib = _lib.get()
        M = ctx.M
        x, head_out, *rest = ctx.saved_tensors
        gate = rest.pop(0) if ctx.has_gate else None
        ln_w, ln_b, fc_w = rest[:M], rest[M:2 * M], rest[2 * M:3 * M]
        dout = dout.float().contiguous()
        _, B, Cc, N, T = x.shape
        Tout = fc_w[0].shape[0]
        d = _lib.HeadsDesc(B, Cc, N, Tout, M, 1 if ctx.x_is_relu else 0, x.stride(0),
                           0 if gate is None else gate.stride(0), 0 if gate is None else gate.stride(1))
        dx = torch.empty_like(x)
        sk = ctx.sinks
        like = list(ln_w) + list(ln_b) + list(fc_w) + [torch.empty(0)] * M
        grads = []
        for i in range(4 * M):
            if sk[i] is not None:
                grads.append(sk[i])
            elif i >= 3 * M:
                grads.append(torch.empty(Tout, dtype=torch.float32, device=x.device))
            else:
                grads.append(torch.empty_like(like[i]))
        g_gate = torch.empty_like(gate) if ctx.gate_needs_grad else None   # same [B,M,N,Tout] strides as the gate
        if g_gate is not None and g_gate.stride() != gate.stride():
            raise RuntimeError("heads backward: gate must be dense")
        scratch = torch.empty(max(int(lib.msgat_heads_scratch_floats(C.byref(d))), 4) * M, dtype=torch.float32, device=x.device)
        _lib.check(lib.msgat_heads_backward(C.byref(d), _lib.ptr(x), _ptr_array(ln_w), _ptr_array(ln_b), _ptr_array(fc_w),
                                            _lib.ptr(gate), _lib.ptr(head_out), _lib.ptr(dout), _lib.ptr(dx),
                                            _ptr_array(grads[:M]), _ptr_array(grads[M:2 * M]), _ptr_array(grads[2 * M:3 * M]),
                                            _ptr_array(grads[3 * M:]), _lib.ptr(g_gate), _lib.ptr(scratch),
                                            _lib.stream_of(x)), "msgat_heads_backward")
        return (dx, g_gate, None, *[None if k is not None else g for g, k in zip(grads, sk)])


def heads_op(x, gate, ln_w, ln_b, fc_w, fc_b, x_is_relu=False):
    """``sum_m head_m(x[m]) * gate[:, m]`` for the M components at once; the parameter arguments are lists of M."""
    with torch.autocast(device_type=x.device.type, enabled=False):
        return _HeadsFunction.apply(x, gate, x_is_relu, *ln_w, *ln_b, *fc_w, *fc_b)
