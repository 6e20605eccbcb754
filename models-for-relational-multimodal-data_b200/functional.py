"""torch.autograd Functions over the C-ABI kernels (include/b200mp.h).

All edge tensors handled here are in *e-storage order* = CSC order of the graph
(``EdgeStructure.fwd``); ``permute_rows`` moves between the caller's edge order and it.
Saved tensors are never modified, so ``retain_graph=True`` + repeated backward (MoCo,
/root/reference/src/nn/weighting/MoCo.py:45) works.
"""
from __future__ import annotations

from typing import NamedTuple, Optional, Sequence

import torch

from . import _lib as L
from .graph import CHUNK, LIGHT_MAX, EdgeStructure, SegmentIndex


class KernelStats:
    """Optional per-kernel accounting (bench.py / profiling): number of kernel launches issued
    through the C ABI and, when ``timing`` is on, a CUDA-event pair around each call recorded on
    the launching (current) stream."""

    def __init__(self, timing: bool = False):
        self.timing = timing
        self.launches = 0
        self.calls = {}        # name -> number of calls
        self.events = {}       # name -> [(start, stop)]

    def summary(self):
        """name -> (calls, total_ms); call after torch.cuda.synchronize()."""
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in self.events.items()}


_STATS = None


def set_kernel_stats(stats):
    global _STATS
    _STATS = stats


class _track:
    def __init__(self, name, n_kernels=1):
        self.name, self.n = name, n_kernels

    def __enter__(self):
        st = _STATS
        self.ev = None
        if st is not None:
            st.launches += self.n
            st.calls[self.name] = st.calls.get(self.name, 0) + 1
            if st.timing:
                self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                self.ev[0].record()

    def __exit__(self, *exc):
        if self.ev is not None:
            self.ev[1].record()
            _STATS.events.setdefault(self.name, []).append(self.ev)
        return False


def _f32c(t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise TypeError(f"b200mp kernels compute in fp32; got {t.dtype}")
    return t.contiguous()


_EDGE_DTYPES = (torch.float32, torch.bfloat16)


def _edge_c(t: torch.Tensor) -> torch.Tensor:
    """an [E,H] edge tensor: fp32, or bf16 STORAGE (config 3: half the bytes of the streams; arithmetic stays fp32 in the kernels)"""
    if t.dtype not in _EDGE_DTYPES:
        raise TypeError(f"b200mp edge tensors are stored as fp32 or bf16; got {t.dtype}")
    return t.contiguous()


def _bf16_ok(t: torch.Tensor):
    if t.shape[-1] != 64 or not row_maps_available():
        raise L.B200MPError("bf16 storage runs on the H = 64 tensor-core kernels only")


def _H(t: torch.Tensor) -> int:
    H = t.shape[-1]
    L.tile_edges(H)      # raises for unsupported widths
    return H


# ----------------------------------------------------------------------------- row permutation
class _PermuteRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, idx, scatter):
        x = _f32c(x)
        L.require_cuda(x, idx)
        ctx.save_for_backward(idx)
        ctx.scatter = scatter
        return _permute(x, idx, scatter)

    @staticmethod
    def backward(ctx, g):
        (idx,) = ctx.saved_tensors
        return _permute(_f32c(g), idx, not ctx.scatter), None, None


def _permute(x, idx, scatter):
    if x.dtype != torch.float32:
        raise TypeError("row permutation passes are fp32 only (bf16 storage uses the row-map kernels, H = 64)")
    n, H = x.shape
    if H % 4:
        raise L.B200MPError("row width must be a multiple of 4")
    out = torch.empty_like(x)
    fn = L.lib().b200mp_scatter_rows if scatter else L.lib().b200mp_gather_rows
    with torch.cuda.device(x.device), _track("permute_rows"):
        L.check(fn(L.ptr(x), L.ptr(idx), L.ptr(out), n, H, L.stream()), "b200mp_permute_rows")
    return out


def row_maps_available() -> bool:
    """The eid / out_eid row maps of the edge update live in the tensor-core kernels (H = 64);
    B200MP_SIMT=1 (A/B runs on the SIMT kernels) switches them off."""
    import os
    return os.environ.get("B200MP_SIMT", "0") != "1"


def scatter_rows_i32(x, idx32):
    """out[idx32[i]] = x[i] (no autograd; rare fallback of the row-map paths)."""
    return _permute(x, idx32.long(), True)


def gather_rows(x: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """out[i] = x[idx[i]] for a permutation idx (backward = scatter, no atomics)."""
    return _PermuteRows.apply(x, idx, False)


def scatter_rows(x: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """out[idx[i]] = x[i] for a permutation idx."""
    return _PermuteRows.apply(x, idx, True)


# ----------------------------------------------------------------------------- node-level linears
def _split_tf32(x):
    x = x.contiguous()
    hi, lo = torch.empty_like(x), torch.empty_like(x)
    with torch.cuda.device(x.device), _track("split_tf32"):
        L.check(L.lib().b200mp_split_tf32(L.ptr(x), L.ptr(hi), L.ptr(lo), x.numel(), L.stream()), "b200mp_split_tf32")
    return hi, lo


def _use_mm3(x, W):
    """3xTF32 pays off only where the fp32 GEMM is FLOP-bound: many rows and a wide contraction
    (the [N,4H] x [4H,3H] post GEMM of PNAConv); H x H node linears are HBM-bound in plain fp32."""
    return x.shape[0] >= 4096 and W.shape[1] >= 256 and W.shape[0] % 4 == 0 and W.shape[1] % 4 == 0


def _mm3(a, b, ta=False, tb=False):
    """op(a) @ op(b) in fp32 accuracy on the tensor cores: 3xTF32 (lo*hi + hi*lo + hi*hi) as three
    TF32 library GEMMs accumulating in fp32.  a, b are (hi, lo) pairs from _split_tf32."""
    (a_hi, a_lo), (b_hi, b_lo) = a, b
    if ta:
        a_hi, a_lo = a_hi.t(), a_lo.t()
    if tb:
        b_hi, b_lo = b_hi.t(), b_lo.t()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        out = torch.mm(a_lo, b_hi)
        out.addmm_(a_hi, b_lo)
        out.addmm_(a_hi, b_hi)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    return out


def _w_image(W):
    """[NO, K] weight -> the pre-split, pre-tiled operand image b200mp_node_linear expects:
    [K/32][2 (hi, lo)][8][NO][4], hi = W rounded to tf32 (same rounding as the kernels), lo = W - hi."""
    W = W.contiguous()
    NO, K = W.shape
    hi = ((W.view(torch.int32) + 0x1000) & -8192).view(torch.float32)
    lo = W - hi
    tile = lambda P: P.view(NO, K // 32, 8, 4).permute(1, 2, 0, 3)
    return torch.stack([tile(hi), tile(lo)], dim=1).contiguous()


def _node_kernel_ok(rows, K, NO):
    """Y[rows, NO] = X[rows, K] W^T on the tcgen05 node kernel (node_gemm_umma.cu)."""
    return rows >= 4096 and K % 32 == 0 and K >= 128 and NO % 64 == 0 and 64 <= NO <= 1024


def _col_blocks(NO):
    """column blocks of <= 256 (the accumulator width of the node kernels) covering NO outputs"""
    return [(c0, min(256, NO - c0)) for c0 in range(0, NO, 256)]


def _node_linear(x, W, b):
    N, K = x.shape
    NO = W.shape[0]
    y = torch.empty((N, NO), dtype=torch.float32, device=x.device)
    for c0, nc in _col_blocks(NO):           # more than 256 outputs (post_nn at H = 128): column blocks of the same output
        img = _w_image(W[c0:c0 + nc])
        with torch.cuda.device(x.device), _track("node_linear"):
            L.check(L.lib().b200mp_node_linear_ld(L.ptr(x), N, K, K, L.ptr(img), nc, L.ptr(None if b is None else b[c0:c0 + nc]),
                                                  L.ptr(y[:, c0:]), NO, L.stream()), "b200mp_node_linear_ld")
    return y


def _node_wgrad_ok(rows, K, NO):
    return rows >= 4096 and K % 128 == 0 and NO % 64 == 0 and 64 <= NO <= 1024


def _node_wgrad(x, y):
    """x^T y  ([K, NO]) over the N rows, per-CTA partials summed here."""
    N, K = x.shape
    NO = y.shape[1]
    npart = torch.cuda.get_device_properties(x.device).multi_processor_count
    outs = []
    for c0, nc in _col_blocks(NO):
        part = torch.empty((npart, K, nc), dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device), _track("node_linear_wgrad"):
            L.check(L.lib().b200mp_node_linear_wgrad(L.ptr(x), K, K, L.ptr(y[:, c0:]), NO, nc, N, L.ptr(part), npart, L.stream()),
                    "b200mp_node_linear_wgrad")
        outs.append(part.sum(0))
    return outs[0] if len(outs) == 1 else torch.cat(outs, 1)


class _Linear3x(torch.autograd.Function):
    """y = x W^T (+ b) for N-row node tensors with fp32 accuracy on the tensor cores: the wide
    post linear runs on the hand-written tcgen05 kernels (one pass over HBM per product); shapes
    they do not cover fall back to three TF32 library GEMMs, narrow ones to plain fp32."""

    @staticmethod
    def forward(ctx, x, W, b):
        x, W = _f32c(x), _f32c(W)
        L.require_cuda(x, W)
        ctx.save_for_backward(x, W)
        ctx.has_bias = b is not None
        if _node_kernel_ok(x.shape[0], W.shape[1], W.shape[0]):
            return _node_linear(x, W, None if b is None else _f32c(b))
        if not _use_mm3(x, W):
            return torch.nn.functional.linear(x, W, b)
        y = _mm3(_split_tf32(x), _split_tf32(W), tb=True)
        if b is not None:
            y += b
        return y

    @staticmethod
    def backward(ctx, g):
        x, W = ctx.saved_tensors
        g = _f32c(g)
        N, (NO, K) = x.shape[0], W.shape
        gb = g.sum(0) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        gx = gW = None
        mm3 = _use_mm3(x, W)
        gs = None
        if ctx.needs_input_grad[0]:
            if _node_kernel_ok(N, NO, K):
                gx = _node_linear(g, W.t(), None)
            elif mm3:
                gs = _split_tf32(g)
                gx = _mm3(gs, _split_tf32(W))
            else:
                gx = g @ W
        if ctx.needs_input_grad[1]:
            if _node_wgrad_ok(N, K, NO):
                gW = _node_wgrad(x, g).t()
            elif _node_wgrad_ok(N, NO, K):
                gW = _node_wgrad(g, x)
            elif mm3:
                gW = _mm3(gs if gs is not None else _split_tf32(g), _split_tf32(x), ta=True)
            else:
                gW = g.t() @ x
        return gx, gW, gb


def _node_pair_wgrad(g0, g1, x):
    """(g0^T x, g1^T x) for [N,64] tensors on the tcgen05 weight-gradient kernel (one launch); g1 None: (g0^T x, zeros), g0 read once."""
    N = x.shape[0]
    npart = torch.cuda.get_device_properties(x.device).multi_processor_count
    part = torch.empty((npart, 128, 64), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device), _track("node_pair_wgrad"):
        L.check(L.lib().b200mp_node_pair_wgrad(L.ptr(g0), L.ptr(g1), L.ptr(x), N, L.ptr(part), npart, L.stream()),
                "b200mp_node_pair_wgrad")
    G = part.sum(0)
    return G[:64], G[64:]


def _node_linear64_colsum(xs, Wcat):
    """([x_0 | x_1 | ..] Wcat^T  [N,64],  column sums of [x_0 | x_1 | ..]  [64 n]) in one pass of the tcgen05 node kernel"""
    N, dev = xs[0].shape[0], xs[0].device
    K = 64 * len(xs)
    y = torch.empty((N, 64), dtype=torch.float32, device=dev)
    npart = 2 * torch.cuda.get_device_properties(dev).multi_processor_count
    part = torch.zeros((npart, K), dtype=torch.float32, device=dev)
    img = _w_image(Wcat)
    px = [L.ptr(t) for t in xs] + [None] * (3 - len(xs))
    with torch.cuda.device(dev), _track("node_linear64"):
        L.check(L.lib().b200mp_node_linear64_colsum(px[0], px[1], px[2], len(xs), N, L.ptr(img), None, L.ptr(y), L.ptr(part), npart,
                                                    L.stream()), "b200mp_node_linear64_colsum")
    return y, part.sum(0)


def _node_linear64(xs, Wcat, bias, ysplit):
    """[x_0 | x_1 | ..] Wcat^T + bias for [N,64] blocks on the tcgen05 node kernel; ysplit: the result as
    Wcat.shape[0] / 64 separate contiguous [N,64] tensors."""
    N, NO = xs[0].shape[0], Wcat.shape[0]
    dev = xs[0].device
    y = torch.empty((NO // 64, N, 64) if ysplit else (N, NO), dtype=torch.float32, device=dev)
    img = _w_image(Wcat)
    px = [L.ptr(t) for t in xs] + [None] * (3 - len(xs))
    with torch.cuda.device(dev), _track("node_linear64"):
        L.check(L.lib().b200mp_node_linear64(px[0], px[1], px[2], len(xs), N, L.ptr(img), NO, L.ptr(bias), L.ptr(y),
                                             1 if ysplit else 0, L.stream()), "b200mp_node_linear64")
    return list(y.unbind(0)) if ysplit else y


class _LinearGroup(torch.autograd.Function):
    """y_i = x W_i^T + b_i for several H x H node linears over the SAME input (PNAConv's hoisted
    pre_nn halves and post_nn self term, the edge update's P | Q).  Forward: fp32 library GEMMs
    (HBM bound).  Backward: the data gradients accumulate in the GEMM epilogue (no separate adds),
    the weight gradients -- reductions over all N rows, which the library runs as slow split-K
    GEMMs -- go two at a time through the tcgen05 weight-gradient kernel."""

    @staticmethod
    def forward(ctx, x, n, *params):
        x = _f32c(x)
        L.require_cuda(x)
        Ws, bs = params[:n], params[n:]
        ctx.n = n
        ctx.save_for_backward(x, *Ws)
        ctx.kernel_ok = x.shape[0] >= 4096 and x.shape[1] == 64 and n <= 3 and all(W.shape == (64, 64) for W in Ws)
        if ctx.kernel_ok:                    # one pass over x for all outputs (tcgen05, 3xTF32)
            bias = None
            if any(b is not None for b in bs):
                bias = torch.cat([b if b is not None else x.new_zeros(64) for b in bs])
            return tuple(_node_linear64([x], torch.cat(list(Ws), 0), bias, True))
        return tuple(torch.nn.functional.linear(x, W, b) for W, b in zip(Ws, bs))

    @staticmethod
    def backward(ctx, *gs):
        x, *Ws = ctx.saved_tensors
        n = ctx.n
        gs = [None if g is None else _f32c(g) for g in gs]
        gx = None
        colsum = None
        if ctx.needs_input_grad[0]:
            if ctx.kernel_ok and all(g is not None for g in gs):
                # gx = [g_0 | g_1 | ..] [W_0; W_1; ..]: one pass over the gradients, no separate adds -- and, where a bias wants
                # its gradient, the column sums of the g_i out of the same pass
                Wt = torch.cat(list(Ws), 0).t()
                if any(ctx.needs_input_grad[2 + n + i] for i in range(n)):
                    gx, colsum = _node_linear64_colsum(gs, Wt)
                else:
                    gx = _node_linear64(gs, Wt, None, False)
            else:
                for g, W in zip(gs, Ws):
                    if g is not None:
                        gx = g @ W if gx is None else gx.addmm_(g, W)
        gW = [None] * n
        want = [i for i in range(n) if gs[i] is not None and ctx.needs_input_grad[2 + i]]
        kernel_ok = x.shape[0] >= 4096 and x.shape[1] == 64 and all(Ws[i].shape == (64, 64) for i in want)
        if kernel_ok:
            for k in range(0, len(want), 2):
                i0 = want[k]
                i1 = want[k + 1] if k + 1 < len(want) else i0
                gW[i0], g1 = _node_pair_wgrad(gs[i0], gs[i1] if i1 != i0 else None, x)
                if i1 != i0:
                    gW[i1] = g1
        else:
            for i in want:
                gW[i] = gs[i].t() @ x
        if colsum is not None:
            gb = [colsum[64 * i:64 * (i + 1)] if ctx.needs_input_grad[2 + n + i] else None for i in range(n)]
        else:
            gb = [gs[i].sum(0) if (gs[i] is not None and ctx.needs_input_grad[2 + n + i]) else None for i in range(n)]
        return (gx, None, *gW, *gb)


def linear_group(x, weights, biases):
    """[x W^T + b for W, b in zip(weights, biases)] (b may be None) as one autograd node."""
    if x.dim() == 2 and x.shape[1] != 64 and x.is_cuda and all(_node_kernel_ok(x.shape[0], W.shape[1], W.shape[0]) for W in weights):
        return tuple(linear3x(x, W, b) for W, b in zip(weights, biases))      # H = 128: the general tcgen05 node kernel, one pass each
    return _LinearGroup.apply(x, len(weights), *weights, *biases)


class _LinearCat(torch.autograd.Function):
    """y = [x_0 | x_1 | x_2] W^T + b for [N,64] blocks WITHOUT building the concatenation (the `lin` of the reverse-message-passing
    convs over [x, a_in, a_out]: gine.py:35, pna.py:46): one pass of the tcgen05 node kernel with the blocks as K segments; backward =
    the split-output form of the same kernel for the data gradients, pair weight-gradient kernels for g^T x_i."""

    @staticmethod
    def forward(ctx, W, b, *xs):
        xs = tuple(_f32c(x) for x in xs)
        W = _f32c(W)
        L.require_cuda(W, *xs)
        ctx.n = len(xs)
        ctx.has_bias = b is not None
        ctx.save_for_backward(W, *xs)
        return _node_linear64(list(xs), W, None if b is None else _f32c(b), False)

    @staticmethod
    def backward(ctx, g):
        W, *xs = ctx.saved_tensors
        g = _f32c(g)
        n = ctx.n
        gxs = [None] * n
        if any(ctx.needs_input_grad[2:]):
            Wt = torch.cat([W[:, 64 * i:64 * (i + 1)].t() for i in range(n)], 0)          # y_i = g W_i
            gxs = _node_linear64([g], Wt, None, True)
        gW = None
        if ctx.needs_input_grad[0]:
            parts = []
            for k in range(0, n, 2):
                a, bb = _node_pair_wgrad(xs[k], xs[min(k + 1, n - 1)], g)                  # (x_k^T g, x_k+1^T g)
                parts.append(a.t())
                if k + 1 < n:
                    parts.append(bb.t())
            gW = torch.cat(parts, 1)
        gb = g.sum(0) if (ctx.has_bias and ctx.needs_input_grad[1]) else None
        return (gW, gb, *[gx if need else None for gx, need in zip(gxs, ctx.needs_input_grad[2:])])


def linear_cat(xs, W, b=None):
    """``F.linear(torch.cat(xs, 1), W, b)`` for up to three [N,64] blocks; N >= 4096 runs on the node kernel without the concatenation"""
    xs = list(xs)
    if 1 <= len(xs) <= 3 and all(x.dim() == 2 and x.shape[1] == 64 and x.dtype == torch.float32 for x in xs) and xs[0].shape[0] >= 4096 \
            and tuple(W.shape) == (64, 64 * len(xs)):
        return _LinearCat.apply(W, b, *xs)
    if all(x.dim() == 2 and x.is_cuda and x.dtype == torch.float32 for x in xs) and _node_kernel_ok(xs[0].shape[0], W.shape[1], W.shape[0]):
        return linear3x(torch.cat(xs, dim=1), W, b)          # H = 128: the general tcgen05 node kernel over the concatenation
    return torch.nn.functional.linear(torch.cat(xs, dim=1), W, b)


def linear3x(x, W, b=None):
    """torch.nn.functional.linear(x, W, b) for node-level [N, K] inputs, fp32-accurate on the
    tensor cores (three TF32 library GEMMs)."""
    return _Linear3x.apply(x, W, b)


# ----------------------------------------------------------------------------- H = 128 on the tensor cores
def wide_tc_available(H: int) -> bool:
    """H = 128 (pna.py:50 default width): the per-edge products run on the tcgen05 row-linear kernel (b200mp_row_linear: the node
    GEMM kernel with gathers / activation / mask / residual in its epilogue) and the weight gradients on b200mp_node_linear_wgrad,
    composed here; B200MP_SIMT=1 or B200MP_TC128=0 selects the fused SIMT kernels of edge_fwd.cu / edge_bwd.cu instead."""
    import os
    return H == 128 and row_maps_available() and os.environ.get("B200MP_TC128", "1") != "0"


def _row_linear(x, W, b=None, *, xrow=None, T1=None, idx1=None, T2=None, idx2=None, relu=False, mask=None, R=None, rrow=None,
                alpha=0.0, beta=1.0, R2=None, out=None, orow=None):
    """out[orow[k]] = alpha R[rrow[k]] + beta act(x[xrow[k]] W^T + b + T1[idx1[k]] + T2[idx2[k]]) * [mask[k] > 0] + R2[k]"""
    N = x.shape[0] if xrow is None else xrow.numel()
    K, NO = x.shape[1], W.shape[0]
    if out is None:
        out = torch.empty((N, NO), dtype=torch.float32, device=x.device)
    img = _w_image(W)
    L.check(L.lib().b200mp_row_linear(L.ptr(x), L.ptr(xrow), N, K, K, L.ptr(img), NO, L.ptr(b), L.ptr(T1), L.ptr(idx1), L.ptr(T2),
                                      L.ptr(idx2), 1 if relu else 0, L.ptr(mask), L.ptr(R), L.ptr(rrow), float(alpha), float(beta),
                                      L.ptr(R2), L.ptr(out), L.ptr(orow), L.stream()), "b200mp_row_linear")
    return out


def _rows_wgrad(a, b, arow=None, brow=None):
    """sum_r a[arow[r]]^T b[brow[r]] over the E rows ([E,Ka], [E,Kb] -> [Ka,Kb]; row maps None = r) on the tcgen05 weight-gradient kernel"""
    E = a.shape[0]
    if E == 0:
        return a.new_zeros((a.shape[1], b.shape[1]))
    npart = torch.cuda.get_device_properties(a.device).multi_processor_count
    part = torch.empty((npart, a.shape[1], b.shape[1]), dtype=torch.float32, device=a.device)
    L.check(L.lib().b200mp_node_linear_wgrad_rows(L.ptr(a), L.ptr(arow), a.shape[1], a.shape[1], L.ptr(b), L.ptr(brow), b.shape[1], b.shape[1],
                                                  E, L.ptr(part), npart, L.stream()), "b200mp_node_linear_wgrad_rows")
    return part.sum(0)


def _edge_linear_fwd_wide(mode, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out_eid):
    name = ("pna_message_fwd", "gine_message_fwd", "edge_update_fwd")[mode]
    with torch.cuda.device(e.device), _track(name, 2 if mode == 2 else 1):
        out1 = _row_linear(e, W1, b1, xrow=eid, T1=T1, idx1=idx1 if T1 is not None else None, T2=T2,
                           idx2=idx2 if T2 is not None else None, relu=mode >= 1)
        out2 = None
        if mode == 2:                            # e'[out_eid[k]] = alpha e[eid[k]] + beta (h_k W2^T + b2)
            out2 = _row_linear(out1, W2, b2, R=e, rrow=eid, alpha=alpha, beta=beta, orow=out_eid)
    return out1, out2


def _message_bwd_wide(kind, seg, m, e, node_in, W, need_T1, N, g_e, accumulate, g_acc, eid):
    E, H = m.shape
    dev = m.device
    det = deterministic_backward()
    with torch.cuda.device(dev), _track("pna_message_bwd" if kind == 0 else "gine_message_bwd", 3):
        G = torch.empty((E, H), dtype=torch.float32, device=dev)
        gT1 = torch.zeros((N, H), dtype=torch.float32, device=dev) if (need_T1 and not det) else None
        L.check(L.lib().b200mp_message_grad_rows(kind, H, E, L.ptr(m), L.ptr(node_in), L.ptr(seg.seg32), L.ptr(seg.nbr32), L.ptr(G),
                                                 L.ptr(gT1), L.stream()), "b200mp_message_grad_rows")
        # g_e[eid[k]] (+)= G_k W (+ g_acc[k])
        _row_linear(G, W.t(), None, R=g_e if accumulate else None, rrow=eid, alpha=1.0, beta=1.0, R2=g_acc, out=g_e, orow=eid)
        gW = _rows_wgrad(G, e, brow=eid)
    if need_T1 and det:
        gT1 = _sum_rows_by_neighbour(seg, G, N)
    return g_e, gT1, gW, G.sum(0)


def _edge_update_bwd_wide(st, g_out, h, e, W1, W2, alpha, beta, need_P, need_Q, nP, nQ, g_eid, e_eid):
    E, H = e.shape
    dev = e.device
    det = deterministic_backward()
    with torch.cuda.device(dev), _track("edge_update_bwd", 4):
        Gh = _row_linear(g_out, W2.t(), None, xrow=g_eid, mask=h, beta=beta)               # beta (g W2) [h > 0]
        g_e = _row_linear(Gh, W1.t(), None, R=g_out, rrow=g_eid, alpha=alpha, beta=1.0)     # alpha g + Gh W1
        gW2 = _rows_wgrad(g_out, h, arow=g_eid) * beta
        gW1 = _rows_wgrad(Gh, e, brow=e_eid)
        gP = gQ = None
        if need_P and not det:
            gP = torch.zeros((nP or st.num_nodes, H), dtype=torch.float32, device=dev)
            L.check(L.lib().b200mp_scatter_add_rows(L.ptr(Gh), L.ptr(st.src32), E, H, L.ptr(gP), L.stream()), "b200mp_scatter_add_rows")
    if need_Q:                                   # destination-sorted: a segmented sum, no atomics
        gQ = _segment_sum_rows(st.fwd, Gh)[: (nQ or st.num_nodes)]
    if need_P and det:
        gP = _sum_rows_by_neighbour(st.fwd, Gh, nP or st.num_nodes)
    return g_e, gP, gQ, gW1, Gh.sum(0), gW2, g_out.sum(0) * beta        # (a column sum does not care about the row map)


# ----------------------------------------------------------------------------- raw launches
def _edge_linear_fwd(mode, e, eid, W1, b1, T1, idx1, T2, idx2, W2=None, b2=None, alpha=0.0, beta=0.0, out_eid=None):
    E, H = e.shape
    if e.dtype == torch.float32 and E > 0 and wide_tc_available(H):
        return _edge_linear_fwd_wide(mode, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out_eid)
    out1 = torch.empty((E, H), dtype=e.dtype, device=e.device)
    out2 = torch.empty((E, H), dtype=e.dtype, device=e.device) if mode == 2 else None
    if e.dtype == torch.bfloat16:
        _bf16_ok(e)
        with torch.cuda.device(e.device), _track(("pna_message_fwd", "gine_message_fwd", "edge_update_fwd")[mode]):
            L.check(L.lib().b200mp_edge_linear_fwd_bf16(mode, E, L.ptr(e), L.ptr(eid), L.ptr(W1), L.ptr(b1), L.ptr(T1), L.ptr(idx1),
                                                        L.ptr(T2), L.ptr(idx2), L.ptr(W2), L.ptr(b2), alpha, beta, L.ptr(out1),
                                                        L.ptr(out2), L.ptr(out_eid), L.stream()), "b200mp_edge_linear_fwd_bf16")
        return out1, out2
    with torch.cuda.device(e.device), _track(("pna_message_fwd", "gine_message_fwd", "edge_update_fwd")[mode]):
        L.check(L.lib().b200mp_edge_linear_fwd(mode, H, E, L.ptr(e), L.ptr(eid), L.ptr(W1), L.ptr(b1),
                                               L.ptr(T1), L.ptr(idx1), L.ptr(T2), L.ptr(idx2),
                                               L.ptr(W2), L.ptr(b2), alpha, beta, L.ptr(out1), L.ptr(out2), L.ptr(out_eid),
                                               L.stream()),
                "b200mp_edge_linear_fwd")
    return out1, out2


def _segment_reduce(kind, seg: SegmentIndex, m, H):
    N, dev = seg.num_nodes, m.device
    agg = torch.empty((N, 4 * H if kind == 0 else H), dtype=torch.float32, device=dev)
    tie = torch.empty((N, 2 * H), dtype=torch.float32, device=dev) if kind == 0 else None
    nh, hnode, hcptr, nc, chv, cbeg = seg.heavy()
    partial = torch.empty((nc, 6 * H), dtype=torch.float32, device=dev) if nh else None
    if m.dtype == torch.bfloat16:
        _bf16_ok(m)
        with torch.cuda.device(dev), _track("segment_reduce_pna" if kind == 0 else "segment_reduce_sum", 3 if nh else 1):
            L.check(L.lib().b200mp_segment_reduce_fwd_bf16(kind, N, L.ptr(seg.ptr), L.ptr(m), L.ptr(agg), L.ptr(tie), LIGHT_MAX, CHUNK,
                                                           nh, L.ptr(hnode), L.ptr(hcptr), nc, L.ptr(chv), L.ptr(cbeg), L.ptr(partial),
                                                           L.stream()), "b200mp_segment_reduce_fwd_bf16")
        return agg, tie
    with torch.cuda.device(dev), _track("segment_reduce_pna" if kind == 0 else "segment_reduce_sum", 3 if nh else 1):
        L.check(L.lib().b200mp_segment_reduce_fwd(kind, H, N, L.ptr(seg.ptr), L.ptr(m), L.ptr(agg), L.ptr(tie),
                                                  LIGHT_MAX, CHUNK, nh, L.ptr(hnode), L.ptr(hcptr), nc, L.ptr(chv),
                                                  L.ptr(cbeg), L.ptr(partial), L.stream()),
                "b200mp_segment_reduce_fwd")
    return agg, tie


def fused_reduce_available(H: int) -> bool:
    """The fused message + segmented-reduction kernel (csrc/msg_reduce_umma.cu) covers H = 64 on the tensor-core path.  OPT-IN
    (B200MP_FUSED_REDUCE=1): measured on B200 it is slower than the two kernels it replaces (hi_small: 3.05 ms against
    1.03 + 0.76 ms) -- with one 8-warp CTA per SM the in-tile reduction is serial latency that nothing hides."""
    import os
    return H == 64 and row_maps_available() and os.environ.get("B200MP_FUSED_REDUCE", "0") == "1"


def _message_fwd_reduce(kind, seg: SegmentIndex, e, eid, W, b, T1, T2):
    """messages + their aggregation over the segments of ``seg``: kind 0 (PNA) m = e W^T + b + T1[nbr] + T2[seg] ->
    (m, agg [N,4H] = mean|min|max|std, tie [N,2H]); kind 1 (GINE) m = relu(e W^T + b + T1[nbr]) -> (m, agg [N,H], None)."""
    E, H = e.shape
    if not fused_reduce_available(H) or e.dtype != torch.float32:
        m, _ = _edge_linear_fwd(kind, e, eid, W, b, T1, seg.nbr32, T2, seg.seg32 if T2 is not None else None)
        agg, tie = _segment_reduce(kind, seg, m, H)
        return m, agg, tie
    N, dev = seg.num_nodes, e.device
    m = torch.empty((E, H), dtype=torch.float32, device=dev)
    agg = torch.empty((N, 4 * H if kind == 0 else H), dtype=torch.float32, device=dev)
    tie = torch.empty((N, 2 * H), dtype=torch.float32, device=dev) if kind == 0 else None
    nf, ni = L.C.c_size_t(0), L.C.c_size_t(0)
    L.check(L.lib().b200mp_message_reduce_workspace(kind, E, L.C.byref(nf), L.C.byref(ni)), "b200mp_message_reduce_workspace")
    part = torch.empty(max(nf.value, 1), dtype=torch.float32, device=dev)
    meta = torch.empty(max(ni.value, 4), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev), _track("pna_message_fwd_reduce" if kind == 0 else "gine_message_fwd_reduce", 3):
        L.check(L.lib().b200mp_message_fwd_reduce(kind, H, E, N, L.ptr(seg.ptr), L.ptr(e), L.ptr(eid), L.ptr(W), L.ptr(b), L.ptr(T1),
                                                  L.ptr(seg.nbr32), L.ptr(T2), L.ptr(seg.seg32), L.ptr(m), L.ptr(agg), L.ptr(tie),
                                                  L.ptr(part), L.ptr(meta), L.stream()), "b200mp_message_fwd_reduce")
    return m, agg, tie


class EdgeGradSink:
    """Lets the two consumers of one edge tensor (the conv's message and the edge update of the
    same layer) produce ONE gradient buffer: the edge update's backward, which autograd always
    runs first (its node inputs depend on the conv output), deposits its g_e here instead of
    returning it, and the message backward accumulates into that buffer inside the kernel --
    autograd would otherwise add the two [E,H] tensors in a separate pass over HBM.  Armed only
    by the backbones' layer loop, where that structure is guaranteed."""
    __slots__ = ("buf",)

    def __init__(self):
        self.buf = None

    def take(self):
        b, self.buf = self.buf, None
        return b


_DETERMINISTIC = [None]


def set_deterministic(flag: Optional[bool]):
    """Atomic-free backward for the H = 64 kernels: per-edge gradient rows are written once and summed per node over the CSC / CSR
    orders instead of being scattered with vector REDs -- bit-identical gradients from run to run, at the price of ~4 extra [E,H]
    passes per kernel.  ``None``: follow the environment variable B200MP_DETERMINISTIC (default off)."""
    _DETERMINISTIC[0] = flag


def deterministic_backward() -> bool:
    if _DETERMINISTIC[0] is not None:
        return bool(_DETERMINISTIC[0])
    import os
    return os.environ.get("B200MP_DETERMINISTIC", "0") == "1"


def _sum_rows_by_neighbour(seg: SegmentIndex, rows, num_rows):
    """rows [E,H] in ``seg`` order -> [num_rows, H]: sum of the rows of every neighbour id, in a fixed order (no atomics)"""
    byn = seg.by_neighbour(num_rows)
    return _segment_sum_rows(byn, _permute(rows, byn.perm, False))


def _message_bwd_rows(kind, seg, m, e, node_in, W, g_e, accumulate, g_acc, eid, npart):
    E, H = m.shape
    dev = m.device
    G = torch.empty((E, H), dtype=torch.float32, device=dev)
    gW = torch.empty((npart, H, H), dtype=torch.float32, device=dev)
    gb = torch.empty((npart, H), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _track("pna_message_bwd" if kind == 0 else "gine_message_bwd"):
        L.check(L.lib().b200mp_message_bwd_rows(kind, E, L.ptr(m), L.ptr(e), L.ptr(eid), L.ptr(node_in), L.ptr(seg.seg32), L.ptr(W),
                                                L.ptr(g_e), accumulate, L.ptr(G), L.ptr(gW), L.ptr(gb), npart, L.ptr(g_acc), L.stream()),
                "b200mp_message_bwd_rows")
    return G, gW, gb


def _message_bwd(kind, seg: SegmentIndex, m, e, node_in, W, need_T1, N, g_e=None, eid=None):
    """g_e: optional [E,H] gradient (segment order) the result is added to.  eid: storage row of e
    (and of the returned gradient) for segment position k (default seg.eid32)."""
    E, H = m.shape
    dev = m.device
    eid = seg.eid32 if eid is None else eid
    accumulate, g_acc = 0, None
    if g_e is not None and eid is not None:      # add in segment order, write in storage order
        g_acc, g_e = g_e, None
    elif g_e is not None:
        accumulate = 1
    if g_e is None:
        g_e = torch.empty((E, H), dtype=m.dtype, device=dev)
    if m.dtype == torch.float32 and E > 0 and wide_tc_available(H):
        return _message_bwd_wide(kind, seg, m, e, node_in, W, need_T1, N, g_e, accumulate, g_acc, eid)
    npart = L.n_part(E, H)
    if m.dtype == torch.bfloat16:
        _bf16_ok(m)
        gT1 = torch.zeros((N, H), dtype=torch.float32, device=dev) if need_T1 else None
        gW = torch.empty((npart, H, H), dtype=torch.float32, device=dev)
        gb = torch.empty((npart, H), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev), _track("pna_message_bwd" if kind == 0 else "gine_message_bwd"):
            L.check(L.lib().b200mp_message_bwd_bf16(kind, E, L.ptr(m), L.ptr(e), L.ptr(eid), L.ptr(node_in), L.ptr(seg.seg32),
                                                    L.ptr(seg.nbr32), L.ptr(W), L.ptr(g_e), accumulate, L.ptr(gT1), L.ptr(gW), L.ptr(gb),
                                                    npart, L.ptr(g_acc), L.stream()), "b200mp_message_bwd_bf16")
        return g_e, gT1, gW.sum(0), gb.sum(0)
    if need_T1 and E > 0 and H == 64 and row_maps_available() and deterministic_backward():
        G, gW, gb = _message_bwd_rows(kind, seg, m, e, node_in, W, g_e, accumulate, g_acc, eid, npart)
        return g_e, _sum_rows_by_neighbour(seg, G, N), gW.sum(0), gb.sum(0)
    gT1 = torch.zeros((N, H), dtype=torch.float32, device=dev) if need_T1 else None
    gW = torch.empty((npart, H, H), dtype=torch.float32, device=dev)
    gb = torch.empty((npart, H), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _track("pna_message_bwd" if kind == 0 else "gine_message_bwd"):
        L.check(L.lib().b200mp_message_bwd(kind, H, E, L.ptr(m), L.ptr(e), L.ptr(eid), L.ptr(node_in),
                                           L.ptr(seg.seg32), L.ptr(seg.nbr32), L.ptr(W), L.ptr(g_e), accumulate, L.ptr(gT1),
                                           L.ptr(gW), L.ptr(gb), npart, L.ptr(g_acc), L.stream()),
                "b200mp_message_bwd")
    return g_e, gT1, gW.sum(0), gb.sum(0)


def _edge_update_bwd(st: EdgeStructure, g_out, h, e, W1, W2, alpha, beta, need_P, need_Q, nP=None, nQ=None,
                     g_eid=None, e_eid=None):
    E, H = e.shape
    dev = e.device
    if e.dtype == torch.float32 and E > 0 and wide_tc_available(H):
        return _edge_update_bwd_wide(st, g_out, h, e, W1, W2, alpha, beta, need_P, need_Q, nP, nQ, g_eid, e_eid)
    g_e = torch.empty((E, H), dtype=e.dtype, device=dev)
    npart = L.n_part(E, H)
    gW1 = torch.empty((npart, H, H), dtype=torch.float32, device=dev)
    gW2 = torch.empty((npart, H, H), dtype=torch.float32, device=dev)
    gb1 = torch.empty((npart, H), dtype=torch.float32, device=dev)
    gb2 = torch.empty((npart, H), dtype=torch.float32, device=dev)
    if e.dtype == torch.bfloat16:
        _bf16_ok(e)
        gP = torch.zeros((nP or st.num_nodes, H), dtype=torch.float32, device=dev) if need_P else None
        gQ = torch.zeros((nQ or st.num_nodes, H), dtype=torch.float32, device=dev) if need_Q else None
        with torch.cuda.device(dev), _track("edge_update_bwd"):
            L.check(L.lib().b200mp_edge_update_bwd_bf16(E, L.ptr(g_out), L.ptr(h), L.ptr(e), L.ptr(st.src32), L.ptr(st.dst32), L.ptr(W1),
                                                        L.ptr(W2), alpha, beta, L.ptr(g_e), L.ptr(gP), L.ptr(gQ), L.ptr(gW1), L.ptr(gb1),
                                                        L.ptr(gW2), L.ptr(gb2), npart, L.ptr(g_eid), L.ptr(e_eid), L.stream()),
                    "b200mp_edge_update_bwd_bf16")
        return g_e, gP, gQ, gW1.sum(0), gb1.sum(0), gW2.sum(0), gb2.sum(0)
    if (need_P or need_Q) and E > 0 and H == 64 and row_maps_available() and deterministic_backward():
        Gh = torch.empty((E, H), dtype=torch.float32, device=dev)
        with torch.cuda.device(dev), _track("edge_update_bwd"):
            L.check(L.lib().b200mp_edge_update_bwd_rows(E, L.ptr(g_out), L.ptr(h), L.ptr(e), L.ptr(W1), L.ptr(W2), alpha, beta, L.ptr(g_e),
                                                        L.ptr(Gh), L.ptr(gW1), L.ptr(gb1), L.ptr(gW2), L.ptr(gb2), npart, L.ptr(g_eid),
                                                        L.ptr(e_eid), L.ptr(st.src32), L.ptr(st.dst32), L.stream()),
                    "b200mp_edge_update_bwd_rows")
        gQ = _segment_sum_rows(st.fwd, Gh)[: (nQ or st.num_nodes)] if need_Q else None
        gP = _sum_rows_by_neighbour(st.fwd, Gh, nP or st.num_nodes) if need_P else None
        return g_e, gP, gQ, gW1.sum(0), gb1.sum(0), gW2.sum(0), gb2.sum(0)
    gP = torch.zeros((nP or st.num_nodes, H), dtype=torch.float32, device=dev) if need_P else None
    gQ = torch.zeros((nQ or st.num_nodes, H), dtype=torch.float32, device=dev) if need_Q else None
    with torch.cuda.device(dev), _track("edge_update_bwd"):
        L.check(L.lib().b200mp_edge_update_bwd(H, E, L.ptr(g_out), L.ptr(h), L.ptr(e), L.ptr(st.src32),
                                               L.ptr(st.dst32), L.ptr(W1), L.ptr(W2), alpha, beta,
                                               L.ptr(g_e), L.ptr(gP), L.ptr(gQ), L.ptr(gW1), L.ptr(gb1),
                                               L.ptr(gW2), L.ptr(gb2), npart, L.ptr(g_eid), L.ptr(e_eid), L.stream()),
                "b200mp_edge_update_bwd")
    return g_e, gP, gQ, gW1.sum(0), gb1.sum(0), gW2.sum(0), gb2.sum(0)


_SCALER_CODE = {"identity": 0, "amplification": 1, "attenuation": 2, "linear": 3, "inverse_linear": 4}


def _pna_scale_fwd(ptr, base, Z, codes, avg_lin, avg_log):
    N, F = base.shape
    out = torch.empty_like(base)
    arr = (L.C.c_int * len(codes))(*codes)
    with torch.cuda.device(base.device), _track("pna_scale_fwd"):
        L.check(L.lib().b200mp_pna_scale_fwd(F, N, L.ptr(ptr), L.ptr(base), L.ptr(Z), len(codes), arr, L.ptr(avg_lin),
                                             L.ptr(avg_log), L.ptr(out), L.stream()), "b200mp_pna_scale_fwd")
    return out


def _pna_scale_bwd(ptr, g, codes, avg_lin, avg_log):
    N, F = g.shape
    gZ = torch.empty((N, len(codes) * F), dtype=torch.float32, device=g.device)
    arr = (L.C.c_int * len(codes))(*codes)
    with torch.cuda.device(g.device), _track("pna_scale_bwd"):
        L.check(L.lib().b200mp_pna_scale_bwd(F, N, L.ptr(ptr), L.ptr(g), len(codes), arr, L.ptr(avg_lin), L.ptr(avg_log),
                                             L.ptr(gZ), L.stream()), "b200mp_pna_scale_bwd")
    return gZ


def _pna_coef(ptr, g_agg, agg, tie, need_gA):
    N, H = agg.shape[0], agg.shape[1] // 4
    coef = torch.empty((N, 7 * H), dtype=torch.float32, device=agg.device)
    gA = torch.empty((N, H), dtype=torch.float32, device=agg.device) if need_gA else None
    with torch.cuda.device(agg.device), _track("pna_coef"):
        L.check(L.lib().b200mp_pna_coef(H, N, L.ptr(ptr), L.ptr(g_agg), L.ptr(agg), L.ptr(tie), L.ptr(coef), L.ptr(gA),
                                        L.stream()), "b200mp_pna_coef")
    return coef, gA


class _PNAScale(torch.autograd.Function):
    """out = base + sum_s f_s(deg) Z_s  (degree scalers commuted past post_nn's Linear)."""

    @staticmethod
    def forward(ctx, base, Z, ptr, codes, avg_lin, avg_log):
        base, Z = _f32c(base), _f32c(Z)
        L.require_cuda(base, Z, ptr)
        avg_lin = avg_lin.detach().to(torch.float32).reshape(1).contiguous()
        avg_log = avg_log.detach().to(torch.float32).reshape(1).contiguous()
        ctx.codes = tuple(codes)
        ctx.save_for_backward(ptr, avg_lin, avg_log)
        return _pna_scale_fwd(ptr, base, Z, ctx.codes, avg_lin, avg_log)

    @staticmethod
    def backward(ctx, g):
        ptr, avg_lin, avg_log = ctx.saved_tensors
        g = _f32c(g)
        gZ = _pna_scale_bwd(ptr, g, ctx.codes, avg_lin, avg_log) if ctx.needs_input_grad[1] else None
        return (g if ctx.needs_input_grad[0] else None), gZ, None, None, None, None


def pna_scale(base, Z, ptr, scalers, avg_lin, avg_log):
    return _PNAScale.apply(base, Z, ptr, [_SCALER_CODE[s] for s in scalers], avg_lin, avg_log)


# ----------------------------------------------------------------------------- PNA aggregation
class _PNAAggregate(torch.autograd.Function):
    """agg4[N,4H] = [mean|min|max|std] over incoming m_k = A[seg_k] + B[nbr_k] + e_k W^T + b."""

    @staticmethod
    def forward(ctx, A, B, e, W, b, seg: SegmentIndex, sink=None, e_eid=None):
        A, B, W, b = map(_f32c, (A, B, W, b))
        e = _edge_c(e)
        L.require_cuda(A, B, e, W, b)
        H = _H(e)
        m, agg, tie = _message_fwd_reduce(0, seg, e, seg.eid32 if e_eid is None else e_eid, W, b, B, A)
        ctx.seg, ctx.nB, ctx.sink, ctx.e_eid = seg, B.shape[0], sink, e_eid
        ctx.save_for_backward(m, e, W, agg, tie)
        ctx.mark_non_differentiable(tie)
        return agg, tie

    @staticmethod
    def backward(ctx, g_agg, _g_tie):
        m, e, W, agg, tie = ctx.saved_tensors
        seg: SegmentIndex = ctx.seg
        g_agg = _f32c(g_agg)
        coef, gA = _pna_coef(seg.ptr, g_agg, agg, tie, ctx.needs_input_grad[0])
        need_B = ctx.needs_input_grad[1]
        acc = ctx.sink.take() if (ctx.sink is not None and seg.eid32 is None) else None
        g_e, gB, gW, gb = _message_bwd(0, seg, m, e, coef, W, need_B, ctx.nB, acc, ctx.e_eid)
        return gA, gB, g_e, gW, gb, None, None, None


def pna_aggregate(A, B, e, W, b, seg: SegmentIndex, sink=None, e_eid=None):
    """Returns (agg4 [N,4H] in the order mean,min,max,std, tie counts [N,2H]).  e_eid: int32 map
    segment position -> row of e when e is NOT stored in segment order (first layer: the caller's
    edge order); the gradient of e comes back in the same order."""
    return _PNAAggregate.apply(A, B, e, W, b, seg, sink, e_eid)


# ----------------------------------------------------------------------------- GINE aggregation
class _GINEAggregate(torch.autograd.Function):
    """agg[N,H] = sum over incoming relu(x[nbr_k] + e_k W^T + b)."""

    @staticmethod
    def forward(ctx, x, e, W, b, seg: SegmentIndex, sink=None, e_eid=None):
        x, W, b = map(_f32c, (x, W, b))
        e = _edge_c(e)
        L.require_cuda(x, e, W, b)
        H = _H(e)
        m, agg, _ = _message_fwd_reduce(1, seg, e, seg.eid32 if e_eid is None else e_eid, W, b, x, None)
        ctx.seg, ctx.nx, ctx.sink, ctx.e_eid = seg, x.shape[0], sink, e_eid
        ctx.save_for_backward(m, e, W)
        return agg

    @staticmethod
    def backward(ctx, g_agg):
        m, e, W = ctx.saved_tensors
        seg: SegmentIndex = ctx.seg
        acc = ctx.sink.take() if (ctx.sink is not None and seg.eid32 is None) else None
        g_e, gx, gW, gb = _message_bwd(1, seg, m, e, _f32c(g_agg), W, ctx.needs_input_grad[0], ctx.nx, acc, ctx.e_eid)
        return gx, g_e, gW, gb, None, None, None


def gine_aggregate(x, e, W, b, seg: SegmentIndex, sink=None, e_eid=None):
    return _GINEAggregate.apply(x, e, W, b, seg, sink, e_eid)


# ----------------------------------------------------------------------------- edge update
class _EdgeUpdate(torch.autograd.Function):
    """e' = alpha e + beta (relu(P[src] + Q[dst] + e W1^T + b1) W2^T + b2), e-storage order."""

    @staticmethod
    def forward(ctx, P, Q, e, W1, b1, W2, b2, st: EdgeStructure, alpha, beta, sink=None, e_eid=None, out_eid=None):
        P, Q, W1, b1, W2, b2 = map(_f32c, (P, Q, W1, b1, W2, b2))
        e = _edge_c(e)
        L.require_cuda(P, Q, e, W1, b1, W2, b2)
        _H(e)
        h, out = _edge_linear_fwd(2, e, e_eid, W1, b1, P, st.src32, Q, st.dst32, W2, b2, float(alpha), float(beta), out_eid)
        ctx.st, ctx.alpha, ctx.beta, ctx.nP, ctx.nQ, ctx.sink = st, float(alpha), float(beta), P.shape[0], Q.shape[0], sink
        ctx.e_eid, ctx.out_eid = e_eid, out_eid
        ctx.save_for_backward(h, e, W1, W2)
        return out

    @staticmethod
    def backward(ctx, g_out):
        h, e, W1, W2 = ctx.saved_tensors
        g_e, gP, gQ, gW1, gb1, gW2, gb2 = _edge_update_bwd(ctx.st, _edge_c(g_out).to(e.dtype), h, e, W1, W2, ctx.alpha, ctx.beta,
                                                           ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.nP, ctx.nQ,
                                                           ctx.out_eid, ctx.e_eid)
        if ctx.sink is not None and ctx.needs_input_grad[2]:
            ctx.sink.buf = g_e            # the same layer's message backward accumulates into it and returns it
            g_e = None
        elif ctx.e_eid is not None and ctx.needs_input_grad[2]:
            g_e = scatter_rows_i32(g_e, ctx.e_eid)     # g_e is in segment order, e was given in storage order
        return gP, gQ, g_e, gW1, gb1, gW2, gb2, None, None, None, None, None, None


def edge_update(P, Q, e, W1, b1, W2, b2, st: EdgeStructure, alpha: float, beta: float, sink=None, e_eid=None, out_eid=None):
    """e_eid / out_eid: int32 maps segment position -> row, for an input e / an output e' that
    live in the caller's edge order instead of segment order (first / last layer)."""
    return _EdgeUpdate.apply(P, Q, e, W1, b1, W2, b2, st, alpha, beta, sink, e_eid, out_eid)


# ----------------------------------------------------------------------------- column attention
def _col_attention_fwd(q, k, v, nhead, p, seed):
    R, S, Cc = q.shape
    out = torch.empty_like(q)
    with torch.cuda.device(q.device), _track("col_attention_fwd"):
        L.check(L.lib().b200mp_col_attention_fwd(L.ptr(q), L.ptr(k), L.ptr(v), R, S, Cc, nhead, float(p), C_u64(seed), L.ptr(out),
                                                 L.stream()), "b200mp_col_attention_fwd")
    return out


def _col_attention_bwd(q, k, v, g, nhead, p, seed):
    R, S, Cc = q.shape
    gq, gk, gv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
    with torch.cuda.device(q.device), _track("col_attention_bwd"):
        L.check(L.lib().b200mp_col_attention_bwd(L.ptr(q), L.ptr(k), L.ptr(v), L.ptr(g), R, S, Cc, nhead, float(p), C_u64(seed),
                                                 L.ptr(gq), L.ptr(gk), L.ptr(gv), L.stream()), "b200mp_col_attention_bwd")
    return gq, gk, gv


def C_u64(x):
    return L.C.c_uint64(int(x) & (2 ** 64 - 1))


def _next_seed(salt: int) -> int:
    """Seed of one dropout mask, drawn from torch's default (CPU) generator: ``torch.manual_seed`` reproduces the masks,
    ``get_rng_state`` / ``set_rng_state`` and activation checkpointing's RNG preservation (torch.utils.checkpoint,
    row_chunked below) replay them, and no device synchronisation is involved.  ``salt`` separates the three mask users."""
    return (int(torch.randint(0, 2 ** 62, (1,)).item()) * 2 + 1) ^ salt



class _ColAttention(torch.autograd.Function):
    """softmax(q k^T / sqrt(d)) v per (row, head) over [R, S, C] column tokens, with the attention
    probability dropout of nn.MultiheadAttention (mask regenerated from the seed in backward)."""

    @staticmethod
    def forward(ctx, q, k, v, nhead, p):
        q, k, v = map(_f32c, (q, k, v))
        L.require_cuda(q, k, v)
        seed = 0
        if p > 0.0:
            seed = _next_seed(0x9E3779B97F4A7C15)
        ctx.nhead, ctx.p, ctx.seed = int(nhead), float(p), seed
        ctx.save_for_backward(q, k, v)
        return _col_attention_fwd(q, k, v, ctx.nhead, ctx.p, seed)

    @staticmethod
    def backward(ctx, g):
        q, k, v = ctx.saved_tensors
        gq, gk, gv = _col_attention_bwd(q, k, v, _f32c(g), ctx.nhead, ctx.p, ctx.seed)
        return gq, gk, gv, None, None


def col_attention(q, k, v, nhead: int, p: float = 0.0):
    return _ColAttention.apply(q, k, v, nhead, p)


def transformer_encoder_layer(layer: torch.nn.TransformerEncoderLayer, x: torch.Tensor) -> torch.Tensor:
    """forward of a post-norm, batch_first nn.TransformerEncoderLayer (the module stays the parameter
    container: same state_dict keys) over column tokens x [R, S, C]: the four projections run as node
    linears (tcgen05 kernels for C = 64, R S >= 4096), the attention core on col_attention.cu; residuals,
    LayerNorms with their dropouts on row_norm.cu (add_dropout_layernorm); the feed-forward's inner
    activation + dropout is relu_dropout (one pass; other activations: torch ops)."""
    if layer.norm_first:
        raise ValueError("b200mp column attention implements the post-norm layer the reference builds")
    R, S, Cc = x.shape
    mha = layer.self_attn
    if S > 8 or Cc % mha.num_heads or (Cc // mha.num_heads) not in (1, 2, 4, 8, 16):
        # shapes the attention kernels do not cover (more than 8 tokens per row, head widths other than 1..16): torch's own
        # layer on the device, as add_dropout_layernorm / bn_relu_residual do for their uncovered shapes
        return layer(x)
    x2 = x.reshape(R * S, Cc)
    Wq, Wk, Wv = mha.in_proj_weight.chunk(3)
    bq, bk, bv = mha.in_proj_bias.chunk(3)
    q, k, v = linear_group(x2, (Wq, Wk, Wv), (bq, bk, bv))
    att = col_attention(q.view(R, S, Cc), k.view(R, S, Cc), v.view(R, S, Cc), mha.num_heads,
                        float(mha.dropout) if layer.training else 0.0)
    sa = linear_group(att.reshape(R * S, Cc), (mha.out_proj.weight,), (mha.out_proj.bias,))[0]
    train = layer.training
    x2 = add_dropout_layernorm(x2, sa, float(layer.dropout1.p) if train else 0.0, layer.norm1)
    hmid = linear_group(x2, (layer.linear1.weight,), (layer.linear1.bias,))[0]
    if layer.activation is torch.nn.functional.relu or isinstance(layer.activation, torch.nn.ReLU):
        hmid = relu_dropout(hmid, float(layer.dropout.p) if train else 0.0)
    else:
        hmid = layer.dropout(layer.activation(hmid))
    ff = linear_group(hmid, (layer.linear2.weight,), (layer.linear2.bias,))[0]
    x2 = add_dropout_layernorm(x2, ff, float(layer.dropout2.p) if train else 0.0, layer.norm2)
    return x2.view(R, S, Cc)


# ----------------------------------------------------------------------------- column (stype-wise) encoders
class EncColumn(NamedTuple):
    """One output token of the stype-wise encoder (struct b200mp_column): ``feat`` is the stype's whole feature
    tensor, this column's value(s) of row r start at element r * stride + offset."""
    kind: int
    feat: torch.Tensor
    offset: int
    stride: int
    weight: torch.Tensor                  # numerical [C] | categorical [ncat + 1, C] | timestamp [7, 8, C]
    bias: Optional[torch.Tensor]          # [C] | None (categorical)
    mean: float = 0.0
    std: float = 1.0
    min_year: float = 0.0
    ncat: int = 0


def _enc_struct(cols: Sequence[EncColumn]):
    arr = (L.Column * max(len(cols), 1))()
    for i, c in enumerate(cols):
        want = torch.float32 if c.kind == L.COL_NUMERICAL else torch.int64
        if c.feat.dtype != want or not c.feat.is_contiguous():
            raise ValueError(f"column {i}: features must be contiguous {want}, got {c.feat.dtype}")
        L.require_cuda(c.feat, c.weight, c.bias)
        arr[i] = L.Column(c.kind, c.offset, c.stride, c.feat.data_ptr(), c.weight.data_ptr(),
                          0 if c.bias is None else c.bias.data_ptr(), c.mean, c.std, c.min_year, c.ncat)
    return arr


def _col_encode_fwd(cols: Sequence[EncColumn], cls: Optional[torch.Tensor], rows: int, Cc: int) -> torch.Tensor:
    S = len(cols) + (cls is not None)
    dev = cols[0].weight.device if cols else cls.device
    out = torch.empty(rows, S, Cc, device=dev, dtype=torch.float32)
    with torch.cuda.device(dev), _track("col_encode_fwd"):
        L.check(L.lib().b200mp_col_encode_fwd(_enc_struct(cols), len(cols), L.ptr(cls), rows, Cc, L.ptr(out), L.stream()),
                "b200mp_col_encode_fwd")
    return out


def _col_encode_bwd(cols: Sequence[EncColumn], has_cls: bool, rows: int, Cc: int, g_out: torch.Tensor,
                    g_weight: Sequence[torch.Tensor], g_bias: Sequence[Optional[torch.Tensor]], g_cls: Optional[torch.Tensor]):
    """accumulates into g_weight[i] / g_bias[i] (views with the layout of cols[i].weight / bias) and g_cls"""
    n = len(cols)
    pw = (L.C.c_void_p * max(n, 1))(*[t.data_ptr() for t in g_weight])
    pb = (L.C.c_void_p * max(n, 1))(*[None if t is None else t.data_ptr() for t in g_bias])
    with torch.cuda.device(g_out.device), _track("col_encode_bwd", n + int(has_cls)):
        L.check(L.lib().b200mp_col_encode_bwd(_enc_struct(cols), n, int(has_cls), rows, Cc, L.ptr(g_out), pw, pb, L.ptr(g_cls),
                                              L.stream()), "b200mp_col_encode_bwd")


class _ColEncode(torch.autograd.Function):
    """plan: per column (kind, feat, offset, stride, (param index, row | None) of the weight, same of the bias | None,
    mean, std, min_year, ncat); params: the parameter tensors the plan indexes (fp32 contiguous)."""

    @staticmethod
    def _columns(plan, tensors):
        pick = lambda ref, ts: None if ref is None else (ts[ref[0]] if ref[1] is None else ts[ref[0]][ref[1]])
        return [EncColumn(k, feat, off, stride, pick(wref, tensors), pick(bref, tensors), mean, std, my, ncat)
                for (k, feat, off, stride, wref, bref, mean, std, my, ncat) in plan]

    @staticmethod
    def forward(ctx, plan, rows, Cc, cls, *params):
        params = tuple(_f32c(p.detach()) for p in params)
        cls_ = None if cls is None else _f32c(cls.detach()).view(-1)
        ctx.plan, ctx.rows, ctx.Cc, ctx.has_cls = plan, int(rows), int(Cc), cls is not None
        ctx.cls_shape = None if cls is None else cls.shape
        ctx.save_for_backward(*params)
        return _col_encode_fwd(_ColEncode._columns(plan, params), cls_, ctx.rows, ctx.Cc)

    @staticmethod
    def backward(ctx, g):
        params = ctx.saved_tensors
        g = _f32c(g)
        grads = tuple(torch.zeros_like(p) for p in params)
        g_cls = torch.zeros(ctx.Cc, device=g.device, dtype=torch.float32) if ctx.has_cls else None
        cols = _ColEncode._columns(ctx.plan, params)
        gcols = _ColEncode._columns(ctx.plan, grads)
        _col_encode_bwd(cols, ctx.has_cls, ctx.rows, ctx.Cc, g, [c.weight for c in gcols], [c.bias for c in gcols], g_cls)
        return (None, None, None, None if g_cls is None else g_cls.view(ctx.cls_shape)) + grads


def col_encode(plan, rows: int, Cc: int, cls: Optional[torch.Tensor], params: Sequence[torch.Tensor]) -> torch.Tensor:
    """[rows, S, C] tokens of all columns in ``plan`` (+ CLS in slot 0) in one launch; see ``_ColEncode``"""
    return _ColEncode.apply(tuple(plan), rows, Cc, cls, *params)


# ----------------------------------------------------------------------------- residual + dropout + LayerNorm
_ROW_NORM_C = (16, 32, 64, 128)


def _row_norm_fwd(x, sub, p, seed, gamma, beta, eps):
    rows, Cc = x.shape
    y = torch.empty_like(x)
    stats = torch.empty(rows, 2, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device), _track("add_dropout_layernorm_fwd"):
        L.check(L.lib().b200mp_add_dropout_layernorm_fwd(L.ptr(x), L.ptr(sub), rows, Cc, float(p), C_u64(seed), L.ptr(gamma), L.ptr(beta),
                                                         float(eps), L.ptr(y), L.ptr(stats), L.stream()), "b200mp_add_dropout_layernorm_fwd")
    return y, stats


def _row_norm_bwd(g, x, sub, stats, p, seed, gamma):
    rows, Cc = x.shape
    gx = torch.empty_like(x)
    gsub = torch.empty_like(x) if (sub is not None and p > 0.0) else None
    gg = torch.zeros(2, Cc, device=x.device, dtype=torch.float32)
    with torch.cuda.device(x.device), _track("add_dropout_layernorm_bwd"):
        L.check(L.lib().b200mp_add_dropout_layernorm_bwd(L.ptr(g), L.ptr(x), L.ptr(sub), L.ptr(stats), rows, Cc, float(p), C_u64(seed),
                                                         L.ptr(gamma), L.ptr(gx), L.ptr(gsub), L.ptr(gg[0]), L.ptr(gg[1]), L.stream()),
                "b200mp_add_dropout_layernorm_bwd")
    return gx, gsub, gg[0], gg[1]


def _row_norm_fwd_ex(x, sub, p, seed, gamma, beta, eps, yscale=1.0, blends=(), out=None):
    """y = yscale * LayerNorm(x + dropout(sub)) + sum_i s_i b_i  (blends: up to two (tensor, scale)); ``out``: where y goes"""
    rows, Cc = x.shape
    y = torch.empty_like(x) if out is None else out
    stats = torch.empty(rows, 2, device=x.device, dtype=torch.float32)
    b = list(blends) + [(None, 0.0)] * (2 - len(blends))
    with torch.cuda.device(x.device), _track("add_dropout_layernorm_fwd"):
        L.check(L.lib().b200mp_add_dropout_layernorm_fwd_ex(L.ptr(x), L.ptr(sub), rows, Cc, float(p), C_u64(seed), L.ptr(gamma), L.ptr(beta),
                                                            float(eps), float(yscale), L.ptr(b[0][0]), float(b[0][1]), L.ptr(b[1][0]),
                                                            float(b[1][1]), L.ptr(y), L.ptr(stats), L.stream()),
                "b200mp_add_dropout_layernorm_fwd_ex")
    return y, stats


def _row_norm_bwd_ex(gterms, x, sub, stats, p, seed, gamma):
    """_row_norm_bwd with the incoming gradient given as a sum of up to three (tensor, scale) terms"""
    rows, Cc = x.shape
    gx = torch.empty_like(x)
    gsub = torch.empty_like(x) if (sub is not None and p > 0.0) else None
    gg = torch.zeros(2, Cc, device=x.device, dtype=torch.float32)
    t = list(gterms) + [(None, 0.0)] * (3 - len(gterms))
    with torch.cuda.device(x.device), _track("add_dropout_layernorm_bwd"):
        L.check(L.lib().b200mp_add_dropout_layernorm_bwd_ex(L.ptr(t[0][0]), float(t[0][1]), L.ptr(t[1][0]), float(t[1][1]), L.ptr(t[2][0]),
                                                            float(t[2][1]), L.ptr(x), L.ptr(sub), L.ptr(stats), rows, Cc, float(p),
                                                            C_u64(seed), L.ptr(gamma), L.ptr(gx), L.ptr(gsub), L.ptr(gg[0]), L.ptr(gg[1]),
                                                            L.stream()), "b200mp_add_dropout_layernorm_bwd_ex")
    return gx, gsub, gg[0], gg[1]


def _scaled_sum(terms, out=None):
    """sum_i s_i t_i over up to six (tensor, scale) terms of one shape, one pass"""
    n = len(terms)
    ref = terms[0][0]
    y = torch.empty_like(ref) if out is None else out
    ptrs = (L.C.c_void_p * n)(*[t.data_ptr() for t, _ in terms])
    scales = (L.C.c_float * n)(*[float(sc) for _, sc in terms])
    with torch.cuda.device(ref.device), _track("scaled_sum"):
        L.check(L.lib().b200mp_scaled_sum(ptrs, scales, n, ref.numel(), L.ptr(y), L.stream()), "b200mp_scaled_sum")
    return y


class _AddDropoutLayerNorm(torch.autograd.Function):
    """y = LayerNorm(x + dropout(sub)) (sub None: LayerNorm(x)) over [rows, C]; mask regenerated from the seed in backward"""

    @staticmethod
    def forward(ctx, x, sub, p, gamma, beta, eps):
        x, gamma, beta = _f32c(x), _f32c(gamma), _f32c(beta)
        sub = None if sub is None else _f32c(sub)
        L.require_cuda(x, sub, gamma, beta)
        p = float(p) if sub is not None else 0.0
        seed = 0
        if p > 0.0:
            seed = _next_seed(0x632BE59BD9B4E019)
        y, stats = _row_norm_fwd(x, sub, p, seed, gamma, beta, eps)
        ctx.p, ctx.seed, ctx.has_sub = p, seed, sub is not None
        ctx.save_for_backward(x, sub, stats, gamma)
        return y

    @staticmethod
    def backward(ctx, g):
        x, sub, stats, gamma = ctx.saved_tensors
        gx, gsub, gg, gb = _row_norm_bwd(_f32c(g), x, sub, stats, ctx.p, ctx.seed, gamma)
        if ctx.has_sub and gsub is None:
            gsub = gx
        return gx, (gsub if ctx.has_sub else None), None, gg, gb, None


def add_dropout_layernorm(x: torch.Tensor, sub: Optional[torch.Tensor], p: float, norm: torch.nn.LayerNorm) -> torch.Tensor:
    """``norm(x + dropout(sub, p))`` (``sub`` None: ``norm(x)``) over the last dimension, one kernel pass each way for
    C in {16, 32, 64, 128}; other widths (not used by the reference's models) go through torch's own ops."""
    Cc = x.shape[-1]
    if Cc not in _ROW_NORM_C or norm.weight is None or norm.bias is None or tuple(norm.normalized_shape) != (Cc,):
        z = x if sub is None else x + torch.nn.functional.dropout(sub, p, p > 0.0)
        return norm(z)
    y = _AddDropoutLayerNorm.apply(x.reshape(-1, Cc), None if sub is None else sub.reshape(-1, Cc), p, norm.weight, norm.bias,
                                   norm.eps)
    return y.view(x.shape)


# ----------------------------------------------------------------------------- BatchNorm + ReLU + residual average
def _bn_stats(c):
    N, H = c.shape
    sums = torch.zeros(2, H, device=c.device, dtype=torch.float64)
    with torch.cuda.device(c.device), _track("bn_stats"):
        L.check(L.lib().b200mp_bn_stats(L.ptr(c), N, H, L.ptr(sums), L.stream()), "b200mp_bn_stats")
    return sums


def _bn_relu_res_fwd(c, x, gamma, beta, mean_rstd):
    N, H = c.shape
    out = torch.empty_like(c)
    with torch.cuda.device(c.device), _track("bn_relu_res_fwd"):
        L.check(L.lib().b200mp_bn_relu_res_fwd(L.ptr(c), L.ptr(x), N, H, L.ptr(gamma), L.ptr(beta), L.ptr(mean_rstd), L.ptr(out),
                                               L.stream()), "b200mp_bn_relu_res_fwd")
    return out


def _bn_bwd_reduce(g, c, gamma, beta, mean_rstd):
    """-> fp64 [2, H]: (sum gr, sum gr chat) over the local rows = (g_beta, g_gamma)"""
    N, H = c.shape
    sums = torch.zeros(2, H, device=c.device, dtype=torch.float64)
    with torch.cuda.device(c.device), _track("bn_relu_res_bwd_reduce"):
        L.check(L.lib().b200mp_bn_relu_res_bwd_reduce(L.ptr(g), L.ptr(c), N, H, L.ptr(gamma), L.ptr(beta), L.ptr(mean_rstd), L.ptr(sums),
                                                      L.stream()), "b200mp_bn_relu_res_bwd_reduce")
    return sums


def _bn_bwd_apply(g, c, gamma, beta, mean_rstd, gmeans):
    """-> g_c, g_x"""
    N, H = c.shape
    gc, gx = torch.empty_like(c), torch.empty_like(c)
    with torch.cuda.device(c.device), _track("bn_relu_res_bwd_apply"):
        L.check(L.lib().b200mp_bn_relu_res_bwd_apply(L.ptr(g), L.ptr(c), N, H, L.ptr(gamma), L.ptr(beta), L.ptr(mean_rstd), L.ptr(gmeans),
                                                     L.ptr(gc), L.ptr(gx), L.stream()), "b200mp_bn_relu_res_bwd_apply")
    return gc, gx


class _BnReluRes(torch.autograd.Function):
    """group / n_total: the rows of ALL ranks of a destination-range partition form the batch (the column sums are
    all-reduced, fp64); the gamma / beta gradients returned are the LOCAL parts (summed with the other weight gradients)."""

    @staticmethod
    def forward(ctx, c, x, gamma, beta, eps, bn, group, n_total):
        c, x, gamma, beta = map(_f32c, (c, x, gamma, beta))
        L.require_cuda(c, x, gamma, beta)
        synced = n_total is not None                                   # group None = the default process group
        N = int(n_total) if synced else c.shape[0]
        sums = _bn_stats(c)
        if synced:
            torch.distributed.all_reduce(sums, group=group)
        mean = sums[0] / N
        var = (sums[1] / N - mean * mean).clamp_min_(0.0)           # biased, fp64
        mean_rstd = torch.stack([mean, torch.rsqrt(var + eps)]).float()
        if bn is not None and bn.track_running_stats and bn.running_mean is not None:
            with torch.no_grad():
                bn.num_batches_tracked += 1
                m = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
                bn.running_mean.mul_(1 - m).add_(mean.to(bn.running_mean.dtype), alpha=m)
                bn.running_var.mul_(1 - m).add_((var * (N / max(N - 1, 1))).to(bn.running_var.dtype), alpha=m)
        ctx.save_for_backward(c, gamma, beta, mean_rstd)
        ctx.group, ctx.N, ctx.synced = group, N, synced
        return _bn_relu_res_fwd(c, x, gamma, beta, mean_rstd)

    @staticmethod
    def backward(ctx, g):
        c, gamma, beta, mean_rstd = ctx.saved_tensors
        g = _f32c(g)
        sums = _bn_bwd_reduce(g, c, gamma, beta, mean_rstd)
        gg, gb = sums[1].float(), sums[0].float()
        if ctx.synced:
            torch.distributed.all_reduce(sums, group=ctx.group)
        gc, gx = _bn_bwd_apply(g, c, gamma, beta, mean_rstd, (sums / ctx.N).float())
        return gc, gx, gg, gb, None, None, None, None


def bn_relu_residual(bn: torch.nn.BatchNorm1d, c: torch.Tensor, x: torch.Tensor, group=None, n_total=None) -> torch.Tensor:
    """``(x + relu(bn(c))) / 2`` (gine.py:87, pna.py:96).  Training mode with affine parameters and H in {16, 32, 64, 128}:
    two kernel passes forward, two backward (csrc/node_norm.cu), running statistics updated as BatchNorm1d does;
    eval mode / other shapes: torch's own ops.  ``group`` / ``n_total``: synchronised statistics over the ranks of a
    destination-range partition ("needs SyncBN-style statistics")."""
    H = c.shape[-1]
    if (bn.training or not bn.track_running_stats) and bn.affine and c.dim() == 2 and c.shape == x.shape \
            and H in _ROW_NORM_C and (c.shape[0] > 1 or n_total is not None):
        return _BnReluRes.apply(c, x, bn.weight, bn.bias, bn.eps, bn, group, n_total)
    if n_total is not None:
        raise ValueError("bn_relu_residual over a partition needs training mode, affine parameters and H in {16, 32, 64, 128}")
    return (x + torch.relu(bn(c))) / 2


# ----------------------------------------------------------------------------- relu + dropout
def _relu_dropout(x_or_y, g, p, seed):
    out = torch.empty_like(x_or_y)
    with torch.cuda.device(x_or_y.device), _track("relu_dropout_bwd" if g is not None else "relu_dropout_fwd"):
        L.check(L.lib().b200mp_relu_dropout(L.ptr(x_or_y), L.ptr(g), x_or_y.numel(), float(p), C_u64(seed), L.ptr(out), L.stream()),
                "b200mp_relu_dropout")
    return out


class _ReluDropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p):
        x = _f32c(x)
        L.require_cuda(x)
        seed = 0
        if p > 0.0:
            seed = _next_seed(0xD1342543DE82EF95)
        y = _relu_dropout(x, None, p, seed)
        ctx.p, ctx.seed = float(p), seed
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        return _relu_dropout(y, _f32c(g), ctx.p, ctx.seed), None


def relu_dropout(x: torch.Tensor, p: float) -> torch.Tensor:
    """dropout(relu(x), p) in one pass each way (numel % 4 == 0; otherwise torch's ops)"""
    if x.numel() % 4:
        return torch.nn.functional.dropout(torch.relu(x), p, p > 0.0)
    return _ReluDropout.apply(x, float(p))


# ----------------------------------------------------------------------------- read-out: pair gather, ranking, pooled write-back
def _pair_gather_act(Rs, Rd, T, src, dst, E, W, relu):
    dev = (T if T is not None else Rs if Rs is not None else Rd).device
    out = torch.empty((E, W), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev), _track("pair_gather_act"):
        L.check(L.lib().b200mp_pair_gather_act(L.ptr(Rs), L.ptr(Rd), L.ptr(T), L.ptr(src), L.ptr(dst), E, W, int(relu), L.ptr(out),
                                               L.stream()), "b200mp_pair_gather_act")
    return out


def _segment_sum_rows(seg: SegmentIndex, rows_sorted):
    """sum of the rows of each segment ([E, W] in segment order -> [N, W]); W in {16, 32, 64, 128}"""
    return _segment_reduce(1, seg, rows_sorted, rows_sorted.shape[1])[0]


_SEG_W = (16, 32, 64, 128)


def _pad_width(W: int) -> int:
    for w in _SEG_W:
        if W <= w:
            return w
    raise L.B200MPError(f"read-out widths up to {_SEG_W[-1]} are supported, got {W}")


class _PairGatherAct(torch.autograd.Function):
    """act(Rs[src] + Rd[dst] + T) for [N, W] node tables and an [E, W] edge table (any may be None).  Backward without
    atomics: the masked gradient is summed per destination over the CSC order and per source over the CSR order of
    edge_index (the structure is built on first use and cached, graph.edge_structure)."""

    @staticmethod
    def forward(ctx, Rs, Rd, T, edge_index, relu):
        Rs, Rd, T = (None if t is None else _f32c(t) for t in (Rs, Rd, T))
        L.require_cuda(Rs, Rd, T, edge_index)
        ref = T if T is not None else (Rs if Rs is not None else Rd)
        if edge_index.dtype != torch.int64 or edge_index.dim() != 2 or edge_index.shape[0] != 2:
            raise TypeError("edge_index must be int64 [2, E]")
        ei = edge_index.contiguous()
        E, W = int(ei.shape[1]), int(ref.shape[1])
        if W not in _SEG_W:
            raise L.B200MPError(f"pair_gather_act needs a width in {_SEG_W} (pad the weights), got {W}")
        out = _pair_gather_act(Rs, Rd, T, ei[0], ei[1], E, W, relu)
        ctx.relu, ctx.ei = bool(relu), ei
        ctx.nS = None if Rs is None else Rs.shape[0]
        ctx.nD = None if Rd is None else Rd.shape[0]
        ctx.has_T = T is not None
        ctx.save_for_backward(out if relu else None)
        return out

    @staticmethod
    def backward(ctx, g):
        from .graph import edge_structure
        (out,) = ctx.saved_tensors
        g = _f32c(g)
        gz = _relu_dropout(out, g, 0.0, 0) if ctx.relu else g          # g where out > 0
        gRs = gRd = None
        n = max(ctx.nS or 0, ctx.nD or 0)
        if (ctx.nS is not None and ctx.needs_input_grad[0]) or (ctx.nD is not None and ctx.needs_input_grad[1]):
            st = edge_structure(ctx.ei, n, reverse=ctx.nS is not None and ctx.needs_input_grad[0])
            if ctx.nD is not None and ctx.needs_input_grad[1]:
                gRd = _segment_sum_rows(st.fwd, _permute(gz, st.fwd.perm, False))[: ctx.nD]
            if ctx.nS is not None and ctx.needs_input_grad[0]:
                gRs = _segment_sum_rows(st.rev, _permute(gz, st.rev.perm, False))[: ctx.nS]
        return gRs, gRd, (gz if ctx.has_T and ctx.needs_input_grad[2] else None), None, None


def pair_gather_act(Rs, Rd, T, edge_index, relu=True):
    """``act(Rs[edge_index[0]] + Rd[edge_index[1]] + T)`` (decoder.py:17-21, :62-71 after hoisting the first Linear)."""
    return _PairGatherAct.apply(Rs, Rd, T, edge_index, relu)


def rank_positive(pos: torch.Tensor, neg: torch.Tensor, num_neg: int) -> torch.Tensor:
    """int32 [B]: 1-based rank of every positive score among its ``num_neg`` negatives (metric.py:30-43), on the device."""
    pos = pos.detach().reshape(-1).float().contiguous()
    neg = neg.detach().reshape(-1).float().contiguous()
    L.require_cuda(pos, neg)
    B = pos.numel()
    if neg.numel() != B * num_neg:
        raise ValueError(f"expected {B} x {num_neg} negative scores, got {neg.numel()}")
    return _rank_positive(pos, neg, B, int(num_neg))


def _rank_positive(pos, neg, B, K):
    rank = torch.empty(B, dtype=torch.int32, device=pos.device)
    with torch.cuda.device(pos.device), _track("rank_positive"):
        L.check(L.lib().b200mp_rank_positive(L.ptr(pos), L.ptr(neg), B, K, L.ptr(rank), L.stream()), "b200mp_rank_positive")
    return rank


def _segment_mean_blend_fwd(x, emb, ptr, perm):
    out = torch.empty_like(x)
    with torch.cuda.device(x.device), _track("segment_mean_blend_fwd"):
        L.check(L.lib().b200mp_segment_mean_blend_fwd(L.ptr(x), L.ptr(emb), L.ptr(ptr), L.ptr(perm), x.shape[0], x.shape[1], L.ptr(out),
                                                      L.stream()), "b200mp_segment_mean_blend_fwd")
    return out


def _segment_mean_blend_bwd(g, ptr, index, R, need_x, need_emb):
    N, W = g.shape
    gx = torch.empty_like(g) if need_x else None
    gemb = torch.empty((R, W), dtype=torch.float32, device=g.device) if need_emb else None
    with torch.cuda.device(g.device), _track("segment_mean_blend_bwd"):
        L.check(L.lib().b200mp_segment_mean_blend_bwd(L.ptr(g), L.ptr(ptr), L.ptr(index), N, R, W, L.ptr(gx), L.ptr(gemb), L.stream()),
                "b200mp_segment_mean_blend_bwd")
    return gx, gemb


class _SegmentMeanBlend(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, emb, index):
        from .graph import build_segment_index
        x, emb = _f32c(x), _f32c(emb)
        L.require_cuda(x, emb, index)
        index = index.contiguous().view(-1)
        if index.numel() != emb.shape[0] or emb.shape[1] != x.shape[1] or x.shape[1] % 4:
            raise ValueError("segment_mean_blend: emb must be [len(index), W] with x [N, W], W % 4 == 0")
        seg = build_segment_index(index, index, x.shape[0])
        ctx.save_for_backward(seg.ptr, index)
        ctx.R = emb.shape[0]
        return _segment_mean_blend_fwd(x, emb, seg.ptr, seg.perm)

    @staticmethod
    def backward(ctx, g):
        ptr, index = ctx.saved_tensors
        gx, gemb = _segment_mean_blend_bwd(_f32c(g), ptr, index, ctx.R, ctx.needs_input_grad[0], ctx.needs_input_grad[1])
        return gx, gemb, None


def segment_mean_blend(x: torch.Tensor, emb: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    """rows of ``x`` named in ``index`` become ``(x[n] + mean of the emb rows with index == n) / 2``, the others are copied:
    the ``unique`` / ``index_add_`` / ``bincount`` write-back of models/fused.py:256-268 as one pass without atomics."""
    return _SegmentMeanBlend.apply(x, emb, index)


# ----------------------------------------------------------------------------- row-chunked recompute
class _RowChunked(torch.autograd.Function):
    """y = fn(x) for a ROW-LOCAL fn (row r of y depends on row r of x and on the parameters only): the column-attention
    layer over [rows, S, C] tokens.  Forward runs chunk by chunk without recording the graph; backward re-runs each chunk
    with the graph, so at most one chunk's intermediates are alive (SURVEY.md section 7 "hard parts": a [E, 6, 64] token tensor
    is 49 GB at HI-Medium and the layer keeps ~10 tensors of that size for backward).  The CPU generator state of every
    chunk is recorded and replayed, so dropout masks match between the two runs."""

    @staticmethod
    def forward(ctx, fn, chunk_rows, x, *params):
        ctx.fn, ctx.chunk_rows = fn, int(chunk_rows)
        ctx.save_for_backward(x, *params)
        ctx.rng = []
        out = None
        with torch.no_grad():
            for r0 in range(0, x.shape[0], ctx.chunk_rows):
                ctx.rng.append(torch.get_rng_state())
                yc = fn(x[r0:r0 + ctx.chunk_rows])
                if out is None:                        # chunks land in place: no concatenation pass over the whole tensor
                    out = torch.empty((x.shape[0], *yc.shape[1:]), dtype=yc.dtype, device=yc.device)
                out[r0:r0 + yc.shape[0]] = yc
        return out

    @staticmethod
    def backward(ctx, g):
        x, *params = ctx.saved_tensors
        gx = torch.empty_like(x) if ctx.needs_input_grad[2] else None
        gp = [None] * len(params)
        keep = torch.get_rng_state()
        for i, r0 in enumerate(range(0, x.shape[0], ctx.chunk_rows)):
            torch.set_rng_state(ctx.rng[i])
            with torch.enable_grad():
                xc = x[r0:r0 + ctx.chunk_rows].detach().requires_grad_(gx is not None)
                yc = ctx.fn(xc)
                need = [p for p in params if p.requires_grad]
                grads = torch.autograd.grad(yc, ([xc] if gx is not None else []) + need, g[r0:r0 + ctx.chunk_rows], allow_unused=True)
            if gx is not None:
                gx[r0:r0 + ctx.chunk_rows] = grads[0]
                grads = grads[1:]
            it = iter(grads)
            for j, p in enumerate(params):
                if p.requires_grad:
                    gj = next(it)
                    if gj is not None:
                        gp[j] = gj if gp[j] is None else gp[j] + gj
        torch.set_rng_state(keep)
        return (None, None, gx, *gp)


def row_chunked(fn, x: torch.Tensor, params, chunk_rows: int) -> torch.Tensor:
    """``fn(x)`` evaluated in chunks of ``chunk_rows`` rows with recompute in backward; ``params``: the tensors fn closes over
    that need gradients (e.g. ``list(module.parameters())``).  One chunk (or fewer rows than a chunk): plain ``fn(x)``."""
    if chunk_rows is None or chunk_rows <= 0 or x.shape[0] <= chunk_rows:
        return fn(x)
    return _RowChunked.apply(fn, chunk_rows, x, *params)
