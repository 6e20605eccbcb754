"""A stack of FT-Transformer column-attention layers evaluated as ONE autograd node.

Reference: /root/reference/src/nn/models/tabgnn.py:116-131 -- per layer ``x_tab = (x_tab + tab_norm(tab_conv(x_tab))) / 2`` with
``tab_conv`` a post-norm nn.TransformerEncoderLayer (tabgnn.py:199-209, 219), then ``(x + x_tab) / 2`` into the embeddings.

functional.transformer_encoder_layer runs the same kernels one autograd node per operator.  At [rows, 6, 64] token tensors of
several GB that leaves autograd's own glue on the critical path (measured on hi_small, 525 ms/step: 36 ms of gradient adds where a
tensor has two consumers, 27 ms of bias-gradient sums, 26 ms of residual blends, 25 ms of chunk copies).  Here forward and backward
are written out by hand over the raw kernel launches:
  * the residual blends are written by the layer's last LayerNorm (b200mp_add_dropout_layernorm_fwd_ex);
  * a gradient with several contributions is carried as a list of (tensor, scale) terms and consumed as such by the next LayerNorm
    backward (.._bwd_ex); it is written once, where it leaves the stack (b200mp_scaled_sum);
  * bias gradients are the column sums the data-gradient pass of each linear delivers on the side (b200mp_node_linear64_colsum);
  * rows are processed in chunks with recompute in backward (as functional.row_chunked), the last kernel of a chunk writing
    straight into the output / input-gradient slice.
Dropout masks are counter hashes of seeds drawn once per chunk in forward and reused by the recompute.
"""
from __future__ import annotations

from typing import List, Sequence

import torch

from . import functional as BF

#: token rows below which the stack is not worth its set-up (and the node kernels fall back to library GEMMs anyway)
MIN_TOKEN_ROWS = 4096
PARAMS_PER_LAYER = 14


def layer_params(conv: torch.nn.TransformerEncoderLayer, norm: torch.nn.LayerNorm) -> List[torch.Tensor]:
    m = conv.self_attn
    return [m.in_proj_weight, m.in_proj_bias, m.out_proj.weight, m.out_proj.bias, conv.norm1.weight, conv.norm1.bias,
            conv.linear1.weight, conv.linear1.bias, conv.linear2.weight, conv.linear2.bias, conv.norm2.weight, conv.norm2.bias,
            norm.weight, norm.bias]


def supported(convs: Sequence[torch.nn.TransformerEncoderLayer], norms: Sequence[torch.nn.LayerNorm], x: torch.Tensor) -> bool:
    """C = 64 tokens, feed-forward width 64, relu, post-norm, affine norms, <= 8 tokens per row, fp32 on the device"""
    if x.dim() != 3 or x.dtype != torch.float32 or not x.is_cuda or x.shape[0] * x.shape[1] < MIN_TOKEN_ROWS:
        return False
    R, S, C = x.shape
    if C != 64 or S > 8 or len(convs) == 0:
        return False
    for conv, norm in zip(convs, norms):
        m = conv.self_attn
        relu = conv.activation is torch.nn.functional.relu or isinstance(conv.activation, torch.nn.ReLU)
        if conv.norm_first or not relu or conv.linear1.out_features != 64 or m.in_proj_weight is None or m.in_proj_bias is None:
            return False
        if C % m.num_heads or (C // m.num_heads) not in (1, 2, 4, 8, 16) or not getattr(m, "_qkv_same_embed_dim", True):
            return False
        if any(n.weight is None or n.bias is None or tuple(n.normalized_shape) != (C,) for n in (conv.norm1, conv.norm2, norm)):
            return False
        if conv.linear1.bias is None or conv.linear2.bias is None or m.out_proj.bias is None:
            return False
    return True


def _seeds(ps):
    """one seed per dropout site of a layer (0 where p == 0): attention probabilities, dropout1, inner dropout, dropout2"""
    salts = (0x9E3779B97F4A7C15, 0x632BE59BD9B4E019, 0xD1342543DE82EF95, 0x632BE59BD9B4E019)
    return tuple(BF._next_seed(s) if p > 0.0 else 0 for p, s in zip(ps, salts))


def _chunk_forward(t0, P, cfg, seeds, out, keep):
    """t0 [rows * S, 64] token rows of one chunk -> out (= T_L(..T_1(t0)) (+ t0)); keep: the tensors backward needs, per layer"""
    S, nhead, ps, eps, add_input = cfg["S"], cfg["nhead"], cfg["ps"], cfg["eps"], cfg["add_input"]
    L = len(ps)
    R = t0.shape[0] // S
    t, saved = t0, []
    for li in range(L):
        Wqkv, bqkv, Wo, bo, g1, be1, W1, b1, W2, b2, g2, be2, g3, be3 = P[li * PARAMS_PER_LAYER:(li + 1) * PARAMS_PER_LAYER]
        pa, p1, pr, p2 = ps[li]
        sa_, s1, sr, s2 = seeds[li]
        q, k, v = BF._node_linear64([t], Wqkv, bqkv, True)
        att = BF._col_attention_fwd(q.view(R, S, 64), k.view(R, S, 64), v.view(R, S, 64), nhead[li], pa, sa_).view(R * S, 64)
        sa = BF._node_linear64([att], Wo, bo, False)
        x1, st1 = BF._row_norm_fwd(t, sa, p1, s1, g1, be1, eps[li][0])
        h = BF._node_linear64([x1], W1, b1, False)
        hm = BF._relu_dropout(h, None, pr, sr)
        del h
        ff = BF._node_linear64([hm], W2, b2, False)
        ye, st2 = BF._row_norm_fwd(x1, ff, p2, s2, g2, be2, eps[li][1])
        last = li == L - 1
        blends = [(t, 0.5)] + ([(t0, 1.0)] if (last and add_input) else [])
        y, st3 = BF._row_norm_fwd_ex(ye, None, 0.0, 0, g3, be3, eps[li][2], 0.5, blends, out if last else None)
        if keep:
            saved.append((t, q, k, v, att, sa, x1, st1, hm, ff, st2, ye, st3))
        t = y
    return t, saved


def _merge(terms):
    """(tensor, scale) terms with the scales of repeated tensors added up, zero scales dropped"""
    out = []
    for t, s in terms:
        for i, (u, su) in enumerate(out):
            if u is t:
                out[i] = (u, su + s)
                break
        else:
            out.append((t, s))
    return [(t, s) for t, s in out if s != 0.0]


def _chunk_backward(g_out, saved, P, cfg, seeds, gx_out, pgrads, need_param):
    """g_out: gradient of the chunk's output; gx_out: where the input gradient goes (None: not needed); pgrads: running parameter
    gradients (list, None = nothing yet), accumulated in place"""
    S, nhead, ps, add_input = cfg["S"], cfg["nhead"], cfg["ps"], cfg["add_input"]
    L = len(ps)
    R = g_out.shape[0] // S

    def acc(idx, val):
        if not need_param[idx] or val is None:
            return
        val = val.reshape(P[idx].shape)
        if pgrads[idx] is None:
            pgrads[idx] = val.clone()
        else:
            pgrads[idx] += val

    G = [(g_out, 1.0)]
    to_input = [(g_out, 1.0)] if add_input else []
    for li in reversed(range(L)):
        base = li * PARAMS_PER_LAYER
        Wqkv, bqkv, Wo, bo, g1, be1, W1, b1, W2, b2, g2, be2, g3, be3 = P[base:base + PARAMS_PER_LAYER]
        pa, p1, pr, p2 = ps[li]
        sa_, s1, sr, s2 = seeds[li]
        t, q, k, v, att, sa, x1, st1, hm, ff, st2, ye, st3 = saved[li]
        G = _merge(G)
        if len(G) > 3:
            G = [(BF._scaled_sum(G), 1.0)]
        half = [(g, 0.5 * s) for g, s in G]                      # y = 0.5 LN3(ye) + 0.5 t
        g_ye, _, gg3, gb3 = BF._row_norm_bwd_ex(half, ye, None, st3, 0.0, 0, g3)
        acc(base + 12, gg3), acc(base + 13, gb3)
        g_x1r, g_ff, gg2, gb2 = BF._row_norm_bwd(g_ye, x1, ff, st2, p2, s2, g2)
        acc(base + 10, gg2), acc(base + 11, gb2)
        del g_ye
        g_ff = g_x1r if g_ff is None else g_ff
        g_hm, cs = BF._node_linear64_colsum([g_ff], W2.t())      # ff = hm W2^T + b2
        acc(base + 9, cs)
        if need_param[base + 8]:
            acc(base + 8, BF._node_pair_wgrad(g_ff, None, hm)[0])
        g_h = BF._relu_dropout(hm, g_hm, pr, sr)
        del g_hm
        g_x1f, cs = BF._node_linear64_colsum([g_h], W1.t())      # h = x1 W1^T + b1
        acc(base + 7, cs)
        if need_param[base + 6]:
            acc(base + 6, BF._node_pair_wgrad(g_h, None, x1)[0])
        del g_h
        g_tr, g_sa, gg1, gb1 = BF._row_norm_bwd_ex([(g_x1r, 1.0), (g_x1f, 1.0)], t, sa, st1, p1, s1, g1)
        acc(base + 4, gg1), acc(base + 5, gb1)
        del g_x1r, g_x1f, g_ff
        g_sa = g_tr if g_sa is None else g_sa
        g_att, cs = BF._node_linear64_colsum([g_sa], Wo.t())     # sa = att Wo^T + bo
        acc(base + 3, cs)
        if need_param[base + 2]:
            acc(base + 2, BF._node_pair_wgrad(g_sa, None, att)[0])
        gq, gk, gv = BF._col_attention_bwd(q.view(R, S, 64), k.view(R, S, 64), v.view(R, S, 64), g_att.view(R, S, 64), nhead[li], pa, sa_)
        gq, gk, gv = gq.view(R * S, 64), gk.view(R * S, 64), gv.view(R * S, 64)
        g_tq, cs = BF._node_linear64_colsum([gq, gk, gv], Wqkv.t())          # [q | k | v] = t Wqkv^T + bqkv
        acc(base + 1, cs)
        if need_param[base]:
            gWq, gWk = BF._node_pair_wgrad(gq, gk, t)
            gWv = BF._node_pair_wgrad(gv, None, t)[0]
            acc(base, torch.cat([gWq, gWk, gWv], 0))
        G = [(g_tr, 1.0), (g_tq, 1.0)] + half
    if gx_out is not None:
        final = _merge(G + to_input)
        while len(final) > 6:                                    # deeper stacks: fold the tail first
            final = final[:5] + [(BF._scaled_sum(final[5:10]), 1.0)] + final[10:]
        BF._scaled_sum(final, out=gx_out)


class _ColumnStack(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cfg, *P):
        x = BF._f32c(x)
        R, S, C = x.shape
        x2 = x.view(R * S, C)
        out = torch.empty_like(x2)
        chunk = cfg["chunk_rows"] if (cfg["chunk_rows"] and cfg["chunk_rows"] > 0) else R
        Pc = [BF._f32c(p) for p in P]
        seeds = []
        for r0 in range(0, R, chunk):
            r1 = min(R, r0 + chunk)
            sd = [_seeds(ps) for ps in cfg["ps"]]
            seeds.append(sd)
            _chunk_forward(x2[r0 * S:r1 * S], Pc, cfg, sd, out[r0 * S:r1 * S], False)
        ctx.cfg, ctx.seeds, ctx.chunk = cfg, seeds, chunk
        ctx.save_for_backward(x, *P)
        return out.view(R, S, C)

    @staticmethod
    def backward(ctx, g):
        x, *P = ctx.saved_tensors
        cfg = ctx.cfg
        R, S, C = x.shape
        x2, g2 = x.view(R * S, C), BF._f32c(g).view(R * S, C)
        Pc = [BF._f32c(p) for p in P]
        need_param = [bool(n) for n in ctx.needs_input_grad[2:]]
        gx = torch.empty_like(x2) if ctx.needs_input_grad[0] else None
        pgrads = [None] * len(P)
        scratch = None
        for i, r0 in enumerate(range(0, R, ctx.chunk)):
            r1 = min(R, r0 + ctx.chunk)
            rows = (r1 - r0) * S
            if scratch is None or scratch.shape[0] < rows:
                scratch = torch.empty((rows, C), dtype=torch.float32, device=x.device)
            _, saved = _chunk_forward(x2[r0 * S:r1 * S], Pc, cfg, ctx.seeds[i], scratch[:rows], True)
            _chunk_backward(g2[r0 * S:r1 * S], saved, Pc, cfg, ctx.seeds[i], None if gx is None else gx[r0 * S:r1 * S], pgrads, need_param)
            del saved
        return (None if gx is None else gx.view(R, S, C), None, *pgrads)


def column_stack(convs, norms, x: torch.Tensor, chunk_rows: int, add_input: bool = False) -> torch.Tensor:
    """``t = x; for conv, norm: t = (t + norm(conv(t))) / 2; return t (+ x if add_input)`` over [rows, S, 64] column tokens"""
    train = convs[0].training
    ps = tuple((float(c.self_attn.dropout) if train else 0.0, float(c.dropout1.p) if train else 0.0,
                float(c.dropout.p) if train else 0.0, float(c.dropout2.p) if train else 0.0) for c in convs)
    cfg = {"S": int(x.shape[1]), "nhead": tuple(int(c.self_attn.num_heads) for c in convs), "ps": ps,
           "eps": tuple((float(c.norm1.eps), float(c.norm2.eps), float(n.eps)) for c, n in zip(convs, norms)),
           "add_input": bool(add_input), "chunk_rows": int(chunk_rows) if torch.is_grad_enabled() else 0}
    P = [p for c, n in zip(convs, norms) for p in layer_params(c, n)]
    return _ColumnStack.apply(x, cfg, *P)
