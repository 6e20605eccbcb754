"""TEST INFRASTRUCTURE ONLY: a plain-PyTorch emulation of the C-ABI kernels (include/b200mp.h),
patched in by CPU tests so that the host-side logic of the drop-in modules (weight hoisting,
coefficient algebra of the PNA backward, permutations, state_dict layout) can be checked against
the golden fixtures without a GPU.  The product never imports this file; GPU tests run the real
kernels and additionally compare them with this emulation op by op."""
from __future__ import annotations

import contextlib
import math

import torch

import b200mp
from b200mp import _lib as L
from b200mp import functional as BF
from b200mp import graph as G

STD_THR = torch.tensor(math.sqrt(1e-5), dtype=torch.float32)


def build_segment_index(key, other_end, num_nodes):
    perm = torch.argsort(key, stable=True)
    ptr = torch.zeros(num_nodes + 1, dtype=torch.int64, device=key.device)
    ptr[1:] = torch.cumsum(torch.bincount(key, minlength=num_nodes), 0)
    other = other_end[perm]
    return G.SegmentIndex(int(num_nodes), int(key.numel()), perm, ptr, other, other.to(torch.int32),
                          key[perm].to(torch.int32))


def permute(x, idx, scatter):
    out = torch.empty_like(x)
    if scatter:
        out[idx] = x
    else:
        out = x[idx]
    return out


def edge_linear_fwd(mode, e, eid, W1, b1, T1, idx1, T2, idx2, W2=None, b2=None, alpha=0.0, beta=0.0, out_eid=None):
    rows = eid.long() if eid is not None else torch.arange(e.shape[0], device=e.device)
    x = e[rows]
    z = x @ W1.t() + b1
    if T1 is not None:
        z = z + T1[idx1.long()]
    if T2 is not None:
        z = z + T2[idx2.long()]
    if mode >= 1:
        z = z.relu()
    out2 = None
    if mode == 2:
        out2 = torch.empty_like(e)
        orows = out_eid.long() if out_eid is not None else torch.arange(e.shape[0], device=e.device)
        out2[orows] = alpha * x + beta * (z @ W2.t() + b2)
    return z, out2


def segment_reduce(kind, seg, m, H):
    N = seg.num_nodes
    idx = seg.seg32.long()
    s = torch.zeros((N, H), dtype=m.dtype).index_add_(0, idx, m)
    if kind == 1:
        return s, None
    deg = (seg.ptr[1:] - seg.ptr[:-1]).to(m.dtype).unsqueeze(1)
    dn = deg.clamp(min=1)
    q = torch.zeros((N, H), dtype=m.dtype).index_add_(0, idx, m * m)
    ix = idx.unsqueeze(1).expand_as(m)
    mn = torch.zeros((N, H), dtype=m.dtype).scatter_reduce_(0, ix, m, "amin", include_self=False)
    mx = torch.zeros((N, H), dtype=m.dtype).scatter_reduce_(0, ix, m, "amax", include_self=False)
    cn = torch.zeros((N, H), dtype=m.dtype).index_add_(0, idx, (m == mn[idx]).to(m.dtype))
    cx = torch.zeros((N, H), dtype=m.dtype).index_add_(0, idx, (m == mx[idx]).to(m.dtype))
    mean = s / dn
    var = q / dn - mean * mean
    std = var.clamp(min=1e-5).sqrt()
    std = torch.where(std <= STD_THR.to(m.dtype), torch.zeros_like(std), std)
    has = deg > 0
    agg = torch.cat([mean, mn, mx, std], 1) * has
    return agg, torch.cat([cn, cx], 1) * has


def message_bwd(kind, seg, m, e, node_in, W, need_T1, N, g_e=None, eid=None):
    H = m.shape[1]
    acc = g_e
    d = seg.seg32.long()
    eid = seg.eid32 if eid is None else eid
    rows = eid.long() if eid is not None else torch.arange(m.shape[0])
    if kind == 0:
        c = node_in[d]
        c0, c1, gmn, mn, gmx, mx, mu = [c[:, i * H:(i + 1) * H] for i in range(7)]
        Gm = c0 + c1 * (m - mu) + torch.where(m == mn, gmn, torch.zeros_like(m)) + torch.where(m == mx, gmx, torch.zeros_like(m))
    else:
        Gm = torch.where(m > 0, node_in[d], torch.zeros_like(m))
    gT1 = torch.zeros((N, H), dtype=m.dtype).index_add_(0, seg.nbr32.long(), Gm) if need_T1 else None
    g_e = torch.empty_like(e)
    if acc is not None and eid is not None:  # kernel contract: add in segment order, write through the row map
        g_e[rows] = Gm @ W + acc
    elif acc is not None:                    # accumulate into the given buffer and return it
        acc += Gm @ W
        g_e = acc
    else:
        g_e[rows] = Gm @ W
    return g_e, gT1, Gm.t() @ e[rows], Gm.sum(0)


def edge_update_bwd(st, g_out, h, e, W1, W2, alpha, beta, need_P, need_Q, nP=None, nQ=None, g_eid=None, e_eid=None):
    N, H = st.num_nodes, e.shape[1]
    if g_eid is not None:
        g_out = g_out[g_eid.long()]
    if e_eid is not None:
        e = e[e_eid.long()]
    Gout = beta * g_out
    Gh = (Gout @ W2) * (h > 0).to(e.dtype)
    gP = torch.zeros((nP or N, H), dtype=e.dtype).index_add_(0, st.src32.long(), Gh) if need_P else None
    gQ = torch.zeros((nQ or N, H), dtype=e.dtype).index_add_(0, st.dst32.long(), Gh) if need_Q else None
    g_e = alpha * g_out + Gh @ W1
    return g_e, gP, gQ, Gh.t() @ e, Gh.sum(0), Gout.t() @ h, Gout.sum(0)


def split_tf32(x):
    hi = ((x.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)
    return hi, x - hi


def _scaler(code, d, avg_lin, avg_log):
    if code == 1:
        return torch.log(d + 1) / avg_log
    if code == 2:
        return avg_log / torch.log(d.clamp(min=1) + 1)
    if code == 3:
        return d / avg_lin
    if code == 4:
        return avg_lin / d.clamp(min=1)
    return torch.ones_like(d)


def pna_scale_fwd(ptr, base, Z, codes, avg_lin, avg_log):
    d = (ptr[1:] - ptr[:-1]).to(base.dtype).unsqueeze(1)
    F = base.shape[1]
    out = base.clone()
    for s, c in enumerate(codes):
        out = out + _scaler(c, d, avg_lin, avg_log) * Z[:, s * F:(s + 1) * F]
    return out


def pna_scale_bwd(ptr, g, codes, avg_lin, avg_log):
    d = (ptr[1:] - ptr[:-1]).to(g.dtype).unsqueeze(1)
    return torch.cat([_scaler(c, d, avg_lin, avg_log) * g for c in codes], 1)


def pna_coef(ptr, g_agg, agg, tie, need_gA):
    H = agg.shape[1] // 4
    deg = (ptr[1:] - ptr[:-1]).to(agg.dtype).unsqueeze(1)
    dn = deg.clamp(min=1.0)
    g_mean, g_min, g_max, g_std = [g_agg[:, i * H:(i + 1) * H] for i in range(4)]
    mean, mn, mx, std = [agg[:, i * H:(i + 1) * H] for i in range(4)]
    valid = std > 0
    c1 = torch.where(valid, g_std / (dn * torch.where(valid, std, torch.ones_like(std))), torch.zeros_like(std))
    coef = torch.cat([g_mean / dn, c1, g_min / tie[:, :H].clamp(min=1.0), mn, g_max / tie[:, H:].clamp(min=1.0), mx, mean], 1)
    gA = (g_mean + g_min + g_max) * (deg > 0).to(agg.dtype) if need_gA else None
    return coef.contiguous(), gA


def col_attention_fwd(q, k, v, nhead, p, seed):
    R, S, C = q.shape
    hd = C // nhead
    sp = lambda t: t.view(R, S, nhead, hd).transpose(1, 2)
    prob = torch.softmax(sp(q) @ sp(k).transpose(-1, -2) / hd ** 0.5, -1)
    if p > 0.0:          # a mask stream of the emulator's own (the kernel's hash is not restated): same seed, same mask
        prob = prob * _drop_mask(prob.detach(), p, seed)
    a = prob @ sp(v)
    return a.transpose(1, 2).reshape(R, S, C)


def col_attention_bwd(q, k, v, g, nhead, p, seed):
    q, k, v = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
    with torch.enable_grad():
        out = col_attention_fwd(q, k, v, nhead, p, seed)
    out.backward(g)
    return q.grad, k.grad, v.grad


def _enc_blocks(cols):
    """EncColumn list -> the oracle's per-column blocks (one column per block keeps the output order)"""
    from oracle import stype_encoders as O  # noqa: F401
    blocks = []
    for c in cols:
        if c.kind == 0:
            v = c.feat.reshape(c.feat.shape[0], -1)[:, c.offset:c.offset + 1]
            blocks.append(('numerical', v, torch.tensor([c.mean]), torch.tensor([c.std]), c.weight.view(1, -1), c.bias.view(1, -1)))
        elif c.kind == 1:
            v = c.feat.reshape(c.feat.shape[0], -1)[:, c.offset:c.offset + 1]
            blocks.append(('categorical', v, [c.weight]))
        else:
            v = c.feat.reshape(c.feat.shape[0], -1)[:, c.offset:c.offset + 7].unsqueeze(1)
            blocks.append(('timestamp', v, torch.tensor([c.min_year]), c.weight.reshape(1, 7, 8, -1), c.bias.view(1, -1)))
    return blocks


def col_encode_fwd(cols, cls, rows, Cc):
    from oracle import stype_encoders as O
    return O.stype_wise_encode(_enc_blocks(cols), cls).contiguous()


def col_encode_bwd(cols, has_cls, rows, Cc, g_out, g_weight, g_bias, g_cls):
    from oracle import stype_encoders as O
    with torch.enable_grad():
        leaves = []
        cols2 = []
        for c in cols:
            w = c.weight.detach().clone().requires_grad_(True)
            b = None if c.bias is None else c.bias.detach().clone().requires_grad_(True)
            leaves.append((w, b))
            cols2.append(c._replace(weight=w, bias=b))
        cls = torch.zeros(Cc, requires_grad=True) if has_cls else None
        out = O.stype_wise_encode(_enc_blocks(cols2), cls)
        out.backward(g_out)
    for (w, b), gw, gb, c in zip(leaves, g_weight, g_bias, cols):
        g = w.grad.clone()
        if c.kind == 1:
            g[0] = 0                       # padding row of nn.Embedding(padding_idx=0)
        gw += g.view_as(gw)
        if b is not None:
            gb += b.grad.view_as(gb)
    if has_cls:
        g_cls += cls.grad


def row_norm_fwd(x, sub, p, seed, gamma, beta, eps):
    z = x if sub is None else x + sub * _drop_mask(sub, p, seed)
    mean = z.mean(-1, keepdim=True)
    var = z.var(-1, unbiased=False, keepdim=True)
    rstd = torch.rsqrt(var + eps)
    return ((z - mean) * rstd * gamma + beta).contiguous(), torch.cat([mean, rstd], -1).contiguous()


def _drop_mask(t, p, seed):
    if p <= 0.0:
        return torch.ones_like(t)
    g = torch.Generator().manual_seed(int(seed) & (2 ** 62 - 1))
    return (torch.rand(t.shape, generator=g) >= p).to(t.dtype) / (1.0 - p)


def row_norm_bwd(g, x, sub, stats, p, seed, gamma):
    mask = None if sub is None else _drop_mask(sub, p, seed)
    z = x if sub is None else x + sub * mask
    mean, rstd = stats[:, :1], stats[:, 1:]
    zh = (z - mean) * rstd
    gh = g * gamma
    gz = rstd * (gh - gh.mean(-1, keepdim=True) - zh * (gh * zh).mean(-1, keepdim=True))
    gsub = (gz * mask).contiguous() if (sub is not None and p > 0.0) else None
    return gz.contiguous(), gsub, (g * zh).sum(0), g.sum(0)


def row_norm_fwd_ex(x, sub, p, seed, gamma, beta, eps, yscale=1.0, blends=(), out=None):
    y, stats = row_norm_fwd(x, sub, p, seed, gamma, beta, eps)
    y = yscale * y
    for b, sc in blends:
        y = y + sc * b
    if out is not None:
        out.copy_(y)
        y = out
    return y, stats


def row_norm_bwd_ex(gterms, x, sub, stats, p, seed, gamma):
    g = sum(sc * t for t, sc in gterms)
    return row_norm_bwd(g, x, sub, stats, p, seed, gamma)


def scaled_sum(terms, out=None):
    y = sum(sc * t for t, sc in terms)
    if out is not None:
        out.copy_(y)
        return out
    return y


def bn_stats(c):
    cd = c.double()
    return torch.stack([cd.sum(0), (cd * cd).sum(0)])


def bn_relu_res_fwd(c, x, gamma, beta, mean_rstd):
    y = (c - mean_rstd[0]) * (mean_rstd[1] * gamma) + beta
    return ((x + torch.relu(y)) * 0.5).contiguous()


def _bn_gr_ch(g, c, gamma, beta, mean_rstd):
    ch = (c - mean_rstd[0]) * mean_rstd[1]
    y = (c - mean_rstd[0]) * (mean_rstd[1] * gamma) + beta
    return torch.where(y > 0, 0.5 * g, torch.zeros_like(g)), ch


def bn_bwd_reduce(g, c, gamma, beta, mean_rstd):
    gr, ch = _bn_gr_ch(g, c, gamma, beta, mean_rstd)
    return torch.stack([gr.double().sum(0), (gr * ch).double().sum(0)])


def bn_bwd_apply(g, c, gamma, beta, mean_rstd, gmeans):
    gr, ch = _bn_gr_ch(g, c, gamma, beta, mean_rstd)
    gc = gamma * mean_rstd[1] * (gr - gmeans[0] - ch * gmeans[1])
    return gc.contiguous(), (0.5 * g).contiguous()


def relu_dropout_emul(x_or_y, g, p, seed):
    m = _drop_mask(x_or_y, p, seed)
    if g is None:
        return (torch.relu(x_or_y) * m).contiguous()
    return torch.where(x_or_y > 0, g * m, torch.zeros_like(g)).contiguous()


def pair_gather_act(Rs, Rd, T, src, dst, E, W, relu):
    out = torch.zeros((E, W), dtype=torch.float32) if T is None else T.clone()
    if Rs is not None:
        out = out + Rs[src]
    if Rd is not None:
        out = out + Rd[dst]
    return (torch.relu(out) if relu else out).contiguous()


def rank_positive(pos, neg, B, K):
    return (1 + (neg.view(B, K) >= pos.view(B, 1)).sum(1)).to(torch.int32)


def segment_mean_blend_fwd(x, emb, ptr, perm):
    cnt = (ptr[1:] - ptr[:-1])
    node = torch.repeat_interleave(torch.arange(x.shape[0]), cnt)
    s = torch.zeros_like(x).index_add_(0, node, emb[perm])
    has = (cnt > 0).unsqueeze(1)
    return torch.where(has, 0.5 * (x + s / cnt.clamp(min=1).unsqueeze(1).float()), x).contiguous()


def segment_mean_blend_bwd(g, ptr, index, R, need_x, need_emb):
    cnt = (ptr[1:] - ptr[:-1])
    gx = torch.where((cnt > 0).unsqueeze(1), 0.5 * g, g).contiguous() if need_x else None
    gemb = (g[index] * (0.5 / cnt[index].float()).unsqueeze(1)).contiguous() if need_emb else None
    return gx, gemb


def node_linear64(xs, Wcat, bias, ysplit):
    y = torch.cat(list(xs), dim=1) @ Wcat.t()
    if bias is not None:
        y = y + bias
    return [c.contiguous() for c in y.split(64, dim=1)] if ysplit else y.contiguous()


def node_linear64_colsum(xs, Wcat):
    x = torch.cat(list(xs), dim=1)
    return (x @ Wcat.t()).contiguous(), x.sum(0)


def node_pair_wgrad(g0, g1, x):
    return g0.t() @ x, (g1.t() @ x if g1 is not None else torch.zeros(64, x.shape[1]))


def node_linear(x, W, b):
    return torch.nn.functional.linear(x, W, b)


def node_wgrad(x, y):
    return x.t() @ y


def _with_storage(fn, edge_args, edge_outs):
    """bf16-stored edge tensors: compute in fp32, round the edge-shaped results to the storage dtype (what the kernels do)"""
    def wrapped(*args, **kw):
        args, dt = list(args), None
        for i in edge_args:
            if i < len(args) and torch.is_tensor(args[i]) and args[i].dtype == torch.bfloat16:
                dt, args[i] = torch.bfloat16, args[i].float()
        for k in ("g_e",):
            if torch.is_tensor(kw.get(k)) and kw[k].dtype == torch.bfloat16:
                dt, kw[k] = torch.bfloat16, kw[k].float()
        out = fn(*args, **kw)
        if dt is None:
            return out
        out = list(out)
        for i in edge_outs:
            if out[i] is not None:
                out[i] = out[i].to(dt)
        return tuple(out)
    return wrapped


edge_linear_fwd = _with_storage(edge_linear_fwd, (1,), (0, 1))
segment_reduce = _with_storage(segment_reduce, (2,), ())
message_bwd = _with_storage(message_bwd, (2, 3, 8), (0,))
edge_update_bwd = _with_storage(edge_update_bwd, (1, 2, 3), (0,))

_PATCHES = [
    (BF, "_bf16_ok", lambda t: None),
    (BF, "fused_reduce_available", lambda H: False),          # the fused kernel's two halves are emulated separately
    (BF, "_node_linear64", node_linear64),
    (BF, "_node_linear64_colsum", node_linear64_colsum),
    (BF, "_row_norm_fwd_ex", row_norm_fwd_ex),
    (BF, "_row_norm_bwd_ex", row_norm_bwd_ex),
    (BF, "_scaled_sum", scaled_sum),
    (BF, "_node_pair_wgrad", node_pair_wgrad),
    (BF, "_node_linear", node_linear),
    (BF, "_node_wgrad", node_wgrad),
    (BF, "_pair_gather_act", pair_gather_act),
    (BF, "_rank_positive", rank_positive),
    (BF, "_segment_mean_blend_fwd", segment_mean_blend_fwd),
    (BF, "_segment_mean_blend_bwd", segment_mean_blend_bwd),
    (BF, "_relu_dropout", relu_dropout_emul),
    (BF, "_bn_stats", bn_stats),
    (BF, "_bn_relu_res_fwd", bn_relu_res_fwd),
    (BF, "_bn_bwd_reduce", bn_bwd_reduce),
    (BF, "_bn_bwd_apply", bn_bwd_apply),
    (BF, "_row_norm_fwd", row_norm_fwd),
    (BF, "_row_norm_bwd", row_norm_bwd),
    (BF, "_col_encode_fwd", col_encode_fwd),
    (BF, "_col_encode_bwd", col_encode_bwd),
    (BF, "_col_attention_fwd", col_attention_fwd),
    (BF, "_col_attention_bwd", col_attention_bwd),
    (BF, "_pna_scale_fwd", pna_scale_fwd),
    (BF, "_pna_scale_bwd", pna_scale_bwd),
    (BF, "_pna_coef", pna_coef),
    (BF, "_split_tf32", split_tf32),
    (G, "build_segment_index", build_segment_index),
    (L, "require_cuda", lambda *a: None),
    (BF, "_permute", permute),
    (BF, "_edge_linear_fwd", edge_linear_fwd),
    (BF, "_segment_reduce", segment_reduce),
    (BF, "_message_bwd", message_bwd),
    (BF, "_edge_update_bwd", edge_update_bwd),
]


@contextlib.contextmanager
def emulated_kernels():
    saved = [(mod, name, getattr(mod, name)) for mod, name, _ in _PATCHES]
    for mod, name, fn in _PATCHES:
        setattr(mod, name, fn)
    G.clear_cache()
    try:
        yield
    finally:
        for mod, name, fn in saved:
            setattr(mod, name, fn)
        G.clear_cache()
