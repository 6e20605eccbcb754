"""-m gpu: the code path bench.py times, against the CPU oracle, at sizes where the tcgen05 node kernels
(node_linear64 split outputs, node_pair_wgrad, node_linear<192> + pna_scale) and the row-map edge kernels are
all active: N >= 8192 / E >= 100 k hub graphs, a 1 M-edge sample drawn by bench.synth_graph at HI-Medium's
edge / node ratio, and the TABGNN / Interleaved / Fused composites at C = H = 64 with >= 4096 token rows.

Per tensor the test reports (and writes to gpurun_out/parity_r02.txt, copied to profiles/) the max error and the
FRACTION of elements further than 1e-5 (relative to the tensor's scale) from the fp64 oracle, next to the same two
numbers for the fp32 oracle itself -- the backbone is piecewise smooth (relu, min / max selection, the std clamp),
so a few elements next to a kink legitimately land on the other side in fp32; what must not happen is a systematic
error in the smooth region, which the fraction bound catches and a widened max-norm tolerance would not.
"""
import os

import pytest
import torch

import b200mp
from oracle import models as OM
from tests.test_gpu_ops import hub_graph

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-5
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPORT = os.path.join(ROOT, "gpurun_out", "parity_r02.txt")


def _stats(got, truth):
    """(max, fraction beyond 1e-5, relative L2, median) of the element-wise error relative to the tensor's scale"""
    got, truth = got.detach().double().cpu(), truth.detach().double().cpu()
    scale = max(float(truth.abs().max()), 1e-30)
    d = (got - truth).abs() / scale
    return (float(d.max()), float((d > TOL).double().mean()), float((got - truth).norm() / max(float(truth.norm()), 1e-30)),
            float(d.flatten().median()) if d.numel() else 0.0)


def _line(case, name, got, refs32, o64, verdict=""):
    k = _stats(got, o64)
    o = [_stats(r, o64) for r in refs32]
    spread = tuple(max(x[i] for x in o) for i in range(4))
    txt = (f"{case:34s} {name:44s} n={o64.numel():>10d}  kernel vs fp64: max {k[0]:.2e} frac>1e-5 {k[1]:.2e} relL2 {k[2]:.2e} median {k[3]:.1e}"
           f"   | fp32 oracle (+{len(refs32) - 1} one-ulp perturbed runs) vs fp64, worst: max {spread[0]:.2e} frac>1e-5 {spread[1]:.2e} "
           f"relL2 {spread[2]:.2e} median {spread[3]:.1e}{verdict}")
    return k, spread, txt


def _emit(txt):
    print(txt)
    try:
        os.makedirs(os.path.dirname(REPORT), exist_ok=True)
        with open(REPORT, "a") as f:
            f.write(txt + "\n")
    except OSError:
        pass


def _check(case, name, got, refs32, o64, forward, flipped=False):
    """Forward values: max-norm at fp32 level.  Gradients: the strict rule -- relative L2 and the share of elements beyond 1e-5 within
    4x what the fp32 oracle itself shows (its plain run and its runs on inputs perturbed by <= 1 ulp) -- OR the signature of a kink
    decision: these networks are piecewise smooth (relu, min / max selection, std clamp), at these sizes a few pre-activations sit within
    an ulp of their kink, and an fp32 evaluation lands on either side (measured: the CPU fp32 oracle's own gradient moves by 3.7e-3 relative
    L2 / 2.5e-2 max between two thread counts on the interleaved case, through ONE hub node).  A flip changes a sparse set of rows by a lot
    and leaves the bulk untouched, so the MEDIAN element-wise error must still be at the oracle's level and the relative L2 error below
    2e-2; a systematic error in the smooth region (1e-4, say) moves the median and fails.  "At the oracle's level" = within 4x the
    oracle's median or 5e-6 (half the 1e-5 tolerance): the 3xTF32 tensor-core products carry ~1e-6 per GEMM against the CPU's 6e-8
    (measured: node linears 4e-7..1.3e-6 relative L2), so the kernels' bulk noise sits at ~2e-6 of the scale after two layers."""
    k, o, txt = _line(case, name, got, refs32, o64)
    strict = k[2] < max(TOL, 4 * o[2]) and k[1] <= max(1e-4, 4 * o[1]) and (not forward or k[0] < max(TOL, 4 * o[0]))
    kink = (not forward) and k[2] < 2e-2 and k[3] <= max(5e-6, 4 * o[3])
    # parameter gradients are sums over ALL rows: a kink decision that an input gradient of the same run has already shown (flipped)
    # shifts every element of them coherently, so only the relative L2 bound applies there
    shifted = (not forward) and flipped and k[2] < 2e-2
    # a gradient the fp32 ORACLE cannot deliver at 1e-5 either (its own relative L2 distance to fp64 exceeds the tolerance: sums with
    # massive cancellation, e.g. a bias in front of a normalisation) amplifies forward round-off by its condition number.  The kernels'
    # forward round-off is the 3xTF32 one, 2^-20 per product against fp32's 2^-24, so up to 16x the oracle's own error is the same
    # conditioning, not a different result; the bulk bound (2e-2) still applies
    ill = (not forward) and o[2] > TOL and k[2] <= 16 * o[2] and k[2] < 2e-2
    _emit(txt + ("" if strict else "   [kink decision: bulk at oracle level]" if kink else
                 "   [sums over rows of a run with a kink decision]" if shifted else
                 "   [ill-conditioned: within 16x the fp32 oracle's own distance to fp64]" if ill else "   [FAIL]"))
    assert strict or kink or shifted or ill, (case, name, k, o)
    return strict


def _jitter(t, gen):
    return t * (1 + (torch.rand(t.shape, generator=gen, dtype=torch.float64) * 2 - 1).to(t.dtype) * 2.0 ** -23)


def _oracle(fn, sd, ins, gin, ws, dtype, jitter_seed=None):
    if jitter_seed is not None:            # every float input / parameter perturbed by <= 1 ulp
        gen = torch.Generator().manual_seed(jitter_seed)
        sd = {k: (_jitter(v, gen) if (v.is_floating_point() and "running" not in k and "avg_deg" not in k) else v) for k, v in sd.items()}
        ins = {k: (_jitter(v, gen) if v.is_floating_point() else v) for k, v in ins.items()}
    p = {k: (v.detach().clone().to(dtype) if v.is_floating_point() else v.clone()) for k, v in sd.items()}
    for k, v in p.items():
        v.requires_grad_(v.is_floating_point() and "running" not in k and "avg_deg" not in k and not k.endswith("eps"))
    for k in list(p):
        if ".conv_back.nn." in k:
            p[k] = p[k.replace(".conv_back.nn.", ".conv_forw.nn.")]
    i = {k: (v.detach().clone().to(dtype) if v.is_floating_point() else v.clone()) for k, v in ins.items()}
    for k in gin:
        i[k].requires_grad_(True)
    outs = fn(p, i)
    sum((o * w.to(dtype)).sum() for o, w in zip(outs, ws)).backward()
    return [o.detach() for o in outs], {k: i[k].grad for k in gin}, {k: v.grad for k, v in p.items() if v.grad is not None}


def _run_case(case, model, oracle_fn, dropin_fn, ins, gin, seed):
    """model: CPU module (moved to the GPU here); oracle_fn(p, ins) / dropin_fn(model, ins) -> tuple of outputs"""
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    model = model.to(DEV).train()
    di = {k: v.to(DEV) for k, v in ins.items()}
    for k in gin:
        di[k].requires_grad_(True)
    outs = dropin_fn(model, di)
    g = torch.Generator().manual_seed(seed)
    ws = [torch.randn(o.shape, generator=g) for o in outs]
    sum((o * w.to(DEV)).sum() for o, w in zip(outs, ws)).backward()
    r32 = [_oracle(oracle_fn, sd, ins, gin, ws, torch.float32)] + [_oracle(oracle_fn, sd, ins, gin, ws, torch.float32, s) for s in (11, 12)]
    o64 = _oracle(oracle_fn, sd, ins, gin, ws, torch.float64)
    for i, o in enumerate(outs):
        _check(case, f"out{i}", o, [r[0][i] for r in r32], o64[0][i], True)
    flipped = False
    for k in gin:
        flipped |= not _check(case, f"grad_in.{k}", di[k].grad, [r[1][k] for r in r32], o64[1][k], False)
    gp = {k: p.grad for k, p in model.named_parameters(remove_duplicate=False) if p.grad is not None}
    # all parameter gradients as ONE vector (per-tensor scales differ by orders of magnitude; biases in front of a
    # BatchNorm have mathematically zero gradients) plus each weight matrix on its own
    keys = [k for k in o64[2] if k in gp]
    assert len(keys) == len(o64[2]), sorted(set(o64[2]) - set(gp))
    cat = lambda d: torch.cat([d[k].detach().double().cpu().reshape(-1) for k in keys])
    _check(case, "grad.params (all, concatenated)", cat(gp), [cat(r[2]) for r in r32], cat(o64[2]), False, flipped)
    gmax = float(cat(o64[2]).abs().max())
    for k in keys:
        if float(o64[2][k].abs().max()) > 1e-2 * gmax:
            _check(case, f"grad.{k}", gp[k], [r[2][k] for r in r32], o64[2][k], False, flipped)


def _backbone(kind, n, ei, H, seed, rev=False):
    deg = torch.bincount(torch.bincount(ei[1], minlength=n))
    torch.manual_seed(seed)
    if kind == "pna":
        return b200mp.PNAS(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=H, deg=deg, reverse_mp=rev)
    return b200mp.GINe(num_features=H, num_gnn_layers=2, n_hidden=H, edge_updates=True, edge_dim=H, reverse_mp=rev)


@pytest.mark.parametrize("H", [64, 128])
@pytest.mark.parametrize("kind", ["pna", "gine"])
def test_backbone_vs_oracle_node_kernels_active(kind, H):
    """full forward (node_emb / edge_emb included) at N = 8192, E = 120 k: every node linear runs on the tcgen05 kernels; H = 128
    (pna.py:50 default width) takes the row-linear composition for the per-edge products (functional.wide_tc_available)"""
    n, e = 8192, 120_000
    ei = hub_graph(n, e, 11, 3000)
    g = torch.Generator().manual_seed(5)
    m = _backbone(kind, n, ei, H, 21)
    ins = {"x": torch.randn(n, H, generator=g), "edge_index": ei, "edge_attr": torch.randn(e, 1, H, generator=g)}
    fn = OM.pnas if kind == "pna" else OM.gine
    _run_case(f"{kind} N=8192 E=120k H={H}", m, lambda p, i: fn(p, i["x"], i["edge_index"], i["edge_attr"], 2, True, False),
              lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), ins, ("x", "edge_attr"), 3)


@pytest.mark.parametrize("H", [64, 128])
@pytest.mark.parametrize("kind", ["pna", "gine"])
def test_reverse_mp_vs_oracle_node_kernels_active(kind, H):
    """reverse message passing (gine.py:28-35, pna.py:37-46) at N = 8192, E = 120 k: both convs through the row maps (no permutation
    passes), `lin` over [x, a_in, a_out] on the node kernel without the concatenation (H = 64); H = 128: permuted storage, the
    row-linear composition, `lin` over the concatenation on the general node kernel"""
    n, e = 8192, 120_000
    ei = hub_graph(n, e, 12, 3000)
    g = torch.Generator().manual_seed(6)
    m = _backbone(kind, n, ei, H, 22, rev=True)
    ins = {"x": torch.randn(n, H, generator=g), "edge_index": ei, "edge_attr": torch.randn(e, 1, H, generator=g)}
    fn = OM.pnas if kind == "pna" else OM.gine
    _run_case(f"{kind} reverse_mp N=8192 E=120k H={H}", m, lambda p, i: fn(p, i["x"], i["edge_index"], i["edge_attr"], 2, True, True),
              lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"]), ins, ("x", "edge_attr"), 5)


@pytest.mark.parametrize("kind", ["pna", "gine"])
def test_bench_sample_vs_oracle(kind):
    """the exact region bench.py times (model.message_passing on embedded x, e) on a 1 M-edge sample of the bench
    generator at HI-Medium's edge / node ratio -- the same sample the CPU baseline leg times"""
    import bench
    n_full, e_full, seed = bench.WORKLOADS["hi_medium"]
    e = bench.CPU_SAMPLE_EDGES
    n = max(2, int(round(n_full * e / e_full)))
    ei = bench.synth_graph(n, e, seed, "cpu")
    H = bench.H
    g = torch.Generator().manual_seed(seed)
    m = _backbone(kind, n, ei, H, 1234)
    ins = {"x": torch.randn(n, H, generator=g), "edge_index": ei, "edge_attr": torch.randn(e, H, generator=g)}
    _run_case(f"{kind} bench sample E=1M N={n}", m,
              lambda p, i: OM.mp_stack(p, kind, i["x"], i["edge_index"], i["edge_attr"], 2, True, False, True),
              lambda mod, i: mod.message_passing(i["x"], i["edge_index"], i["edge_attr"]), ins, ("x", "edge_attr"), 4)


@pytest.mark.parametrize("model", ["tabgnn", "interleaved", "fused_mcm", "fused_lp"])
def test_tab_models_h64_vs_oracle(model):
    """TABGNN / TABGNNInterleaved / TABGNNFused at C = H = 64, nhead 8, K = 5 edge columns, 6 000 nodes / 24 000 edges /
    4 096 target rows: column attention projections and all node linears on the tcgen05 kernels, H = 64 edge kernels"""
    n, e, C, K, L, B = 6000, 24_000, 64, 5, 2, 4096
    ei = hub_graph(n, e, 13, 2000)
    deg = torch.bincount(torch.bincount(ei[1], minlength=n))
    g = torch.Generator().manual_seed(9)
    torch.manual_seed(17)
    kw = dict(channels=C, num_layers=L, deg=deg, nhidden=C, edge_dim=K * C, nhead=8, dropout=0.0)
    ea = torch.randn(e, K, C, generator=g)
    if model == "tabgnn":
        m = b200mp.TABGNN(node_dim=C, **kw)
        ins = {"x": torch.randn(n, 1, C, generator=g), "edge_index": ei, "edge_attr": ea}
        ofn = lambda p, i: OM.tabgnn(p, i["x"], i["edge_index"], i["edge_attr"], L, 8)
        dfn = lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"])
        gin = ("x", "edge_attr")
    elif model == "interleaved":
        m = b200mp.TABGNNInterleaved(node_dim=C, **kw)
        ins = {"x": torch.randn(n, C, generator=g), "edge_index": ei, "edge_attr": ea}
        ofn = lambda p, i: OM.tabgnn_interleaved(p, i["x"], i["edge_index"], i["edge_attr"], L, C, 8)
        dfn = lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"])
        gin = ("x", "edge_attr")
    else:
        lp = model == "fused_lp"
        m = b200mp.TABGNNFused(node_dim=C, **kw)
        tei = torch.stack([torch.randint(0, n, (B,), generator=g), torch.randint(0, n, (B,), generator=g)])
        tei[:, :64] = tei[:, 64:128]                      # repeated target nodes: the pooled write-back averages them
        ins = {"x": torch.randn(n, C, generator=g), "edge_index": ei, "edge_attr": ea, "target_edge_index": tei,
               "target_edge_attr": torch.randn(B, K, C, generator=g)}
        ofn = lambda p, i: OM.tabgnn_fused(p, i["x"], i["edge_index"], i["edge_attr"], i["target_edge_index"],
                                           i["target_edge_attr"], L, C, lp=lp, nhead=8)
        dfn = lambda mod, i: mod(i["x"], i["edge_index"], i["edge_attr"], i["target_edge_index"], i["target_edge_attr"], lp=lp)
        gin = ("x", "edge_attr", "target_edge_attr")
    _run_case(f"{model} C=H=64 N={n} E={e}", m, ofn, dfn, ins, gin, 6)
