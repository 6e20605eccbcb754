"""Read-out side: heads, self-supervised losses, ranking metric, pooled write-back.
CPU: host logic with the emulated kernels against the golden fixtures (reference's own decoder.py / loss.py / metric.py
outputs).  -m gpu: the same through the real kernels, plus oracle comparisons at larger sizes."""
import numpy as np
import pytest
import torch

import b200mp
from b200mp import functional as BF
from oracle import models as OM
from tests.emulator import emulated_kernels
from tests.util import GOLDEN, load_golden, rel_err


def _head_case(name, device):
    rec = load_golden(name)
    H = rec["in"]["x"].shape[1]
    m = (b200mp.ClassifierHead(n_classes=2, n_hidden=H, dropout=0.0) if name == "classifier_head"
         else b200mp.LinkPredHead(n_classes=1, n_hidden=H, dropout=0.0))
    m.load_state_dict(rec["sd"], strict=True)
    m = m.to(device).train()
    ins = {k: v.clone().to(device) for k, v in rec["in"].items()}
    for k in rec["gin"]:
        ins[k].requires_grad_(True)
    outs = ((m(ins["x"], ins["edge_index"], ins["edge_attr"]),) if name == "classifier_head"
            else m(ins["x"], ins["pos_ei"], ins["pos_ea"], ins["neg_ei"], ins["neg_ea"]))
    sum((o * rec["w"][str(i)].to(device)).sum() for i, o in enumerate(outs)).backward()
    for i, o in enumerate(outs):
        assert rel_err(o, rec["out"][str(i)]) < 1e-5, (name, i)
    for k in rec["gin"]:
        assert rel_err(ins[k].grad, rec["gin"][k]) < 1e-5, (name, k)
    for k, p in m.named_parameters():
        assert rel_err(p.grad, rec["g"][k]) < 1e-5, (name, k)


def _loss_metric_case(device):
    z = np.load(f"{GOLDEN}/ss_loss_metric.npz")
    t = lambda k: torch.from_numpy(z[k]).to(device)
    nn_, K = int(z["num_numerical"]), int(z["num_neg"])
    L = b200mp.SSLoss(device, num_numerical=nn_)
    assert abs(float(L.lp_loss(t("pos"), t("neg"))) - float(z["lp_loss"])) < 1e-5
    cat_out = [t(f"cat_out_{c}").requires_grad_(True) for c in range(3)]
    num_out = t("num_out").requires_grad_(True)
    total, (cl, tc, acc), (nl, tn) = L.mcm_loss(cat_out, num_out, t("y"))
    assert (tc, tn, int(acc)) == (int(z["mcm_t_c"]), int(z["mcm_t_n"]), int(z["mcm_acc"]))
    for got, key in ((total, "mcm_total"), (cl, "mcm_cat_loss"), (nl, "mcm_num_loss")):
        assert abs(float(got) - float(z[key])) < 1e-5 * max(1.0, abs(float(z[key]))), key
    # gradients of the vectorised loss == gradients of the per-sample loop (oracle restatement of loss.py:41-72)
    total.backward()
    oc = [t(f"cat_out_{c}").cpu().requires_grad_(True) for c in range(3)]
    on = t("num_out").cpu().requires_grad_(True)
    OM.mcm_loss(oc, on, t("y").cpu(), nn_)[0].backward()
    for a, b in zip(cat_out + [num_out], oc + [on]):
        assert rel_err(a.grad, b.grad) < 1e-5
    # async variant: counts stay on the device
    _, (_, tc2, _), (_, tn2) = L.mcm_loss([c.detach() for c in cat_out], num_out.detach(), t("y"), sync_counts=False)
    assert torch.is_tensor(tc2) and int(tc2) == tc and int(tn2) == tn
    # ranking: every rank equals the stable-sort restatement; MRR / Hits@k equal the reference's up to its unstable tie
    ranks = b200mp.SSMetric.ranks(t("pos"), t("neg"), K).cpu().numpy()
    _, _, want = OM.mrr(t("pos").cpu(), t("neg").cpu(), [1, 2, 5, 10], K)
    assert np.array_equal(ranks, want)
    mrr, hits = b200mp.SSMetric.mrr(t("pos"), t("neg"), [1, 2, 5, 10], K)
    B = len(want)
    assert abs(mrr - float(z["mrr"])) <= 1.0 / B and np.allclose([hits[f"hits@{k}"] for k in (1, 2, 5, 10)], z["hits"], atol=1.0 / B)
    S = b200mp.SSMetric(nn_)
    mv = torch.randn(32, 6, generator=torch.Generator().manual_seed(0)).to(device)
    yy = torch.stack([torch.zeros(32), mv.argmax(1).cpu().float()], 1).to(device)
    yy[:8, 1] = (yy[:8, 1] + 1) % 6
    assert abs(S.mv_accuracy(mv, yy) - 0.75) < 1e-12
    assert abs(float(L.mv_loss(mv, yy)) - float(torch.nn.functional.cross_entropy(mv, yy[:, 1].long()))) < 1e-6


def _blend_case(device, n, r, W, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, W, generator=g).to(device).requires_grad_(True)
    emb = torch.randn(r, W, generator=g).to(device).requires_grad_(True)
    index = torch.randint(0, max(n // 2, 1), (r,), generator=g).to(device)          # upper half of the nodes never named
    w = torch.randn(n, W, generator=g).to(device)
    out = BF.segment_mean_blend(x, emb, index)
    (out * w).sum().backward()
    # models/fused.py:260-268 restated with unique / index_add_ / bincount
    xd, ed = x.detach().double().cpu().requires_grad_(True), emb.detach().double().cpu().requires_grad_(True)
    uniq, inv = torch.unique(index.cpu(), return_inverse=True)
    pooled = torch.zeros(uniq.numel(), W, dtype=torch.float64).index_add_(0, inv, ed) / torch.bincount(inv).unsqueeze(1)
    ref = xd.clone()
    ref[uniq] = (xd[uniq] + pooled) / 2
    (ref * w.double().cpu()).sum().backward()
    assert rel_err(out, ref) < 1e-6 and rel_err(x.grad, xd.grad) < 1e-6 and rel_err(emb.grad, ed.grad) < 1e-6


@pytest.mark.parametrize("name", ["classifier_head", "linkpred_head"])
def test_heads_host_logic(name):
    with emulated_kernels():
        _head_case(name, "cpu")


def test_losses_and_metric_host_logic():
    with emulated_kernels():
        _loss_metric_case("cpu")


def test_pooled_write_back_host_logic():
    with emulated_kernels():
        _blend_case("cpu", 50, 64, 16, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["classifier_head", "linkpred_head"])
def test_heads_gpu(name):
    _head_case(name, "cuda")


@pytest.mark.gpu
def test_losses_and_metric_gpu():
    _loss_metric_case("cuda")


@pytest.mark.gpu
@pytest.mark.parametrize("n,r,W", [(50, 64, 16), (100_000, 300_000, 64), (7, 0, 32)])
def test_pooled_write_back_gpu(n, r, W):
    _blend_case("cuda", n, r, W, n)


@pytest.mark.gpu
def test_heads_vs_oracle_at_scale_and_determinism():
    """H = 64, 20 000 nodes / 200 000 edges with a hub: node linears on the tcgen05 kernels, gathers on readout.cu; two
    backward passes are bit-identical (no atomics anywhere on the read-out path)."""
    from tests.test_gpu_ops import hub_graph
    n, e, H = 20_000, 200_000, 64
    ei = hub_graph(n, e, 3, 5000)
    g = torch.Generator().manual_seed(1)
    x, ea = torch.randn(n, H, generator=g), torch.randn(e, H, generator=g)
    w = torch.randn(e, 2, generator=g)
    torch.manual_seed(2)
    m = b200mp.ClassifierHead(n_classes=2, n_hidden=H, dropout=0.0)
    p64 = {k: v.detach().double().requires_grad_(True) for k, v in m.state_dict().items()}
    x64, ea64 = x.double().requires_grad_(True), ea.double().requires_grad_(True)
    (OM.classifier_head(p64, x64, ei, ea64) * w.double()).sum().backward()
    m = m.cuda().train()
    grads = []
    for _ in range(2):
        xd, ed = x.cuda().requires_grad_(True), ea.cuda().requires_grad_(True)
        m.zero_grad(set_to_none=True)
        out = m(xd, ei.cuda(), ed)
        (out * w.cuda()).sum().backward()
        grads.append([xd.grad.clone(), ed.grad.clone()] + [p.grad.clone() for p in m.parameters()])
    assert rel_err(out, OM.classifier_head(p64, x64, ei, ea64)) < 1e-5
    # gradients: the kink-aware rule of tests/test_gpu_parity_scale.py (a relu decision within an ulp of its kink moves single rows by
    # O(1): measured 2.9e-4 max-norm / 1e-4 relative L2 on x.grad), against the fp32 oracle's own distance from fp64
    from tests.test_gpu_parity_scale import _check
    p32 = {k: v.detach().float().requires_grad_(True) for k, v in p64.items()}
    x32, ea32 = x.clone().requires_grad_(True), ea.clone().requires_grad_(True)
    (OM.classifier_head(p32, x32, ei, ea32) * w).sum().backward()
    _check("ClassifierHead N=20000 E=200k", "grad_in.x", grads[0][0], [x32.grad], x64.grad, False)
    _check("ClassifierHead N=20000 E=200k", "grad_in.edge_attr", grads[0][1], [ea32.grad], ea64.grad, False)
    for (k, _), gp in zip(m.named_parameters(), grads[0][2:]):
        _check("ClassifierHead N=20000 E=200k", f"grad.{k}", gp, [p32[k].grad], p64[k].grad, False)
    for a, b in zip(*grads):
        assert torch.equal(a, b)
