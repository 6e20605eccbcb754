"""Pins oracle/models.py (the functional restatement) to the fixtures produced by the
reference's own model files (oracle/gen_golden.py).  CPU only."""
import os

import pytest
import torch

from oracle import models as OM
from tests.util import GOLDEN, load_golden, rel_err

TOL = 2e-6   # same arithmetic, same library kernels: only reassociation noise is expected


@pytest.fixture(autouse=True)
def _single_thread():
    """The fixtures were generated single-threaded; PNA's std gradient (1/std near the 1e-5
    clamp) amplifies summation-order noise to ~4e-5, so reproduce the generating conditions."""
    n = torch.get_num_threads()
    torch.set_num_threads(1)
    yield
    torch.set_num_threads(n)


def _eval(rec, fn, dtype):
    p = {k: (v.detach().clone().to(dtype) if v.is_floating_point() else v.clone()) for k, v in rec["sd"].items()}
    for k, v in p.items():
        trainable = v.is_floating_point() and "running" not in k and "avg_deg" not in k and not k.endswith("eps")
        v.requires_grad_(trainable)
    for k in list(p):   # GINEConvHetero hands ONE `network` to both convs (gine.py:50-55): tie them
        if ".conv_back.nn." in k:
            p[k] = p[k.replace(".conv_back.nn.", ".conv_forw.nn.")]
    ins = {k: (v.detach().clone().to(dtype) if v.is_floating_point() else v.clone()) for k, v in rec["in"].items()}
    for k in rec["gin"]:
        ins[k].requires_grad_(True)
    outs = fn(p, ins, rec["meta"])
    loss = sum((o * rec["w"][str(i)].to(dtype)).sum() for i, o in enumerate(outs))
    loss.backward()
    return p, ins, outs


def _run(name, fn):
    """oracle(fp32) must agree with the reference-generated fixture to 1e-5, or - where PNA's
    std gradient (1/std near the 1e-5 clamp) amplifies summation-order noise - to within 4x the
    fixture's own distance from the fp64 oracle (the round-off yardstick)."""
    rec = load_golden(name)
    p, ins, outs = _eval(rec, fn, torch.float32)
    p64, ins64, outs64 = _eval(rec, fn, torch.float64)

    def check(got, want, truth, what, floor=1e-30):
        tol = max(1e-5, 4 * rel_err(want, truth, floor))
        assert tol < 1e-3, f"{name} {what}: fixture itself is {tol / 4:.2e} from fp64"
        assert rel_err(got, want, floor) < tol, f"{name} {what}: {rel_err(got, want, floor):.3e} (tol {tol:.1e})"

    for i, o in enumerate(outs):
        check(o, rec["out"][str(i)], outs64[i], f"out {i}")
    for k, g in rec["gin"].items():
        check(ins[k].grad, g, ins64[k].grad, f"grad input {k}")
    gmax = max(float(g.abs().max()) for g in rec["g"].values())
    for k, g in rec["g"].items():
        if p[k].grad is None:
            continue
        # a bias that only shifts a BatchNorm input has an exactly-zero true gradient (a sum that
        # cancels to round-off in both implementations): judge biases against the gradient scale.
        floor = (0.2 if k.endswith("bias") else 1e-3) * gmax
        check(p[k].grad, g, p64[k].grad, f"grad {k}", floor)
    for k, v in rec["sd_after"].items():
        assert rel_err(p[k].detach().to(v.dtype), v) < TOL, f"{name} running stat {k}"
    return rec


@pytest.mark.parametrize("name", ["gine_eu", "gine_eu_rev", "gine_noeu"])
def test_gine(name):
    _run(name, lambda p, i, m: OM.gine(p, i["x"], i["edge_index"], i["edge_attr"], m["L"], m["edge_updates"], m["reverse_mp"]))


@pytest.mark.parametrize("name", ["pnas_eu", "pnas_eu_rev", "pnas_h64"])
def test_pnas(name):
    _run(name, lambda p, i, m: OM.pnas(p, i["x"], i["edge_index"], i["edge_attr"], m["L"], m["edge_updates"], m["reverse_mp"]))


def test_tabgnn():
    _run("tabgnn", lambda p, i, m: OM.tabgnn(p, i["x"], i["edge_index"], i["edge_attr"], m["L"], m["nhead"]))


def test_interleaved():
    _run("interleaved", lambda p, i, m: OM.tabgnn_interleaved(p, i["x"], i["edge_index"], i["edge_attr"], m["L"], m["node_dim"], m["nhead"]))


@pytest.mark.parametrize("name", ["fused_mcm", "fused_lp"])
def test_fused(name):
    _run(name, lambda p, i, m: OM.tabgnn_fused(p, i["x"], i["edge_index"], i["edge_attr"], i["target_edge_index"],
                                                i["target_edge_attr"], m["L"], m["node_dim"], lp=m["lp"], nhead=m["nhead"]))


@pytest.mark.parametrize("name", ["cpna", "cpnatab"])
def test_column_stacks(name):
    fn = OM.cpna if name == "cpna" else OM.cpnatab
    _run(name, lambda p, i, m: fn(p, i["x"], i["edge_index"], i["edge_attr"], m["L"], m["edge_updates"], m["reverse_mp"]))


def test_losses_and_ranking_metric_pinned_to_reference_vectors():
    """tests/golden/ss_loss_metric.npz holds outputs of the reference's OWN loss.py / metric.py (oracle/gen_golden_r2.py)"""
    import numpy as np
    z = np.load(os.path.join(GOLDEN, "ss_loss_metric.npz"))
    t = lambda k: torch.from_numpy(z[k])
    nn_, K = int(z["num_numerical"]), int(z["num_neg"])
    assert abs(float(OM.lp_loss(t("pos"), t("neg"))) - float(z["lp_loss"])) < 1e-6
    cat_out = [t(f"cat_out_{c}") for c in range(3)]
    total, (cl, tc, acc), (nl, tn) = OM.mcm_loss(cat_out, t("num_out"), t("y"), nn_)
    assert (tc, tn, acc) == (int(z["mcm_t_c"]), int(z["mcm_t_n"]), int(z["mcm_acc"]))
    for got, key in ((total, "mcm_total"), (cl, "mcm_cat_loss"), (nl, "mcm_num_loss")):
        assert abs(float(got) - float(z[key])) < 1e-5 * max(1.0, abs(float(z[key])))
    mrr, hits, _ = OM.mrr(t("pos"), t("neg"), [1, 2, 5, 10], K)
    # the reference's numpy argsort is unstable on the tied scores of sample 0: every other rank is pinned exactly
    B = t("pos").shape[0]
    assert abs(mrr - float(z["mrr"])) <= 1.0 / B and np.allclose([hits[f"hits@{k}"] for k in (1, 2, 5, 10)], z["hits"], atol=1.0 / B)


def test_heads():
    _run("classifier_head", lambda p, i, m: (OM.classifier_head(p, i["x"], i["edge_index"], i["edge_attr"]),))
    _run("linkpred_head", lambda p, i, m: OM.link_pred_head(p, i["x"], i["pos_ei"], i["pos_ea"], i["neg_ei"], i["neg_ea"]))
