"""ORACLE (test infrastructure): load the reference's OWN model files, unmodified and in place,
from /root/reference.  Only usable in the build container (the GPU box has
no /root/reference) -- used by oracle/gen_golden.py to produce tests/golden/*.npz and by the
CPU tests that re-check the oracle against the live reference when it is present.

``src/__init__.py`` is never imported (it drags in datasets + the compiled sampler,
src/__init__.py:3-5); the six model files are loaded by path under a private ``refsrc`` package
with ``oracle/shim`` providing torch_geometric / torch_frame.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("B200MP_REFERENCE", "/root/reference")
_SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shim")
_REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "src", "nn", "gnn", "pna.py"))


def _pkg(name: str):
    if name not in sys.modules:
        m = types.ModuleType(name)
        m.__path__ = []  # namespace-like
        sys.modules[name] = m
    return sys.modules[name]


def _load(modname: str, relpath: str):
    if modname in sys.modules:
        return sys.modules[modname]
    spec = importlib.util.spec_from_file_location(modname, os.path.join(REF_ROOT, relpath))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[modname] = mod
    spec.loader.exec_module(mod)
    return mod


def load():
    """Returns a namespace with the reference classes: GINe, PNAS, TABGNN, TABGNNFused,
    TABGNNInterleaved, ClassifierHead, LinkPredHead, ..."""
    if not available():
        raise RuntimeError(f"reference not found under {REF_ROOT}")
    for p in (_REPO, _SHIM):
        if p not in sys.path:
            sys.path.insert(0, p)
    for n in ("refsrc", "refsrc.nn", "refsrc.nn.gnn", "refsrc.nn.models"):
        _pkg(n)
    gine = _load("refsrc.nn.gnn.gine", "src/nn/gnn/gine.py")
    pna = _load("refsrc.nn.gnn.pna", "src/nn/gnn/pna.py")
    dec = _load("refsrc.nn.gnn.decoder", "src/nn/gnn/decoder.py")
    tabgnn = _load("refsrc.nn.models.tabgnn", "src/nn/models/tabgnn.py")
    fused = _load("refsrc.nn.models.fused", "src/nn/models/fused.py")
    inter = _load("refsrc.nn.models.inteleaved", "src/nn/models/inteleaved.py")
    ns = types.SimpleNamespace(
        GINe=gine.GINe, GINEConvHetero=gine.GINEConvHetero,
        PNAS=pna.PNAS, PNAConvHetero=pna.PNAConvHetero, CPNA=pna.CPNA, CPNATAB=pna.CPNATAB,
        ClassifierHead=dec.ClassifierHead, LinkPredHead=dec.LinkPredHead,
        NodeClassificationHead=dec.NodeClassificationHead,
        TABGNN=tabgnn.TABGNN, TABGNNFused=fused.TABGNNFused,
        TABGNNInterleaved=inter.TABGNNInterleaved,
    )
    return ns


def load_loss_metric():
    """The reference's own SSLoss / SSMetric (src/utils/loss.py, src/utils/metric.py: plain torch / numpy files, loaded by
    path so that src/utils/__init__.py is never executed)."""
    loss = _load("refsrc_utils_loss", "src/utils/loss.py")
    metric = _load("refsrc_utils_metric", "src/utils/metric.py")
    return loss.SSLoss, metric.SSMetric
