"""CPU: the C-ABI library loads, exports every entry point include/b200mp.h declares, the ctypes table binds exactly
those, the by-value struct matches the header's layout, and the product path refuses CPU tensors (no fallback).
No kernel is launched here."""
import ctypes as C
import os
import re

import pytest
import torch

import b200mp
from b200mp import _lib as L
from b200mp import functional as BF


@pytest.fixture(scope="module", autouse=True)
def _built_library():
    """a fresh checkout has no libb200mp.so (git-ignored): build it (nvcc cross-compiles without a GPU, ~40 s)"""
    if not os.path.isfile(L.LIB_PATH):
        import __graft_entry__ as g
        g.build()


def test_library_exports_every_declared_entry_point():
    lib = L.lib()
    declared = L.declared_symbols()
    assert len(declared) >= 40
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    bound = set(L._SIGS)
    assert bound == set(declared), (sorted(bound - set(declared)), sorted(set(declared) - bound))


def test_argument_counts_match_the_header():
    text = open(L.HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    for name, args in L._SIGS.items():
        m = re.search(r"\bint\s+" + name + r"\s*\((.*?)\)\s*;", text, flags=re.S)
        assert m, name
        params = [p for p in m.group(1).split(",") if p.strip() and p.strip() != "void"]
        assert len(params) == len(args), (name, len(params), len(args))


def test_column_struct_layout():
    # struct b200mp_column: int32 kind, int32 offset, int64 stride, 3 pointers, 3 floats, int32 ncat
    assert C.sizeof(L.Column) == 56
    offs = {f[0]: getattr(L.Column, f[0]).offset for f in L.Column._fields_}
    assert offs == {"kind": 0, "offset": 4, "stride": 8, "feat": 16, "weight": 24, "bias": 32, "mean": 40, "std": 44,
                    "min_year": 48, "ncat": 52}


def test_product_path_refuses_cpu_tensors():
    q = torch.randn(4, 6, 64)
    with pytest.raises(L.B200MPError):
        BF.col_attention(q, q, q, 8)
    with pytest.raises(L.B200MPError):
        BF.add_dropout_layernorm(torch.randn(8, 64), None, 0.0, torch.nn.LayerNorm(64))
    with pytest.raises(L.B200MPError):
        BF.bn_relu_residual(torch.nn.BatchNorm1d(64), torch.randn(8, 64), torch.randn(8, 64))
    with pytest.raises((L.B200MPError, RuntimeError)):
        b200mp.edge_structure(torch.zeros(2, 4, dtype=torch.long), 3)


def test_header_cites_the_reference():
    text = open(L.HEADER).read()
    for cite in ("negative_sampling.cpp", "gine.py", "pna.py", "tabgnn.py", "batch_processing.py"):
        assert cite in text, cite
    assert os.path.isfile(L.LIB_PATH)
