"""-m gpu: the fused message-forward + segmented-reduction kernel (csrc/msg_reduce_umma.cu) against the two-kernel path it
replaces (edge_fwd_umma.cu + seg_reduce.cu, themselves checked against the emulation / oracle elsewhere): messages bit for
bit, min / max / tie counts bit for bit, mean / std / sums to fp32 re-association.  Graphs: hubs crossing thousands of tiles,
empty segments, runs ending exactly on tile and quarter boundaries, E below one tile."""
import os

import pytest
import torch

from b200mp import functional as BF
from b200mp import graph as G
from tests.util import rel_err

pytestmark = pytest.mark.gpu
H = 64


def _graph(kind, n, e, seed):
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    if kind == "uniform":
        dst = torch.randint(0, n, (e,), generator=g)
    elif kind == "hubs":                       # heavy tail: a few destinations own most edges, many own none
        dst = (torch.rand(e, generator=g) ** 6 * n).long().clamp_(max=n - 1)
    elif kind == "blocks":                     # every segment exactly 32 rows: runs end on quarter and tile boundaries
        dst = torch.arange(e) // 32
    else:                                      # one segment
        dst = torch.full((e,), n // 2)
    return torch.stack([src, dst]).cuda()


def _both(kind_id, seg, e, eid, W, b, T1, T2):
    os.environ["B200MP_FUSED_REDUCE"] = "0"
    ref = BF._message_fwd_reduce(kind_id, seg, e, eid, W, b, T1, T2)
    os.environ["B200MP_FUSED_REDUCE"] = "1"
    got = BF._message_fwd_reduce(kind_id, seg, e, eid, W, b, T1, T2)
    return ref, got


@pytest.mark.parametrize("graph,n,e", [("uniform", 5000, 60000), ("hubs", 3000, 400001), ("blocks", 4096, 131072), ("single", 10, 70000),
                                       ("uniform", 50, 37), ("hubs", 200000, 3000000)])
@pytest.mark.parametrize("kind_id", [0, 1])
@pytest.mark.parametrize("rowmap", [False, True])
def test_fused_equals_two_kernels(graph, n, e, kind_id, rowmap):
    if rowmap and e > 500000:
        pytest.skip("row-map variant covered at the smaller sizes")
    ei = _graph(graph, n, e, n + e)
    seg = G.edge_structure(ei, n).fwd
    g = torch.Generator(device="cuda").manual_seed(1)
    ea = torch.randn(e, H, device="cuda", generator=g)
    W = torch.randn(H, H, device="cuda", generator=g) / 8
    b = torch.randn(H, device="cuda", generator=g)
    T1 = torch.randn(n, H, device="cuda", generator=g)
    T2 = torch.randn(n, H, device="cuda", generator=g) if kind_id == 0 else None
    if rowmap:
        e_in, eid = ea, seg.perm32()                       # e in the caller's order, read through the CSC permutation
    else:
        e_in, eid = BF.gather_rows(ea, seg.perm), None
    (m0, a0, t0), (m1, a1, t1) = _both(kind_id, seg, e_in, eid, W, b, T1, T2)
    assert torch.equal(m0, m1)
    if kind_id == 0:
        assert torch.equal(a0[:, H:3 * H], a1[:, H:3 * H])                      # min | max
        assert torch.equal(t0, t1)                                              # tie counts
        # mean: fp32 re-association.  std = sqrt(E[m^2] - E[m]^2): the cancellation amplifies the re-association of the two sums
        # (both paths are equally far from fp64; on hubs with 1e5+ rows they differ from EACH OTHER by up to ~1e-3 of the scale)
        assert rel_err(a1[:, :H], a0[:, :H]) < 2e-6 and rel_err(a1[:, 3 * H:], a0[:, 3 * H:]) < 2e-3
        sd = (a1[:, 3 * H:].double() - a0[:, 3 * H:].double())
        assert float(sd.norm() / a0[:, 3 * H:].double().norm()) < 2e-5
        deg = seg.ptr[1:] - seg.ptr[:-1]
        assert float(a1[deg == 0].abs().max() if bool((deg == 0).any()) else 0.0) == 0.0
    else:
        assert rel_err(a1, a0) < 2e-6
    # run-to-run identical (no atomics)
    again = BF._message_fwd_reduce(kind_id, seg, e_in, eid, W, b, T1, T2)
    assert torch.equal(again[1], a1) and (kind_id == 1 or torch.equal(again[2], t1))


@pytest.mark.parametrize("e,n", [(37, 50), (128, 10), (100_001, 9000), (2_000_000, 150_000)])
@pytest.mark.parametrize("maps", [False, True])
def test_update_forward_pipelined_equals_serial(e, n, maps):
    """the software-pipelined edge-update forward (two operand sets, next tile's first product issued behind this tile's second)
    computes exactly what the one-tile-at-a-time kernel computes: same operands, same MMA order per tile -> bit for bit"""
    g = torch.Generator(device="cuda").manual_seed(e)
    ei = torch.stack([torch.randint(0, n, (e,), device="cuda", generator=g), torch.randint(0, n, (e,), device="cuda", generator=g)])
    st = G.edge_structure(ei, n)
    ea = torch.randn(e, H, device="cuda", generator=g)
    W1, W2 = torch.randn(H, H, device="cuda", generator=g) / 8, torch.randn(H, H, device="cuda", generator=g) / 8
    b1, b2 = torch.randn(H, device="cuda", generator=g), torch.randn(H, device="cuda", generator=g)
    P, Q = torch.randn(n, H, device="cuda", generator=g), torch.randn(n, H, device="cuda", generator=g)
    eid = st.fwd.perm32() if maps else None
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_UPD_PIPE"] = flag
        outs.append(BF._edge_linear_fwd(2, ea, eid, W1, b1, P, st.src32, Q, st.dst32, W2, b2, 0.5, 0.5, eid))
    os.environ.pop("B200MP_UPD_PIPE")
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("kind_id", [0, 1])
@pytest.mark.parametrize("e,n", [(37, 50), (64, 10), (100_001, 9000), (2_000_000, 150_000)])
def test_message_backward_64_edge_tiles_equals_128(kind_id, e, n):
    """B200MP_MSG_BWD64=1 (64-edge tiles, 256 threads, two CTAs per SM) computes what the 128-edge kernel computes: g_e row by row with
    the same MMA order (bit for bit), node / weight / bias gradients to re-association"""
    g = torch.Generator(device="cuda").manual_seed(e + kind_id)
    ei = torch.stack([torch.randint(0, n, (e,), device="cuda", generator=g), (torch.rand(e, device="cuda", generator=g) ** 3 * n).long().clamp_(max=n - 1)])
    seg = G.edge_structure(ei, n).fwd
    m = torch.randn(e, H, device="cuda", generator=g)
    ea = torch.randn(e, H, device="cuda", generator=g)
    W = torch.randn(H, H, device="cuda", generator=g) / 8
    node_in = torch.randn(n, 7 * H if kind_id == 0 else H, device="cuda", generator=g)
    acc = torch.randn(e, H, device="cuda", generator=g)
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_MSG_BWD64"] = flag
        outs.append(BF._message_bwd(kind_id, seg, m, ea, node_in, W, True, n, acc.clone(), seg.perm32()))
    os.environ.pop("B200MP_MSG_BWD64")
    (ge0, gT0, gW0, gb0), (ge1, gT1, gW1, gb1) = outs
    assert torch.equal(ge0, ge1)
    assert rel_err(gT1, gT0) < 1e-5 and rel_err(gW1, gW0) < 1e-5 and rel_err(gb1, gb0) < 1e-5


@pytest.mark.parametrize("e,n", [(37, 50), (64, 10), (129, 40), (100_001, 9000), (2_000_000, 150_000)])
def test_edge_update_backward_split_barriers_equals_single(e, n):
    """B200MP_UPD_BWD_SPLIT=1 (data / weight-gradient products on separate mbarriers, accumulator rows staged in their own tile) runs the
    same MMAs in the same order as the single-barrier phases: g_e bit for bit, the scattered node gradients and the per-CTA partial sums
    to re-association"""
    g = torch.Generator(device="cuda").manual_seed(e)
    ei = torch.stack([torch.randint(0, n, (e,), device="cuda", generator=g), (torch.rand(e, device="cuda", generator=g) ** 3 * n).long().clamp_(max=n - 1)])
    st = G.edge_structure(ei, n)
    g_out = torch.randn(e, H, device="cuda", generator=g)
    h = torch.randn(e, H, device="cuda", generator=g).relu_()
    ea = torch.randn(e, H, device="cuda", generator=g)
    W1 = torch.randn(H, H, device="cuda", generator=g) / 8
    W2 = torch.randn(H, H, device="cuda", generator=g) / 8
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_UPD_BWD_SPLIT"] = flag
        outs.append(BF._edge_update_bwd(st, g_out, h, ea, W1, W2, 0.5, 0.5, True, True, g_eid=st.fwd.perm32(), e_eid=None))
    os.environ.pop("B200MP_UPD_BWD_SPLIT")
    a, b = outs
    assert torch.equal(a[0], b[0])
    for x, y in zip(a[1:], b[1:]):
        assert rel_err(y, x) < 1e-5


@pytest.mark.parametrize("maps", [False, True])
@pytest.mark.parametrize("e,n", [(37, 50), (128, 10), (129, 40), (100_001, 9000), (2_000_000, 150_000)])
def test_edge_update_forward_gather_prefetch_equals_serial(e, n, maps):
    """B200MP_UPD_GPF=1 (the next tile's P[src] / Q[dst] rows gathered behind this tile's second product, indices double buffered) is
    the same arithmetic on the same operands as the kernel that gathers between its first product and its epilogue: bit for bit"""
    g = torch.Generator(device="cuda").manual_seed(e + 3)
    ei = torch.stack([torch.randint(0, n, (e,), device="cuda", generator=g), torch.randint(0, n, (e,), device="cuda", generator=g)])
    st = G.edge_structure(ei, n)
    ea = torch.randn(e, H, device="cuda", generator=g)
    W1, W2 = torch.randn(H, H, device="cuda", generator=g) / 8, torch.randn(H, H, device="cuda", generator=g) / 8
    b1, b2 = torch.randn(H, device="cuda", generator=g), torch.randn(H, device="cuda", generator=g)
    P, Q = torch.randn(n, H, device="cuda", generator=g), torch.randn(n, H, device="cuda", generator=g)
    eid = st.fwd.perm32() if maps else None
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_UPD_GPF"] = flag
        outs.append(BF._edge_linear_fwd(2, ea, eid, W1, b1, P, st.src32, Q, st.dst32, W2, b2, 0.5, 0.5, eid))
    os.environ.pop("B200MP_UPD_GPF")
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("e,n,dt", [(37, 50, torch.float32), (100_001, 9000, torch.float32), (2_000_000, 150_000, torch.float32),
                                    (100_001, 9000, torch.bfloat16)])
def test_segment_reduce_two_pass_equals_one_pass(e, n, dt):
    """B200MP_SEG_TWOPASS=1 (sums / extrema first, tie counts against the final extrema in a second sweep over the same rows) against the
    one-pass bookkeeping: same summation order, same extrema, integer counts -> bit for bit, ties included (quantised messages)"""
    g = torch.Generator(device="cuda").manual_seed(e)
    ei = torch.stack([torch.randint(0, n, (e,), device="cuda", generator=g), (torch.rand(e, device="cuda", generator=g) ** 3 * n).long().clamp_(max=n - 1)])
    seg = G.edge_structure(ei, n).fwd
    m = (torch.randn(e, H, device="cuda", generator=g) * 4).round() / 4          # plenty of exact ties
    m = m.to(dt)
    outs = []
    for flag in ("0", "1"):
        os.environ["B200MP_SEG_TWOPASS"] = flag
        outs.append(BF._segment_reduce(0, seg, m, H))
    os.environ.pop("B200MP_SEG_TWOPASS")
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert float(outs[0][1].max()) > 1.0            # ties did occur
