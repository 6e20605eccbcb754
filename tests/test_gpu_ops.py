"""-m gpu: every C-ABI kernel against the oracle (integer work: bit-exact vs oracle/index_ops.py)
or against the plain-PyTorch emulation of the same op on the same inputs (tests/emulator.py,
itself pinned to the golden fixtures by tests/test_dropin_emulated.py)."""
import numpy as np
import pytest
import torch

import b200mp
from b200mp import functional as BF
from b200mp import graph as G
from oracle import index_ops as OI
from tests import emulator as EM
from tests.util import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


def hub_graph(n, e, seed, hub_edges):
    """Unsorted multigraph: one hub with `hub_edges` in-edges (> LIGHT_MAX -> chunked path), a
    second hub on the source side, isolated destinations, self loops, duplicates."""
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (e,), generator=g)
    dst = torch.randint(0, n - n // 4, (e,), generator=g)
    dst[:hub_edges] = 5
    src[hub_edges:hub_edges + hub_edges // 2] = 7
    src[-20:] = dst[-20:]
    src[-40:-20], dst[-40:-20] = src[:20].clone(), dst[:20].clone()
    p = torch.randperm(e, generator=g)
    return torch.stack([src[p], dst[p]])


def cpu_seg(seg):
    c = lambda t: None if t is None else t.cpu()
    return G.SegmentIndex(seg.num_nodes, seg.num_edges, c(seg.perm), c(seg.ptr), c(seg.other), c(seg.nbr32), c(seg.seg32), c(seg.eid32))


def cpu_struct(st):
    return G.EdgeStructure(st.num_nodes, st.num_edges, cpu_seg(st.fwd), cpu_seg(st.rev) if st.rev is not None else None)


@pytest.mark.parametrize("n,e", [(1, 0), (7, 1), (50, 333), (3000, 50000), (515, 200001)])
def test_csc_build_bit_exact(n, e):
    g = torch.Generator().manual_seed(n + e)
    ei = torch.stack([torch.randint(0, n, (e,), generator=g), torch.randint(0, n, (e,), generator=g)])
    perm, colptr, row = OI.csc_build(ei.numpy(), n)
    ptr_d, row_d, perm_d = b200mp.to_csc(ei.to(DEV), n)
    assert np.array_equal(perm_d.cpu().numpy(), perm)
    assert np.array_equal(ptr_d.cpu().numpy(), colptr)
    assert np.array_equal(row_d.cpu().numpy(), row)
    st = b200mp.edge_structure(ei.to(DEV), n, reverse=True)
    perm2, rowptr, col = OI.csr_build(ei.numpy(), n)
    assert np.array_equal(st.rev.perm.cpu().numpy(), perm2)
    assert np.array_equal(st.rev.ptr.cpu().numpy(), rowptr)
    assert np.array_equal(st.rev.other.cpu().numpy(), col)
    if e:
        # eid maps CSR position -> CSC storage row of the same edge
        assert torch.equal(st.fwd.perm[st.rev.eid32.long()], st.rev.perm)


def test_degree_hist_and_relabel_bit_exact():
    n, e = 4000, 60000
    ei = hub_graph(n, e, 3, 900)
    deg, hist = OI.degree_hist(ei[1].numpy(), n)
    st = b200mp.edge_structure(ei.to(DEV), n)
    deg_d, hist_d = b200mp.degree_histogram(st.fwd.ptr)
    assert np.array_equal(deg_d.cpu().numpy(), deg) and np.array_equal(hist_d.cpu().numpy(), hist)
    assert np.array_equal(b200mp.degree(ei[1].to(DEV), n, dtype=torch.long).cpu().numpy(), deg)
    ids = torch.randint(0, 10 ** 9, (30000,), generator=torch.Generator().manual_seed(5))
    ids[100:200] = ids[:100]
    nodes, loc = OI.relabel(ids[:15000].numpy(), ids[15000:].numpy())
    nodes_d, loc_d = b200mp.relabel(ids.to(DEV))
    assert np.array_equal(nodes_d.cpu().numpy(), nodes)
    assert np.array_equal(loc_d.cpu().numpy(), loc.reshape(-1))


def test_cpu_tensors_are_refused():
    ei = torch.tensor([[0, 1], [1, 0]])
    with pytest.raises(Exception):
        b200mp.edge_structure(ei, 2)


@pytest.mark.parametrize("H", [16, 32, 64, 128])
def test_edge_kernels_vs_emulation(H):
    n, e = 700, 9000
    ei = hub_graph(n, e, 11 + H, 1300).to(DEV)
    st = b200mp.edge_structure(ei, n, reverse=True)
    stc = cpu_struct(st)
    g = torch.Generator().manual_seed(H)
    rnd = lambda *s: torch.randn(*s, generator=g)
    E_, A, B = rnd(e, H), rnd(n, H), rnd(n, H)
    W1, b1, W2, b2 = rnd(H, H) / H ** 0.5, rnd(H), rnd(H, H) / H ** 0.5, rnd(H)
    d = lambda t: t.to(DEV)

    for seg, segc in ((st.fwd, stc.fwd), (st.rev, stc.rev)):
        # forward, all three modes
        for mode in (0, 1, 2):
            args = (W2, b2, 0.5, 0.25) if mode == 2 else ()
            o1, o2 = BF._edge_linear_fwd(mode, d(E_), seg.eid32, d(W1), d(b1), d(B), seg.nbr32, d(A), seg.seg32, *map(lambda t: d(t) if torch.is_tensor(t) else t, args))
            r1, r2 = EM.edge_linear_fwd(mode, E_, segc.eid32, W1, b1, B, segc.nbr32, A, segc.seg32, *args)
            assert rel_err(o1, r1) < 2e-6
            if mode == 2:
                assert rel_err(o2, r2) < 2e-6
        # segmented reductions on the SAME messages
        m = o1_msg = BF._edge_linear_fwd(0, d(E_), seg.eid32, d(W1), d(b1), d(B), seg.nbr32, d(A), seg.seg32)[0]
        agg, tie = BF._segment_reduce(0, seg, m, H)
        ragg, rtie = EM.segment_reduce(0, segc, m.cpu(), H)
        assert torch.equal(tie.cpu(), rtie)
        assert torch.equal(agg[:, H:3 * H].cpu(), ragg[:, H:3 * H])            # min / max exact
        assert rel_err(agg[:, :H], ragg[:, :H]) < 2e-6
        # std = sqrt(relu(E[m^2] - mean^2) + 1e-5): where the variance vanishes (one-edge segments) the square root turns a 1e-7
        # rounding difference of the variance into ~1e-4 of the std, so the 1e-5 bound is checked on the variance, the std loosely
        assert rel_err(agg[:, 3 * H:] ** 2, ragg[:, 3 * H:] ** 2) < 1e-5
        assert rel_err(agg[:, 3 * H:], ragg[:, 3 * H:]) < 1e-3
        s, _ = BF._segment_reduce(1, seg, m, H)
        rs, _ = EM.segment_reduce(1, segc, m.cpu(), H)
        assert rel_err(s, rs) < 2e-6
        # message backward, PNA and GINE kinds
        coef = torch.cat([rnd(n, H), rnd(n, H), rnd(n, H), ragg[:, H:2 * H], rnd(n, H), ragg[:, 2 * H:3 * H], ragg[:, :H]], 1)
        for kind, node_in in ((0, coef), (1, rnd(n, H))):
            got = BF._message_bwd(kind, seg, m, d(E_), d(node_in), d(W1), True, n)
            ref = EM.message_bwd(kind, segc, m.cpu(), E_, node_in, W1, True, n)
            for a, b, nm in zip(got, ref, ("g_e", "gT1", "gW", "gb")):
                assert rel_err(a, b) < 1e-5, (kind, nm, rel_err(a, b))
    # edge update backward
    h = BF._edge_linear_fwd(2, d(E_), None, d(W1), d(b1), d(B), st.src32, d(A), st.dst32, d(W2), d(b2), 1.0, 0.5)[0]
    gout = rnd(e, H)
    got = BF._edge_update_bwd(st, d(gout), h, d(E_), d(W1), d(W2), 1.0, 0.5, True, True)
    ref = EM.edge_update_bwd(stc, gout, h.cpu(), E_, W1, W2, 1.0, 0.5, True, True)
    for a, b, nm in zip(got, ref, ("g_e", "gP", "gQ", "gW1", "gb1", "gW2", "gb2")):
        assert rel_err(a, b) < 1e-5, (nm, rel_err(a, b))
    # row maps (first / last layer read and write the caller's edge order)
    p32 = st.fwd.perm32()
    o1, o2 = BF._edge_linear_fwd(2, d(E_), p32, d(W1), d(b1), d(B), st.src32, d(A), st.dst32, d(W2), d(b2), 1.0, 0.5, p32)
    r1, r2 = EM.edge_linear_fwd(2, E_, p32.cpu(), W1, b1, B, stc.src32, A, stc.dst32, W2, b2, 1.0, 0.5, p32.cpu())
    assert rel_err(o1, r1) < 2e-6 and rel_err(o2, r2) < 2e-6
    if H in (64, 128):          # row maps in the backward: the tensor-core paths (fused kernels at 64, the row-linear composition at 128)
        got = BF._edge_update_bwd(st, d(gout), h, d(E_), d(W1), d(W2), 1.0, 0.5, True, True, None, None, p32, p32)
        ref = EM.edge_update_bwd(stc, gout, h.cpu(), E_, W1, W2, 1.0, 0.5, True, True, None, None, p32.cpu(), p32.cpu())
        for a, b, nm in zip(got, ref, ("g_e", "gP", "gQ", "gW1", "gb1", "gW2", "gb2")):
            assert rel_err(a, b) < 1e-5, ("mapped", nm, rel_err(a, b))
        acc = rnd(e, H)
        m = BF._edge_linear_fwd(0, d(E_), p32, d(W1), d(b1), d(B), st.fwd.nbr32, d(A), st.fwd.seg32)[0]
        for g_acc, eid in ((acc, p32), (acc, None), (None, p32)):
            got = BF._message_bwd(1, st.fwd, m, d(E_), d(A), d(W1), True, n, None if g_acc is None else d(g_acc).clone(), eid)
            ref = EM.message_bwd(1, stc.fwd, m.cpu(), E_, A, W1, True, n, None if g_acc is None else g_acc.clone(),
                                 None if eid is None else eid.cpu())
            for a, b, nm in zip(got, ref, ("g_e", "gT1", "gW", "gb")):
                assert rel_err(a, b) < 1e-5, ("acc/map", nm, rel_err(a, b))


def test_permute_roundtrip_and_empty():
    x = torch.randn(1000, 64, device=DEV, requires_grad=True)
    p = torch.randperm(1000, device=DEV)
    y = BF.gather_rows(x, p)
    assert torch.equal(y, x[p])
    assert torch.equal(BF.scatter_rows(y, p), x)
    y.sum().backward()
    assert torch.equal(x.grad, torch.ones_like(x))
    # empty graph: no edges at all
    ei = torch.zeros((2, 0), dtype=torch.long, device=DEV)
    st = b200mp.edge_structure(ei, 5)
    agg, tie = BF.pna_aggregate(torch.zeros(5, 32, device=DEV), torch.zeros(5, 32, device=DEV), torch.zeros(0, 32, device=DEV),
                                torch.eye(32, device=DEV), torch.zeros(32, device=DEV), st.fwd)
    assert agg.shape == (5, 128) and float(agg.abs().max()) == 0.0


@pytest.mark.parametrize("e,n", [(1, 3), (127, 40), (128, 40), (4097, 300), (300_001, 20_000)])
@pytest.mark.parametrize("det", [False, True])
def test_wide_tensor_core_path_equals_simt(e, n, det):
    """H = 128: the tcgen05 composition (b200mp_row_linear + b200mp_message_grad_rows + b200mp_node_linear_wgrad; functional.wide_tc_available)
    against the fused SIMT kernels it replaces (B200MP_TC128=0) on the same inputs: forward modes 0 / 1 / 2, both message backward kinds with
    and without an incoming gradient, the edge-update backward; then the row-mapped forms of the composition (the SIMT backward kernels take
    none) against the unmapped ones on pre-permuted inputs -- 3xTF32 against fp32 FFMA"""
    import os
    H = 128
    g = torch.Generator(device=DEV).manual_seed(e)
    rnd = lambda *s: torch.randn(*s, device=DEV, generator=g)
    ei = torch.stack([torch.randint(0, n, (e,), device=DEV, generator=g), (torch.rand(e, device=DEV, generator=g) ** 3 * n).long().clamp_(max=n - 1)])
    st = b200mp.edge_structure(ei, n)
    seg = st.fwd
    p32 = seg.perm32()
    pl = p32.long()
    E_, A, B = rnd(e, H), rnd(n, H), rnd(n, H)
    W1, b1, W2, b2 = rnd(H, H) / H ** 0.5, rnd(H), rnd(H, H) / H ** 0.5, rnd(H)
    coef = torch.cat([rnd(n, H) for _ in range(7)], 1)
    acc, gout = rnd(e, H), rnd(e, H)

    def run(ea, go, maps, saved=None):
        """every output in segment order; maps: read / write the caller's order through p32 (inputs un-permuted).  saved: the (m, h) the
        backward kernels are given -- shared between the runs, because the backward rules branch on m == min / max and h > 0, and at
        300 k edges x 128 columns a last-bit difference of the forward flips a handful of those (a whole gradient term moves to
        another edge: 2e-2 of the tensor's scale, measured; tests/test_gpu_parity_scale.py has the kink-aware statistics)"""
        eid = p32 if maps else None
        out = []
        for mode in (0, 1, 2):
            o1, o2 = BF._edge_linear_fwd(mode, ea, eid, W1, b1, B, seg.nbr32, A, seg.seg32, W2 if mode == 2 else None,
                                         b2 if mode == 2 else None, 0.5, 0.25, eid if mode == 2 else None)
            out += [o1] + ([o2[pl] if maps else o2] if mode == 2 else [])
        m, h = (out[0], out[2]) if saved is None else saved
        agg, _ = BF._segment_reduce(0, seg, m, H)          # PNA min / max rows tied to the messages: the equality branches fire
        cf = coef.clone()
        cf[:, 3 * H:4 * H] = agg[:, H:2 * H]
        cf[:, 5 * H:6 * H] = agg[:, 2 * H:3 * H]
        for kind, node_in in ((0, cf), (1, A)):
            for g_e in (None, acc):
                r = list(BF._message_bwd(kind, seg, m, ea, node_in, W1, True, n, None if g_e is None else g_e.clone(), eid))
                if maps:
                    r[0] = r[0][pl]
                out += r
        out += list(BF._edge_update_bwd(st, go, h, ea, W1, W2, 1.0, 0.5, True, True, None, None, eid, eid))
        return out

    BF.set_deterministic(det)
    try:
        os.environ["B200MP_TC128"] = "0"
        ref = run(E_[pl].contiguous(), gout[pl].contiguous(), False)
        os.environ["B200MP_TC128"] = "1"
        got = run(E_[pl].contiguous(), gout[pl].contiguous(), False, (ref[0], ref[2]))
        got_mapped = run(E_, gout, True, (ref[0], ref[2]))
    finally:
        os.environ.pop("B200MP_TC128", None)
        BF.set_deterministic(None)
    assert len(ref) == len(got) == len(got_mapped)
    # arbiter for the first message backward (PNA, no incoming gradient): the rule in fp64 torch on the composition's own m
    m = ref[0]
    cf = coef.clone()
    agg, _ = BF._segment_reduce(0, seg, m, H)
    cf[:, 3 * H:4 * H] = agg[:, H:2 * H]
    cf[:, 5 * H:6 * H] = agg[:, 2 * H:3 * H]
    c = cf[seg.seg32.long()].double().view(e, 7, H)
    md = m.double()
    G64 = c[:, 1] * (md - c[:, 6]) + c[:, 0] + (md == c[:, 3]) * c[:, 2] + (md == c[:, 5]) * c[:, 4]
    gW64 = G64.t() @ E_[pl].double()
    errs = {"arbiter g_e (wide)": rel_err(got[4], G64 @ W1.double()), "arbiter g_e (simt)": rel_err(ref[4], G64 @ W1.double()),
            "arbiter gW (wide)": rel_err(got[6], gW64), "arbiter gW (simt)": rel_err(ref[6], gW64),
            "arbiter gb (wide)": rel_err(got[7], G64.sum(0)), "arbiter gb (simt)": rel_err(ref[7], G64.sum(0))}
    # (a sum kept in tensor memory loses accuracy linearly in the number of MMAs added into one accumulator -- 1.06e-5 of max|gW| on this
    # case with 1024 rows between two drains of the weight-gradient kernel, hence its 512: profiles/r02_wgrad_chain_and_node_precision.txt)
    bad = {k: v for k, v in errs.items() if v > 1e-5 and "wide" in k}
    for i, (a, b, c_) in enumerate(zip(got, ref, got_mapped)):
        assert a.shape == b.shape == c_.shape
        if rel_err(a, b) > 2e-5:
            bad[i] = rel_err(a, b)
        if rel_err(c_, a) > 2e-5:
            bad[("mapped", i)] = rel_err(c_, a)
    assert not bad, (bad, errs)
