"""Debug helper (GPU): per-op error table of the kernels vs the emulation at a multi-tile size."""
import sys, torch
sys.path.insert(0, ".")
import b200mp
from b200mp import functional as BF
from tests import emulator as EM
from tests.test_gpu_ops import hub_graph, cpu_struct
from tests.util import rel_err
DEV = "cuda"
for H, n, e in [(64, 1500, 24000), (128, 1500, 24000), (32, 1500, 60000)]:
    ei = hub_graph(n, e, 7 + H, 1400).to(DEV)
    st = b200mp.edge_structure(ei, n, reverse=True); stc = cpu_struct(st)
    g = torch.Generator().manual_seed(H); rnd = lambda *s: torch.randn(*s, generator=g)
    E_, A, B = rnd(e, H), rnd(n, H), rnd(n, H)
    W1, b1, W2, b2 = rnd(H, H) / H ** 0.5, rnd(H), rnd(H, H) / H ** 0.5, rnd(H)
    d = lambda t: t.to(DEV)
    seg, segc = st.fwd, stc.fwd
    m = BF._edge_linear_fwd(0, d(E_), None, d(W1), d(b1), d(B), seg.nbr32, d(A), seg.seg32)[0]
    ragg, rtie = EM.segment_reduce(0, segc, m.cpu(), H)
    coef = torch.cat([rnd(n, H), rnd(n, H), rnd(n, H), ragg[:, H:2*H], rnd(n, H), ragg[:, 2*H:3*H], ragg[:, :H]], 1)
    for kind, node_in in ((0, coef), (1, rnd(n, H))):
        got = BF._message_bwd(kind, seg, m, d(E_), d(node_in), d(W1), True, n)
        ref = EM.message_bwd(kind, segc, m.cpu(), E_, node_in, W1, True, n)
        print(H, "msg_bwd kind", kind, [(nm, "%.1e" % rel_err(a, b)) for a, b, nm in zip(got, ref, ("g_e", "gT1", "gW", "gb"))])
    h = BF._edge_linear_fwd(2, d(E_), None, d(W1), d(b1), d(B), st.src32, d(A), st.dst32, d(W2), d(b2), 1.0, 0.5)[0]
    gout = rnd(e, H)
    got = BF._edge_update_bwd(st, d(gout), h, d(E_), d(W1), d(W2), 1.0, 0.5, True, True)
    ref = EM.edge_update_bwd(stc, gout, h.cpu(), E_, W1, W2, 1.0, 0.5, True, True)
    print(H, "upd_bwd", [(nm, "%.1e" % rel_err(a, b)) for a, b, nm in zip(got, ref, ("g_e", "gP", "gQ", "gW1", "gb1", "gW2", "gb2"))])
    bad = (got[2].cpu() - ref[2]).abs().max(1).values
    print("   gQ worst rows", bad.argsort(descending=True)[:5].tolist(), bad.max().item(), "deg", (seg.ptr[1:] - seg.ptr[:-1])[bad.argmax()].item())
