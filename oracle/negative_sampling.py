"""TEST INFRASTRUCTURE (never imported by the product path): CPU restatement of the reference's
negative sampler, /root/reference/src/primitives/negative_sampling/negative_sampling.cpp:10-76, and a
checker for the properties any correct sampler output has.  The reference seeds std::mt19937 from
std::random_device, so outputs cannot be compared value by value: parity = the constraints below +
the distribution of the drawn nodes (uniform over the available nodes of each positive).

Pinned against the reference itself: oracle/build_ref.sh compiles the reference's own .cpp into
oracle/_ref/ (tests/test_negative_sampling.py checks this port and the compiled reference satisfy the
same checker on the reference's own fixture graph, tests/golden/negsamp_fixture.npz)."""
import numpy as np


def adjacency(edge_index):
    """undirected neighbour sets over the batch graph (cpp:20-31)"""
    adj = {}
    for s, d in zip(edge_index[0], edge_index[1]):
        adj.setdefault(int(s), set()).add(int(d))
        adj.setdefault(int(d), set()).add(int(s))
    return adj


def generate_negative_samples(edge_index, pos_edge_index, num_neg_samples, rng=None):
    if num_neg_samples <= 0:
        raise ValueError("num_neg_samples must be greater than 0")          # cpp:13-15
    rng = rng or np.random.default_rng(0)
    adj = adjacency(edge_index)
    n = len(set(map(int, edge_index[0])) | set(map(int, edge_index[1])))      # nodeset.size(), cpp:34-35
    srcs, dsts = [], []
    for s, d in zip(pos_edge_index[0], pos_edge_index[1]):
        s, d = int(s), int(d)
        unavail = {s, d} | adj.get(s, set()) | adj.get(d, set())              # cpp:41-45
        for part in range(2):
            for _ in range(num_neg_samples // 2):                              # cpp:59-74
                while True:
                    c = int(rng.integers(0, n))
                    if c not in unavail:
                        break
                srcs.append(s if part == 0 else c)
                dsts.append(c if part == 0 else d)
    return [srcs, dsts]


def check(edge_index, pos_edge_index, neg, num_neg_samples):
    """Raises AssertionError unless `neg` is a valid output for these inputs; returns the drawn nodes
    as an array [B, 2 * (num_neg // 2)]."""
    adj = adjacency(edge_index)
    n = len(set(map(int, edge_index[0])) | set(map(int, edge_index[1])))
    half = num_neg_samples // 2
    B = len(pos_edge_index[0])
    neg = np.asarray(neg)
    assert neg.shape == (2, B * 2 * half), neg.shape
    drawn = np.empty((B, 2 * half), dtype=np.int64)
    for i in range(B):
        s, d = int(pos_edge_index[0][i]), int(pos_edge_index[1][i])
        unavail = {s, d} | adj.get(s, set()) | adj.get(d, set())
        blk = neg[:, i * 2 * half:(i + 1) * 2 * half]
        assert (blk[0, :half] == s).all() and (blk[1, half:] == d).all(), "layout: (src, n) x half then (n, dst) x half"
        c = np.concatenate([blk[1, :half], blk[0, half:]])
        assert ((c >= 0) & (c < n)).all(), "drawn node outside [0, |nodeset|)"
        assert not (set(c.tolist()) & unavail), "drawn node is src, dst or a neighbour of them"
        drawn[i] = c
    return drawn
