"""ORACLE (test infrastructure, never on the product path): CPU restatement of the stype-wise column
encoders the reference calls through ``StypeWiseFeatureEncoder`` (/root/reference/src/utils.py:139-140,357-359;
fused.py:71-72; encoder choice at src/datasets/ibm_transactions_for_aml.py:283-294,313-319).

PARITY UNPINNED: the arithmetic lives in the pytorch-frame fork ``Atahanak/pytorch-frame@ac933fa`` (an
un-vendored submodule, absent from /root/reference), and the reference holds no golden vectors for it.  What is
restated here is upstream pytorch-frame's published behaviour:

* ``LinearEncoder``      x[r, j, :] = ((v[r, j] - mean_j) / std_j) * W[j, :] + b[j, :]        (std_j = stat std + 1e-6)
* ``EmbeddingEncoder``   x[r, j, :] = table_j[id[r, j] + 1, :]       (table_j has |categories| + 1 rows, row 0 is the
                         padding row that a missing value, id -1, selects)
* ``TimestampEncoder``   feat [rows, cols, 7] = (year, month, day, weekday, hour, minute, second);
                         year - min_year -> PositionalEncoding(8), the other six / (12, 31, 7, 24, 60, 60) ->
                         CyclicEncoding(8); x[r, j, :] = einsum(enc[r, j] [7, 8], W[j] [7, 8, C]) + b[j]
* ``ProjectionEncoder``  (fork-only, stype.relation; inferred from its use on constant / id columns) x = v * W + b
* ``StypeWiseFeatureEncoder``  concatenates the per-stype blocks along the column axis in feat_dict order.

Plain torch on purpose: fp64 inputs give the fp64 oracle.
"""
import math

import torch

CYCLIC_MAX = (12, 31, 7, 24, 60, 60)


def positional_encoding(x: torch.Tensor, out_size: int = 8) -> torch.Tensor:
    mult = torch.pow(torch.tensor(1.0 / 10000.0, dtype=x.dtype), torch.arange(0, out_size, 2, dtype=x.dtype) / out_size)
    m = x.unsqueeze(-1) * mult
    return torch.cat([torch.sin(m), torch.cos(m)], dim=-1)


def cyclic_encoding(x: torch.Tensor, out_size: int = 8) -> torch.Tensor:
    mult = torch.arange(1, out_size // 2 + 1, dtype=x.dtype)
    m = x.unsqueeze(-1) * mult
    return torch.cat([torch.sin(m * math.pi), torch.cos(m * 2 * math.pi)], dim=-1)


def linear_encode(feat, mean, std, W, b):
    z = (feat.to(W.dtype) - mean) / std
    return torch.einsum('ij,jk->ijk', z, W) + b


def embedding_encode(feat, tables):
    return torch.stack([t[feat[:, j] + 1] for j, t in enumerate(tables)], dim=1)


def timestamp_encode(feat, min_year, W, b):
    f = feat.to(W.dtype)
    year = f[..., :1] - min_year.to(W.dtype).view(1, -1, 1)
    rest = f[..., 1:] / torch.tensor(CYCLIC_MAX, dtype=W.dtype).view(1, 1, -1)
    enc = torch.cat([positional_encoding(year), cyclic_encoding(rest)], dim=2)        # [rows, cols, 7, 8]
    return torch.einsum('ijkl,jklm->ijm', enc, W) + b


def stype_wise_encode(blocks, cls=None):
    """blocks: list of ('numerical', feat, mean, std, W, b) | ('categorical', feat, tables) |
    ('timestamp', feat, min_year, W, b) in output order; cls: optional [C] token for slot 0."""
    xs = []
    for blk in blocks:
        kind = blk[0]
        if kind == 'numerical':
            xs.append(linear_encode(*blk[1:]))
        elif kind == 'categorical':
            xs.append(embedding_encode(*blk[1:]))
        elif kind == 'timestamp':
            xs.append(timestamp_encode(*blk[1:]))
        else:
            raise ValueError(kind)
    x = torch.cat(xs, dim=1)
    if cls is not None:
        x = torch.cat([cls.view(1, 1, -1).expand(x.shape[0], 1, -1), x], dim=1)
    return x
