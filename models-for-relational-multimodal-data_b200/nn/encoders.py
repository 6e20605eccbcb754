"""Stype-wise column encoders: the plugin surface the reference takes from the
pytorch-frame fork -- ``StypeWiseFeatureEncoder(out_channels, col_stats, col_names_dict, stype_encoder_dict)`` with
``EmbeddingEncoder()`` / ``LinearEncoder()`` / ``TimestampEncoder()`` / fork-only ``ProjectionEncoder()``
(/root/reference/src/datasets/ibm_transactions_for_aml.py:283-294,313-319; eth_transactions.py:298-304,334-340),
called as ``x, col_names = encoder(tensor_frame)`` (src/utils.py:139-140,357-359; fused.py:71-72) -- and the minimal
``TensorFrame`` those calls need (``feat_dict``, ``col_names_dict``, ``y``, ``__getitem__``, ``to``, ``num_cols``,
``get_col_feat``).  The fork is absent from the reference tree; semantics follow upstream pytorch-frame as restated
in SURVEY.md App. A.4 and oracle/stype_encoders.py (parity unpinned).

The modules hold the parameters and buffers under upstream's names (``encoder_dict.<stype>.weight|bias|mean|std|
embs.<j>.weight|min_year``); all columns of all stypes are encoded by ONE kernel launch that writes the
[rows, K, C] token tensor once (csrc/col_encoder.cu), optionally with a CLS token in slot 0.
"""
from __future__ import annotations

import enum
from typing import Any, Dict, List, Optional, Sequence

import torch
from torch import Tensor, nn

from .. import _lib as L
from .. import functional as BF


class stype(enum.Enum):
    numerical = 'numerical'
    categorical = 'categorical'
    timestamp = 'timestamp'
    relation = 'relation'          # fork-only: numeric pass-through columns (node_attr, EgoID, link ...)

    def __str__(self):
        return self.value


class StatType(enum.Enum):
    MEAN = 'MEAN'
    STD = 'STD'
    COUNT = 'COUNT'                # (categories, counts)
    YEAR_RANGE = 'YEAR_RANGE'      # [min_year, max_year]


def _stat(stats: Dict[Any, Any], key: StatType):
    for k in (key, key.value, key.name):
        if k in stats:
            return stats[k]
    raise KeyError(f"column statistics lack {key.name}")


class TensorFrame:
    """Columns of one table grouped by stype: ``feat_dict[stype]`` is [rows, n_cols] (float32 numerical / relation,
    int64 categorical) or [rows, n_cols, 7] int64 calendar fields (timestamp)."""

    def __init__(self, feat_dict: Dict[stype, Tensor], col_names_dict: Dict[stype, List[str]], y: Optional[Tensor] = None):
        for st, names in col_names_dict.items():
            if st not in feat_dict or feat_dict[st].shape[1] != len(names):
                raise ValueError(f"feat_dict[{st}] does not match its {len(names)} column names")
        rows = {int(t.shape[0]) for t in feat_dict.values()}
        if len(rows) > 1 or (y is not None and rows and int(y.shape[0]) not in rows):
            raise ValueError("all stypes (and y) must have the same number of rows")
        self.feat_dict, self.col_names_dict, self.y = dict(feat_dict), {k: list(v) for k, v in col_names_dict.items()}, y

    @property
    def stype_list(self) -> List[stype]:
        return list(self.feat_dict.keys())

    @property
    def num_rows(self) -> int:
        return int(next(iter(self.feat_dict.values())).shape[0]) if self.feat_dict else 0

    def __len__(self):
        return self.num_rows

    @property
    def num_cols(self) -> int:
        return sum(len(v) for v in self.col_names_dict.values())

    @property
    def device(self):
        return next(iter(self.feat_dict.values())).device if self.feat_dict else None

    def get_col_feat(self, name: str) -> Tensor:
        for st, names in self.col_names_dict.items():
            if name in names:
                j = names.index(name)
                return self.feat_dict[st][:, j:j + 1]
        raise KeyError(name)

    def __getitem__(self, idx) -> "TensorFrame":
        if isinstance(idx, int):
            idx = [idx]
        if isinstance(idx, (list, tuple, range)):
            idx = torch.as_tensor(list(idx), dtype=torch.long, device=self.device)
        return TensorFrame({k: v[idx] for k, v in self.feat_dict.items()}, self.col_names_dict,
                           None if self.y is None else self.y[idx.to(self.y.device) if isinstance(idx, Tensor) else idx])

    def to(self, *args, **kwargs) -> "TensorFrame":
        return TensorFrame({k: v.to(*args, **kwargs) for k, v in self.feat_dict.items()}, self.col_names_dict,
                           None if self.y is None else self.y.to(*args, **kwargs))


class StypeEncoder(nn.Module):
    """Per-stype encoder: constructed empty (as the reference does), completed by ``StypeWiseFeatureEncoder``
    which sets ``out_channels`` / ``stats_list`` / ``stype`` and calls ``post_init``."""
    supported_stypes: Sequence[stype] = ()

    def __init__(self, out_channels: Optional[int] = None, stats_list: Optional[List[Dict]] = None,
                 stype: Optional[stype] = None):
        super().__init__()
        self.out_channels, self.stats_list, self.stype = out_channels, stats_list, stype
        self._initialised = False
        if out_channels is not None and stats_list is not None:
            self.post_init()

    def post_init(self):
        if self.stype is not None and self.supported_stypes and self.stype not in self.supported_stypes:
            raise ValueError(f"{type(self).__name__} does not encode stype.{self.stype}")
        self.init_modules()
        self.reset_parameters()
        self._initialised = True

    def init_modules(self):
        raise NotImplementedError

    def reset_parameters(self):
        raise NotImplementedError

    def plan(self, feat: Tensor, params: list) -> list:
        """column entries for functional.col_encode; appends this encoder's parameters to ``params``"""
        raise NotImplementedError

    def forward(self, feat: Tensor, col_names: Optional[List[str]] = None) -> Tensor:
        params: list = []
        plan = self.plan(feat, params)
        return BF.col_encode(plan, feat.shape[0], self.out_channels, None, params)


class LinearEncoder(StypeEncoder):
    """numerical column j: ((v - mean_j) / std_j) * weight[j] + bias[j]"""
    supported_stypes = (stype.numerical,)

    def init_modules(self):
        n, C = len(self.stats_list), self.out_channels
        self.register_buffer('mean', torch.tensor([float(_stat(s, StatType.MEAN)) for s in self.stats_list], dtype=torch.float32))
        self.register_buffer('std', torch.tensor([float(_stat(s, StatType.STD)) for s in self.stats_list], dtype=torch.float32) + 1e-6)
        self.weight = nn.Parameter(torch.empty(n, C))
        self.bias = nn.Parameter(torch.empty(n, C))
        self._norm = None

    def reset_parameters(self):
        nn.init.normal_(self.weight, std=0.01)
        nn.init.zeros_(self.bias)

    def _load_from_state_dict(self, *a, **k):
        super()._load_from_state_dict(*a, **k)
        self._norm = None

    def _normalisation(self):
        if self._norm is None:           # host copies of the per-column constants (kernel arguments by value)
            self._norm = (self.mean.tolist(), self.std.tolist())
        return self._norm

    def plan(self, feat, params):
        n = feat.shape[1]
        mean, std = self._normalisation()
        w, b = len(params), len(params) + 1
        params += [self.weight, self.bias]
        return [(L.COL_NUMERICAL, feat, j, n, (w, j), (b, j), mean[j], std[j], 0.0, 0) for j in range(n)]


class ProjectionEncoder(LinearEncoder):
    """fork-only encoder of ``stype.relation`` columns (constant / id pass-through columns such as ``node_attr``,
    ibm_transactions_for_aml.py:313-319): v * weight[j] + bias[j], no normalisation (inferred from its use)."""
    supported_stypes = (stype.relation, stype.numerical)

    def init_modules(self):
        n, C = len(self.stats_list), self.out_channels
        self.register_buffer('mean', torch.zeros(n))
        self.register_buffer('std', torch.ones(n))
        self.weight = nn.Parameter(torch.empty(n, C))
        self.bias = nn.Parameter(torch.empty(n, C))
        self._norm = None


class EmbeddingEncoder(StypeEncoder):
    """categorical column j: embs[j](id + 1); row 0 of each table is the padding row of missing values (id -1)"""
    supported_stypes = (stype.categorical,)

    def init_modules(self):
        self.embs = nn.ModuleList()
        for s in self.stats_list:
            ncat = len(_stat(s, StatType.COUNT)[0])
            self.embs.append(nn.Embedding(ncat + 1, self.out_channels, padding_idx=0))

    def reset_parameters(self):
        for emb in self.embs:
            emb.reset_parameters()

    def plan(self, feat, params):
        n = feat.shape[1]
        out = []
        for j, emb in enumerate(self.embs):
            out.append((L.COL_CATEGORICAL, feat, j, n, (len(params), None), None, 0.0, 1.0, 0.0, emb.num_embeddings - 1))
            params.append(emb.weight)
        return out


class TimestampEncoder(StypeEncoder):
    """timestamp column j ([rows, 7] calendar fields): positional encoding of year - min_year and cyclic encodings
    of month / day / weekday / hour / minute / second, 8 values each, contracted with weight[j] [7, 8, C]"""
    supported_stypes = (stype.timestamp,)
    out_size = 8

    def init_modules(self):
        n, C = len(self.stats_list), self.out_channels
        self.register_buffer('min_year', torch.tensor([float(_stat(s, StatType.YEAR_RANGE)[0]) for s in self.stats_list]))
        self.register_buffer('max_values', torch.tensor([12., 31., 7., 24., 60., 60.]))
        self.weight = nn.Parameter(torch.empty(n, 7, self.out_size, C))
        self.bias = nn.Parameter(torch.empty(n, C))
        self._my = None

    def reset_parameters(self):
        nn.init.normal_(self.weight, std=0.01)
        nn.init.zeros_(self.bias)

    def _load_from_state_dict(self, *a, **k):
        super()._load_from_state_dict(*a, **k)
        self._my = None

    def plan(self, feat, params):
        if feat.dim() != 3 or feat.shape[2] != 7:
            raise ValueError("timestamp features are [rows, cols, 7] calendar fields")
        if self._my is None:
            self._my = self.min_year.tolist()
        n = feat.shape[1]
        w, b = len(params), len(params) + 1
        params += [self.weight, self.bias]
        return [(L.COL_TIMESTAMP, feat, 7 * j, 7 * n, (w, j), (b, j), 0.0, 1.0, self._my[j], 0) for j in range(n)]


class StypeWiseFeatureEncoder(nn.Module):
    """``encoder(tf) -> (x [rows, K, C], col_names)``: columns grouped by stype in ``tf.feat_dict`` order.
    ``forward(tf, cls=...)`` also writes a CLS token [C] into slot 0 (x becomes [rows, K + 1, C]) -- the concat the
    TABGNN models do next (src/nn/models/tabgnn.py:214, fused.py:149) folded into the same pass."""

    def __init__(self, out_channels: int, col_stats: Dict[str, Dict], col_names_dict: Dict[stype, List[str]],
                 stype_encoder_dict: Dict[stype, StypeEncoder]):
        super().__init__()
        self.out_channels = out_channels
        self.col_names_dict = {k: list(v) for k, v in col_names_dict.items()}
        self.encoder_dict = nn.ModuleDict()
        for st, enc in stype_encoder_dict.items():
            if st not in col_names_dict:
                continue
            if enc._initialised:
                raise ValueError("stype encoders are passed unconstructed (EmbeddingEncoder()), as the reference does")
            enc.out_channels, enc.stype = out_channels, st
            enc.stats_list = [col_stats[name] for name in col_names_dict[st]]
            enc.post_init()
            self.encoder_dict[st.value] = enc
        missing = [st for st in col_names_dict if st.value not in self.encoder_dict]
        if missing:
            raise ValueError(f"no encoder given for stypes {[str(m) for m in missing]}")

    def reset_parameters(self):
        for enc in self.encoder_dict.values():
            enc.reset_parameters()

    def forward(self, tf: TensorFrame, cls: Optional[Tensor] = None):
        plan, params, names = [], [], []
        for st in tf.stype_list:
            feat = tf.feat_dict[st]
            if st is stype.relation and feat.dtype != torch.float32:
                feat = feat.to(torch.float32)
            plan += self.encoder_dict[st.value].plan(feat.contiguous(), params)
            names += self.col_names_dict[st]
        x = BF.col_encode(plan, tf.num_rows, self.out_channels, cls, params)
        return x, names
