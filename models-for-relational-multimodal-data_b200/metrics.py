"""Ranking metrics of the reference's self-supervised loop (/root/reference/src/utils/metric.py:7-63, ``SSMetric``):
``mrr`` ranks every positive among its negatives with a CUDA kernel (csrc/readout.cu: one warp per positive) instead of
copying the scores to the host and running a per-sample numpy argsort loop (:21-43); ``mv_accuracy`` is one argmax."""
from __future__ import annotations

from typing import Dict, Sequence, Tuple

import torch

from . import functional as BF


class SSMetric:
    def __init__(self, num_numerical):
        self.num_numerical = num_numerical

    @staticmethod
    def ranks(pos_pred, neg_pred, num_neg_samples) -> torch.Tensor:
        """int32 [B] on the device: 1-based rank of each positive among its ``num_neg_samples`` negatives (ties: positive last)"""
        return BF.rank_positive(pos_pred, neg_pred, int(num_neg_samples))

    @staticmethod
    def mrr_on_device(pos_pred, neg_pred, ks: Sequence[int], num_neg_samples) -> Tuple[torch.Tensor, torch.Tensor]:
        """(mrr 0-d tensor, hits [len(ks)] tensor) without any device -> host transfer"""
        r = SSMetric.ranks(pos_pred, neg_pred, num_neg_samples)
        mrr = (1.0 / r.double()).mean()
        hits = torch.stack([(r <= int(k)).double().mean() for k in ks]) if len(ks) else torch.zeros(0, device=r.device, dtype=torch.float64)
        return mrr, hits

    @staticmethod
    def mrr(pos_pred, neg_pred, ks, num_neg_samples) -> Tuple[float, Dict[str, float]]:
        """Mean reciprocal rank and Hits@k with the reference's return types (metric.py:11-52): one small device read at the end"""
        mrr, hits = SSMetric.mrr_on_device(pos_pred, neg_pred, ks, num_neg_samples)
        vals = torch.cat([mrr.reshape(1), hits]).tolist()
        return vals[0], {f"hits@{k}": v for k, v in zip(ks, vals[1:])}

    def mv_accuracy(self, mv_out, y, tol=0.1):
        target = y[:, 1] if torch.is_tensor(y) else torch.as_tensor([m for _, m in y], device=mv_out.device)
        return float((mv_out.argmax(dim=1) == target.to(mv_out.device).long()).double().mean())
