"""Self-supervised losses of the reference (/root/reference/src/utils/loss.py:6-78, ``SSLoss``) on the device:
``lp_loss`` (:11-12), ``mcm_loss`` (:41-72) without the per-sample Python loop (:54-57) and without host copies inside the
step (fused.py:77-78, 289-290 move the targets to the CPU first), ``mv_loss`` (:74-78).  Same call signatures and return
structure."""
from __future__ import annotations

import torch
import torch.nn.functional as F


class SSLoss:
    def __init__(self, device, num_numerical=None):
        self.num_numerical = num_numerical
        self.device = device

    @staticmethod
    def lp_loss(pos_pred, neg_pred):
        return -torch.log(pos_pred + 1e-12).mean() - torch.log(1 - neg_pred + 1e-12).mean()

    def mcm_loss(self, cat_out, num_out, y, sync_counts: bool = True):
        """cat_out: one [B, n_classes_c] logit tensor per categorical column; num_out [B, num_numerical]; y [B, 2] =
        (value, masked column index).  Returns ``(loss, (cat_loss, t_c, acc), (num_loss, t_n))`` as the reference does.
        The per-column cross entropies are masked sums over the whole batch (no per-sample loop, no host round trip
        per sample); ``t_c`` / ``t_n`` are Python ints after ONE 2-element device read, or 0-d tensors with
        ``sync_counts=False`` (fully asynchronous step)."""
        dev = num_out.device if num_out is not None else cat_out[0].device
        y = y.to(dev)
        y_val, y_idx = y[:, 0], y[:, 1].long()
        nn_ = int(self.num_numerical)
        cat_mask = y_idx >= nn_
        num_mask = ~cat_mask
        cat_loss = torch.zeros((), device=dev)
        acc = torch.zeros((), device=dev)
        col = y_idx - nn_
        tgt = y_val.long()
        for c, logits in enumerate(cat_out):                       # columns (a handful), not samples
            m = cat_mask & (col == c)
            t = torch.where(m, tgt, torch.zeros_like(tgt)).clamp_(0, logits.shape[1] - 1)
            ce = F.cross_entropy(logits, t, reduction="none")
            cat_loss = cat_loss + torch.where(m, ce, torch.zeros_like(ce)).sum()
            acc = acc + (m & (logits.argmax(dim=1) == t)).sum()
        if num_out is not None and num_out.shape[1] > 0:
            pred = num_out.gather(1, y_idx.clamp(0, num_out.shape[1] - 1).unsqueeze(1)).squeeze(1)
            num_loss = torch.where(num_mask, (pred - y_val) ** 2, torch.zeros_like(pred)).sum()
        else:
            num_loss = torch.zeros((), device=dev)
        counts = torch.stack([cat_mask.sum(), num_mask.sum()])
        t_c, t_n = counts[0], counts[1]
        # branch-free total: each term appears only when its count is positive (loss.py:67-72)
        cat_term = torch.where(t_c > 0, cat_loss / t_c.clamp(min=1), torch.zeros_like(cat_loss))
        num_term = torch.where(t_n > 0, torch.sqrt(num_loss / t_n.clamp(min=1)), torch.zeros_like(num_loss))
        loss = cat_term + num_term
        if sync_counts:
            t_c, t_n = (int(v) for v in counts.tolist())
        return loss, (cat_loss, t_c, acc), (num_loss, t_n)

    def mv_loss(self, mv_out, y):
        return F.cross_entropy(mv_out, y[:, 1].long())
