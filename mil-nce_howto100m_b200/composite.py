"""Drop-ins for the reference's composite soft-DTW losses (reference loss.py:20-134), the only in-repo
users of ``SoftDTW``.

Same constructors (``args`` with ``.rank``), same ``forward`` signatures and values; the soft-DTW calls go
through :class:`SoftDTW` of this package, and the operand copies the reference makes before them
(``expand(b, b, n, d).reshape(-1, n, d)`` for all b*b pairs, ``repeat(8, 1, 1)``) are replaced by index
lists (``SoftDTW.forward_pairs``).  The small elementwise parts around the soft-DTW calls stay in torch.
"""
from __future__ import annotations

import torch

from .soft_dtw_cuda import SoftDTW


class CDTW(torch.nn.Module):
    """reference loss.py:20-32: the rank's own pair against its video paired with all 8 texts."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.sdtw = SoftDTW(use_cuda=True, gamma=1e-5, dist_func="cosine")

    def forward(self, video_embd, text_embd):
        device = self.args.rank
        nt = text_embd.shape[0]     # the reference hard-codes 8 = its batch size (repeat(8, 1, 1), :29)
        idx = torch.full((1,), device, dtype=torch.int32, device=video_embd.device)
        pos = self.sdtw.forward_pairs(video_embd, text_embd, idx, idx)
        neg = self.sdtw.forward_pairs(video_embd, text_embd, idx.expand(nt).contiguous(), None)
        return (pos - torch.logsumexp(neg, 0)).unsqueeze(0)


class SDTW_CIDM(torch.nn.Module):
    """reference loss.py:34-68: soft-DTW plus the two intra-sequence contrastive terms."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.sdtw = SoftDTW(use_cuda=True, gamma=1e-1, dist_func="cosine")

    @staticmethod
    def _cosine_dist_func(x, y):
        # 1 - cos for all frame pairs of a sequence with itself: one batched matmul instead of the
        # (B, n, m, d) broadcast of the reference (:40-47); same per-norm clamp as cosine_similarity
        xn = x / x.norm(dim=2, keepdim=True).clamp_min(1e-8)
        yn = y / y.norm(dim=2, keepdim=True).clamp_min(1e-8)
        return 1 - torch.bmm(xn, yn.transpose(1, 2))

    @staticmethod
    def _pairwise_dist_func(x, y):
        return torch.abs(x.unsqueeze(2) - y.unsqueeze(1))

    def forward(self, video_embd, text_embd, start, end):
        lam = 1.
        sigma = 10
        distance = self._pairwise_dist_func(start, start)
        y = torch.where(distance > sigma, 1., 0.)
        w_ = distance + 1
        w = 1 / w_
        D_x = self._cosine_dist_func(video_embd, video_embd)
        D_y = self._cosine_dist_func(text_embd, text_embd)
        relu = torch.nn.functional.relu
        I_x = (y * w_ * relu(lam - D_x) + (1 - y) * w * D_x).sum(1).sum(1)
        I_y = (y * w_ * relu(lam - D_y) + (1 - y) * w * D_y).sum(1).sum(1)
        dtw = self.sdtw(video_embd, text_embd)
        return torch.mean(I_x + I_y + dtw)


class SDTW_negative(torch.nn.Module):
    """reference loss.py:70-91: soft-DTW of the matching pairs plus, per video sequence, the sum of
    exp(<frame, word>) over all its frames and ALL text frames - with the entries against its own text
    sequence set to 0 before the exp (so each contributes exp(0) = 1, as in the reference, :84) - over b - 1.

    The reference hard-codes b = 160 sequences of L = 8 frames and d = 512 (chunk(160), the 1288 * i + j mask,
    view(160, 8), / 159); here b, L and d come from the inputs, which is the same computation at that size.
    The chunk / cat / mask / chunk / cat shuffle of :81-86 is a (b, L, b, L) view with its diagonal blocks zeroed."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.sdtw = SoftDTW(use_cuda=True, gamma=1e-1, dist_func="cosine")
        self.device = self.args.rank

    def forward(self, video_embd, text_embd):
        b, L, d = video_embd.shape
        sdtw_loss = self.sdtw(video_embd, text_embd)
        pairwise = torch.matmul(video_embd.reshape(-1, d), text_embd.reshape(-1, d).t()).view(b, L, b, -1)
        own = torch.eye(b, dtype=torch.bool, device=pairwise.device).view(b, 1, b, 1)
        pairwise = pairwise.masked_fill(own, 0.)
        negative_loss = torch.exp(pairwise).sum(dim=(1, 2, 3))
        return torch.mean(sdtw_loss + negative_loss / (b - 1))


class SDTW_3(torch.nn.Module):
    """reference loss.py:93-134: three contrastive terms, each over all b*b sequence pairs."""

    def __init__(self, args):
        super().__init__()
        self.args = args
        self.sdtw = SoftDTW(use_cuda=True, gamma=1e-1, dist_func="negative_dot")
        self.device = self.args.rank
        self._idx = {}

    def _pairs(self, b, device):
        key = (b, str(device))
        if key not in self._idx:
            p = torch.arange(b * b, device=device, dtype=torch.int32)
            # pair p = (row operand p % b, column operand p // b): what expand(b, b, ...).reshape gives for
            # unsqueeze(0) and unsqueeze(1) (:103-104)
            self._idx[key] = ((p % b).contiguous(), (p // b).contiguous())
        return self._idx[key]

    def _contrast(self, a, c):
        b = a.shape[0]
        pos = -self.sdtw(a, c)
        ix, iy = self._pairs(b, a.device)
        neg = -self.sdtw.forward_pairs(a, c, ix, iy).reshape(b, b)
        neg = torch.logsumexp(neg, 1)
        return torch.mean(neg - pos)

    def video_video(self, video_embd):
        return self._contrast(video_embd, video_embd)

    def video_text(self, video_embd, text_embd):
        return self._contrast(video_embd, text_embd)

    def text_text(self, text_embd):
        return self._contrast(text_embd, text_embd)

    def forward(self, video_embd, text_embd):
        return self.video_video(video_embd), self.video_text(video_embd, text_embd), self.text_text(text_embd)
