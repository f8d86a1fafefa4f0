"""Conditioning containers with the reference's interface (comfy/conds.py:9-78)."""
from __future__ import annotations

import math

import torch


def repeat_to_batch_size(tensor, batch_size):
    """comfy/utils.py:252-257."""
    if tensor.shape[0] > batch_size:
        return tensor[:batch_size]
    if tensor.shape[0] < batch_size:
        return tensor.repeat([math.ceil(batch_size / tensor.shape[0])] + [1] * (len(tensor.shape) - 1))[:batch_size]
    return tensor


class CONDRegular:
    def __init__(self, cond):
        self.cond = cond

    def _copy_with(self, cond):
        return self.__class__(cond)

    def process_cond(self, batch_size, device, **kwargs):
        return self._copy_with(repeat_to_batch_size(self.cond, batch_size).to(device))

    def can_concat(self, other):
        return self.cond.shape == other.cond.shape

    def concat(self, others):
        return torch.cat([self.cond] + [x.cond for x in others])


class CONDCrossAttn(CONDRegular):
    """Text context; mismatched token counts are repeat-padded to their LCM (77 vs 154 -> 154)."""
    def can_concat(self, other):
        s1, s2 = self.cond.shape, other.cond.shape
        if s1 != s2:
            if s1[0] != s2[0] or s1[2] != s2[2]:
                return False
            mult_min = abs(s1[1] * s2[1]) // math.gcd(s1[1], s2[1])
            if mult_min // min(s1[1], s2[1]) > 4:
                return False
        return True

    def concat(self, others):
        conds = [self.cond] + [x.cond for x in others]
        max_len = self.cond.shape[1]
        for c in conds[1:]:
            max_len = abs(max_len * c.shape[1]) // math.gcd(max_len, c.shape[1])
        out = []
        for c in conds:
            if c.shape[1] < max_len:
                c = c.repeat(1, max_len // c.shape[1], 1)
            out.append(c)
        return torch.cat(out)


class CONDConstant(CONDRegular):
    def process_cond(self, batch_size, device, **kwargs):
        return self._copy_with(self.cond)

    def can_concat(self, other):
        return self.cond == other.cond

    def concat(self, others):
        return self.cond
