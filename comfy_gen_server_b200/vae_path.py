"""VAE decode, the step after the denoise path (comfy/sd.py:288-305,
comfy/ldm/modules/diffusionmodules/model.py:88-160, 244-290), as far as this library's kernels reach:

  * ResnetBlock.forward: GroupNorm -> SiLU as one kernel (cgs_groupnorm_fwd, NCHW), the block's `x + h` as the
    vectorised add kernel; the 3x3 convolutions stay on cuDNN.
  * AttnBlock.forward: its GroupNorm on the kernel, `x + h_` on the add kernel. The attention itself is a single head
    of d = 512 over h*w tokens -- beyond the flash-attention kernel's TMEM plan (O alone would need all 512 columns) --
    and stays on the reference's own VAE attention function.

Bound onto the reference's classes by integration.install(); anything but contiguous bf16 CUDA images (the
reference decodes in bf16 on this class of GPU: model_management.vae_dtype) goes to the reference's own methods.
The text encoders are left alone on purpose: the reference runs them with manual-cast operators on fp32 embeddings,
i.e. in fp32 (comfy/clip_model.py:84, comfy/sd1_clip.py:81); bf16 kernels would change their results.
"""
from __future__ import annotations

import torch

from . import _native
from . import layers
from .layers import ORIGINAL, on_fast_device


def _fast_image(x) -> bool:
    return on_fast_device(x) and x.dim() == 4 and x.is_contiguous() and (x.shape[2] * x.shape[3]) % 8 == 0


def _norm(norm, x, silu):
    """torch.nn.GroupNorm (model.py:Normalize) with the SiLU that follows it."""
    if _fast_image(x) and norm.affine and norm.weight.dtype == layers.ACT_DTYPE and x.shape[1] % norm.num_groups == 0:
        return _native.group_norm(x, norm.weight, norm.bias, norm.num_groups, norm.eps, silu)
    h = torch.nn.functional.group_norm(x, norm.num_groups, norm.weight, norm.bias, norm.eps)
    return torch.nn.functional.silu(h) if silu else h


def resnet_forward(self, x, temb=None):
    if temb is not None or not _fast_image(x):
        return ORIGINAL["vae.ResnetBlock.forward"](self, x, temb)
    h = self.conv1(_norm(self.norm1, x, True))
    h = self.conv2(self.dropout(_norm(self.norm2, h, True)))
    if self.in_channels != self.out_channels:
        x = self.conv_shortcut(x) if self.use_conv_shortcut else self.nin_shortcut(x)
    return _native.residual_add(x, h)


def attn_forward(self, x):
    if not _fast_image(x):
        return ORIGINAL["vae.AttnBlock.forward"](self, x)
    h = _norm(self.norm, x, False)
    h = self.optimized_attention(self.q(h), self.k(h), self.v(h))
    return _native.residual_add(x, self.proj_out(h))
