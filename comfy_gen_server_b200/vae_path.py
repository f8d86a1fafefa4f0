"""VAE decode, the step after the denoise path (comfy/sd.py:288-305,
comfy/ldm/modules/diffusionmodules/model.py:48-160, 244-294), on this library's kernels:

  * the decoder runs channels-last from its first ResnetBlock on (the weights are converted in place once);
  * ResnetBlock.forward: GroupNorm -> SiLU as one kernel (channels-last, 4..16 channels per group), both 3x3
    convolutions on the implicit-GEMM kernel, conv1's bias and conv2's bias + `x + h` in their epilogues;
  * Upsample.forward: interpolate + conv as four 2x2 phase convolutions of the small image (no 4x intermediate);
  * AttnBlock.forward: its GroupNorm on the kernel, `x + h_` on the add kernel. The attention itself is a single head
    of d = 512 over h*w tokens -- beyond the flash-attention kernel's TMEM plan (O alone would need all 512 columns) --
    and stays on the reference's own VAE attention function.

Bound onto the reference's classes by integration.install(); anything but bf16 CUDA images (the reference decodes
in bf16 on this class of GPU: model_management.vae_dtype) goes to the reference's own methods.
The text encoders are left alone on purpose: the reference runs them with manual-cast operators on fp32 embeddings,
i.e. in fp32 (comfy/clip_model.py:84, comfy/sd1_clip.py:81); bf16 kernels would change their results.
"""
from __future__ import annotations

import torch

from . import _native
from . import layers
from .layers import ORIGINAL, on_fast_device, conv_plain, conv3x3, upsample2x_conv3x3

_CL = torch.channels_last


def _fast_image(x) -> bool:
    return on_fast_device(x) and x.dim() == 4 and (x.shape[2] * x.shape[3]) % 8 == 0


def _channels_last(module, x):
    """x and the module's convolution weights in channels-last (weights converted in place, once)."""
    for p in module.parameters(recurse=True):
        if p.dim() == 4 and not p.is_contiguous(memory_format=_CL):
            with torch.inference_mode(False):        # the parameter must stay usable outside inference mode
                p.data = p.data.contiguous(memory_format=_CL)
    return x.contiguous(memory_format=_CL)


def _norm(norm, x, silu):
    """torch.nn.GroupNorm (model.py:Normalize) with the SiLU that follows it."""
    if (norm.affine and norm.weight.dtype == layers.ACT_DTYPE and layers.groupnorm_kernel_ok(norm, x)):
        return _native.group_norm(x, norm.weight, norm.bias, norm.num_groups, norm.eps, silu)
    h = torch.nn.functional.group_norm(x, norm.num_groups, norm.weight, norm.bias, norm.eps)
    return torch.nn.functional.silu(h) if silu else h


def resnet_forward(self, x, temb=None):
    convs = [self.conv1, self.conv2]
    if temb is not None or not _fast_image(x) or not all(conv_plain(c) for c in convs):
        return ORIGINAL["vae.ResnetBlock.forward"](self, x, temb)
    x = _channels_last(self, x)
    h = conv3x3(self.conv1, _norm(self.norm1, x, True), self.conv1.bias)
    h = self.dropout(_norm(self.norm2, h, True))
    if self.in_channels != self.out_channels:
        x = self.conv_shortcut(x) if self.use_conv_shortcut else self.nin_shortcut(x)
    return conv3x3(self.conv2, h, self.conv2.bias, x)                  # x + conv2(h) in the convolution's epilogue


def upsample_forward(self, x):
    if not (self.with_conv and _fast_image(x) and conv_plain(self.conv)):
        return ORIGINAL["vae.Upsample.forward"](self, x)
    x = _channels_last(self, x)
    y = upsample2x_conv3x3(self.conv, x, self.conv.bias)
    if y is None:
        y = self.conv(_native.upsample_nearest2x(x))
    return y


def attn_forward(self, x):
    if not _fast_image(x):
        return ORIGINAL["vae.AttnBlock.forward"](self, x)
    h = _norm(self.norm, x, False)
    h = self.optimized_attention(self.q(h), self.k(h), self.v(h))
    return _native.residual_add(x, self.proj_out(h))
