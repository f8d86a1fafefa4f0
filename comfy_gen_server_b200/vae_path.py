"""VAE decode, the step after the denoise path (comfy/sd.py:288-305,
comfy/ldm/modules/diffusionmodules/model.py:48-160, 244-294), on this library's kernels:

  * the decoder runs channels-last from its first ResnetBlock on (the weights are converted in place once);
  * ResnetBlock.forward: GroupNorm -> SiLU as one kernel (channels-last, 4..16 channels per group), both 3x3
    convolutions on the implicit-GEMM kernel, conv1's bias and conv2's bias + `x + h` in their epilogues;
  * Upsample.forward: interpolate + conv as four 2x2 phase convolutions of the small image (no 4x intermediate);
  * AttnBlock.forward: GroupNorm on the kernel; q / k / v as one GEMM; the attention itself -- a single head of
    d = 512 over h*w tokens, beyond the flash-attention kernel's TMEM plan -- as the reference's `normal_attention`
    algorithm (scores = q k^T in the compute dtype, row softmax, scores v) on the library GEMM and a row-softmax kernel;
    proj_out with `x + h_` in its epilogue.

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
            with torch.inference_mode(p.is_inference()):    # an inference tensor stays one (the server loads checkpoints
                # under torch.inference_mode()), a normal tensor stays usable outside inference mode
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


MAX_SCORE_TOKENS = 32768          # the [tokens, tokens] score matrix of one image: 2 GiB in bf16 at this size


def _qkv_weight(self):
    """The three 1x1 convolutions q, k, v as one [3C, C] GEMM weight (+ bias), derived once per set of weights."""
    ver = layers.tensor_version
    key = tuple((m.weight.data_ptr(), ver(m.weight), m.bias.data_ptr(), ver(m.bias)) for m in (self.q, self.k, self.v))
    cached = self.__dict__.get("_cgs_vae_qkv")
    if cached is None or cached[0] != key:
        c = self.in_channels
        w = torch.cat([m.weight.reshape(c, c) for m in (self.q, self.k, self.v)]).contiguous()
        b = torch.cat([m.bias for m in (self.q, self.k, self.v)]).contiguous()
        cached = (key, w, b)
        self.__dict__["_cgs_vae_qkv"] = cached
    return cached[1], cached[2]


def _attn_on_gemm_ok(self, x) -> bool:
    c = self.in_channels
    n = x.shape[2] * x.shape[3]
    return (c % 64 == 0 and n % 32 == 0 and n <= MAX_SCORE_TOKENS
            and all(conv_plain(m) and m.kernel_size == (1, 1) for m in (self.q, self.k, self.v, self.proj_out)))


def attn_forward(self, x):
    """AttnBlock.forward (model.py:281-294). The attention is one head of d = C = 512 over h*w tokens -- beyond the
    flash-attention kernel's TMEM plan (O alone would need all 512 columns) -- so it runs the reference's own
    `normal_attention` algorithm (model.py:160-209: scores = bmm(q, k) * scale, softmax, bmm with v, in the compute
    dtype) with both matrix products, the fused q / k / v projection and proj_out + `x + h_` on the library GEMM and
    the row softmax on its own kernel."""
    if not _fast_image(x):
        return ORIGINAL["vae.AttnBlock.forward"](self, x)
    if not _attn_on_gemm_ok(self, x):
        h = _norm(self.norm, x, False)
        h = self.optimized_attention(self.q(h), self.k(h), self.v(h))
        return _native.residual_add(x, self.proj_out(h))
    x = x.contiguous(memory_format=_CL)
    b, c, hh, ww = x.shape
    n = hh * ww
    tokens = lambda t: t.permute(0, 2, 3, 1).reshape(b, n, c)            # a view of a channels-last image
    h = _norm(self.norm, x, False)
    w_qkv, b_qkv = _qkv_weight(self)
    qkv = _native.linear(tokens(h), w_qkv, b_qkv)                         # [b, n, 3c]
    out = torch.empty_like(x)                                             # channels-last
    out_t, x_t = tokens(out), tokens(x)
    for i in range(b):
        q, k, v = qkv[i, :, :c], qkv[i, :, c:2 * c], qkv[i, :, 2 * c:]
        scores = _native.linear(q, k)                                     # [n, n] = q k^T
        _native.softmax_rows_(scores, float(c) ** -0.5)
        o = _native.linear(scores, v.t().contiguous())                    # [n, c] = softmax(..) v
        _native.linear(o, self.proj_out.weight.reshape(c, c), self.proj_out.bias, residual=x_t[i], out=out_t[i])
    return out
