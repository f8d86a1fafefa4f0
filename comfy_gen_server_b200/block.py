"""Fast paths of BasicTransformerBlock._forward and SpatialTransformer.forward (comfy/ldm/modules/attention.py:462-574,
619-639), as functions bound onto the reference's classes by integration.install().

A block takes the fast path only when nothing can observe the difference: no `patches` / `patches_replace` in
transformer_options, plain resident bf16 weights, a contiguous bf16 CUDA token tensor. Otherwise the call is handed
to the reference's own `_forward` -- hook for hook the reference's code, which still reaches the kernels through the
rebound attention function, LayerNorm and GEGLU.

What the fast path changes (results equal up to bf16 rounding order):
  * the in-place residual add `x += n` and the LayerNorm that follows it are one kernel (cgs_add_layernorm_fwd);
  * self-attention q/k/v come from ONE [3C, C] projection whose column slices feed the attention kernel through
    strided tensor maps (no copies);
  * cross-attention K/V depend only on the text context: computed once per context tensor, not once per step;
  * ff.net.2 runs on the tcgen05 GEMM with bias + the block's residual add in its epilogue, written in place;
  * image <-> token rearranges are views for channels-last activations, SMEM-tile transposes otherwise, and the
    SpatialTransformer's `+ x_in` rides along.
Per-module derived state (the fused QKV weight, the K/V cache) lives in attributes prefixed `_cgs_` on the
reference's module instances.
"""
from __future__ import annotations

import torch

from . import _native
from . import layers
from .layers import ORIGINAL, on_fast_device, linear_plain, groupnorm_apply


_ver = layers.tensor_version


# ----------------------------------------------------------------------------------------------
# CrossAttention helpers
# ----------------------------------------------------------------------------------------------
def _attn_plain(attn) -> bool:
    return all(linear_plain(m) for m in (attn.to_q, attn.to_k, attn.to_v))


def _qkv_weight(attn):
    ws = (attn.to_q.weight, attn.to_k.weight, attn.to_v.weight)
    key = tuple((w.data_ptr(), _ver(w)) for w in ws)
    if attn.__dict__.get("_cgs_qkv_key") != key:
        attn.__dict__["_cgs_qkv_w"] = torch.cat([w.detach() for w in ws], dim=0)
        attn.__dict__["_cgs_qkv_key"] = key
    return attn.__dict__["_cgs_qkv_w"]


# Which Linear layers of the fast path run on the tcgen05 GEMM of this library and which stay on cuBLASLt is decided by
# measurement (profiles/r02_kernel_bench.json, profiles/r02_linear_routing.json; switches in _native.FEATURES):
#   * to_out and ff.net.2 take the library GEMM because its epilogue fuses bias AND the block's residual add (`x += n`
#     leaves the LayerNorm that follows); stand-alone the GEMM is 8-12 % behind cuBLASLt on these shapes, in the step
#     the fusion wins 1 ms of 112;
#   * the fused QKV projection is 2-7 % ahead of cuBLASLt stand-alone (CTA pairs) but not inside the power-capped
#     step: it stays on cuBLASLt, as do to_q and the K / V projections of the text context (once per schedule);
#   * below ~8 k rows (one or two images per GPU) there are too few tiles for 148 SMs and cuBLASLt's split-K kernels
#     win by up to 2x: everything stays on cuBLASLt there.
NATIVE_LINEAR_MIN_ROWS = 8192


def _rows(x) -> int:
    return x.numel() // x.shape[-1]


def _gemm_shape_ok(weight) -> bool:
    return weight.shape[0] % 32 == 0 and weight.shape[1] % 8 == 0


def project(x, weight, bias=None, prefer_native=False):
    """x W^T (+ bias): the library GEMM where it is the faster one (see above), cuBLASLt otherwise."""
    if prefer_native and _rows(x) >= NATIVE_LINEAR_MIN_ROWS and weight.shape[1] >= 1024 and _gemm_shape_ok(weight) \
            and _native.FEATURES["native_qkv"]:
        return _native.linear(x, weight, bias)
    return torch.nn.functional.linear(x, weight, bias)


def project_add_(x, a, lin, feature="native_to_out"):
    """x += lin(a) in place (the block's residual update, attention.py:527, 567, 572). With enough rows the bias and
    the residual ride in the GEMM's epilogue; returns True then, False when the caller still has to add."""
    if _rows(a) >= NATIVE_LINEAR_MIN_ROWS and linear_plain(lin) and _gemm_shape_ok(lin.weight) and x.is_contiguous() \
            and _native.FEATURES[feature]:
        _native.linear(a, lin.weight, lin.bias, residual=x, out=x)
        return True
    return False


def self_attention_fused(attn, x):
    """attention(x Wq^T, x Wk^T, x Wv^T) with one projection GEMM (before to_out)."""
    import comfy.ldm.modules.attention as ref
    inner = attn.heads * attn.dim_head
    qkv = project(x, _qkv_weight(attn), None, prefer_native=True)
    return ref.optimized_attention(qkv[..., :inner], qkv[..., inner:2 * inner], qkv[..., 2 * inner:], attn.heads)


def _kv_key(attn, context):
    return (context.data_ptr(), _ver(context), tuple(context.shape), attn.to_k.weight.data_ptr(),
            _ver(attn.to_k.weight), attn.to_v.weight.data_ptr(), _ver(attn.to_v.weight))


def set_kv_cache(attn, context, k, v):
    """Install K/V computed by the caller for `context` (the step-graph code keeps them at fixed addresses)."""
    attn.__dict__["_cgs_kv"] = (_kv_key(attn, context), k, v, context)      # holds `context`: its address stays taken


def get_kv_cache(attn):
    return attn.__dict__.get("_cgs_kv")


def cross_attention_cached(attn, x, context):
    """attention(x Wq^T, K(context), V(context)) with K / V cached per context tensor (before to_out)."""
    import comfy.ldm.modules.attention as ref
    key = _kv_key(attn, context)
    kv = attn.__dict__.get("_cgs_kv")
    if kv is None or kv[0] != key:
        kv = attn.__dict__["_cgs_kv"] = (key, attn.to_k(context), attn.to_v(context), context)
    return ref.optimized_attention(attn.to_q(x), kv[1], kv[2], attn.heads)


def derived_state(module: torch.nn.Module) -> list:
    """Every derived object this package keeps on the sub-modules of `module` (fused QKV weights, cached K / V,
    pre-summed convolution weights). A captured step graph holds on to this list: the graph has the tensors' addresses
    baked in, while the modules replace them whenever the weights they were derived from move (an unload / reload
    through a LoRA clone frees them) -- and a later reload can put every parameter back at the address the graph's
    key remembers."""
    return [v for m in module.modules() for k, v in m.__dict__.items() if k.startswith("_cgs_")]


def release_caches(module: torch.nn.Module) -> None:
    """Drop every derived tensor this package attached to the modules of `module` (fused QKV weights, K/V caches)."""
    for m in module.modules():
        for k in [k for k in m.__dict__ if k.startswith("_cgs_")]:
            del m.__dict__[k]


# ----------------------------------------------------------------------------------------------
# BasicTransformerBlock._forward
# ----------------------------------------------------------------------------------------------
def _block_fast_ok(blk, x, context, transformer_options) -> bool:
    if transformer_options.get("patches") or transformer_options.get("patches_replace"):
        return False
    if blk.ff_in or blk.disable_self_attn or blk.attn2 is None or blk.switch_temporal_ca_to_sa or not blk.is_res:
        return False
    if not (on_fast_device(x) and x.dim() == 3 and x.is_contiguous()):
        return False
    for n in (blk.norm1, blk.norm2, blk.norm3):
        if n.comfy_cast_weights or n.weight is None or n.weight.dtype != layers.ACT_DTYPE:
            return False
    if not (_attn_plain(blk.attn1) and _attn_plain(blk.attn2) and blk.attn1.to_k.weight.shape == blk.attn1.to_q.weight.shape
            and linear_plain(blk.attn1.to_out[0]) and linear_plain(blk.attn2.to_out[0])):
        return False
    return context is not None and on_fast_device(context)


def _attend_add_norm(x, a, to_out, norm):
    """x += to_out(a); return norm(x). Either the GEMM's epilogue does the add and a plain LayerNorm follows, or
    cuBLASLt projects and the add rides in the LayerNorm kernel (cgs_add_layernorm_fwd)."""
    lin = to_out[0]
    if project_add_(x, a, lin):
        return norm(x)
    return _native.add_layer_norm_(x, to_out(a), norm.weight, norm.bias, norm.eps)


def transformer_block_forward(self, x, context=None, transformer_options={}):
    if not _block_fast_ok(self, x, context, transformer_options):
        return ORIGINAL["BasicTransformerBlock._forward"](self, x, context, transformer_options)
    n = self.norm1(x)
    a = self_attention_fused(self.attn1, n)
    n = _attend_add_norm(x, a, self.attn1.to_out, self.norm2)                                       # x += attn1(...)
    a = cross_attention_cached(self.attn2, n, context)
    n = _attend_add_norm(x, a, self.attn2.to_out, self.norm3)                                       # x += attn2(...)
    ff_in, ff_out = self.ff.net[0], self.ff.net[2]
    h = ff_in(n)
    if hasattr(ff_in, "proj") and project_add_(x, h, ff_out, "native_ff_out"):
        return x                                     # x += ff.net.2(geglu(n)): bias and residual in the GEMM epilogue
    h = ff_out(self.ff.net[1](h))
    h += x
    return h


# ----------------------------------------------------------------------------------------------
# SpatialTransformer.forward
# ----------------------------------------------------------------------------------------------
def _dense(x) -> bool:
    return on_fast_device(x) and x.is_contiguous()


def is_channels_last(x) -> bool:
    return x.dim() == 4 and not x.is_contiguous() and x.is_contiguous(memory_format=torch.channels_last)


def _to_tokens(x):
    """'b c h w -> b (h w) c' (attention.py:628). Free for channels-last activations: NHWC *is* token-major."""
    b, c, h, w = x.shape
    if is_channels_last(x):
        return x.permute(0, 2, 3, 1).reshape(b, h * w, c)
    if _dense(x) and c % 8 == 0 and (h * w) % 8 == 0:
        return _native.transpose_add(x.view(b, c, h * w))
    return x.permute(0, 2, 3, 1).reshape(b, h * w, c).contiguous()


def _to_image(x, h, w, residual):
    """'b (h w) c -> b c h w' (attention.py:636), plus the `+ x_in` of attention.py:639 when `residual` is given."""
    b, n, c = x.shape
    if residual is not None and is_channels_last(residual) and x.is_contiguous():
        # channels-last pipeline: the token tensor already is the NHWC image; the add stays a plain vectorised add
        return _native.residual_add(x.view(b, h, w, c).permute(0, 3, 1, 2), residual)
    if _dense(x) and c % 8 == 0 and n % 8 == 0 and (residual is None or _dense(residual)):
        return _native.transpose_add(x, residual).view(b, c, h, w)
    x = x.reshape(b, h, w, c).permute(0, 3, 1, 2).contiguous()
    return x if residual is None else x + residual


def spatial_transformer_forward(self, x, context=None, transformer_options={}):
    if not (on_fast_device(x) and x.dim() == 4):
        return ORIGINAL["SpatialTransformer.forward"](self, x, context, transformer_options)
    if not isinstance(context, list):
        context = [context] * len(self.transformer_blocks)
    b, c, h, w = x.shape
    x_in = x
    x = groupnorm_apply(self.norm, x)
    if not self.use_linear:
        x = self.proj_in(x)
    x = _to_tokens(x)
    if self.use_linear:
        x = self.proj_in(x)
    for i, block in enumerate(self.transformer_blocks):
        transformer_options["block_index"] = i
        x = block(x, context=context[i], transformer_options=transformer_options)
    if self.use_linear:
        x = self.proj_out(x)
        return _to_image(x, h, w, x_in)                     # layout change and `+ x_in` in one kernel
    if is_channels_last(x_in) and x.is_contiguous():
        x = x.view(b, h, w, -1).permute(0, 3, 1, 2)          # NHWC view, no copy
    else:
        x = _to_image(x, h, w, None)
    x = self.proj_out(x)
    return _native.residual_add(x, x_in)
