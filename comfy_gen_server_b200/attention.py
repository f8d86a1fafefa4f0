"""SpatialTransformer blocks on the B200 kernels, with the reference's module interface
(comfy/ldm/modules/attention.py): same class names, constructor arguments, state-dict keys,
`optimized_attention(q, k, v, heads, mask=None)` signature and transformer_options hook points.

Two execution paths inside BasicTransformerBlock._forward:
  * patched path -- any `patches` / `patches_replace` in transformer_options: the reference's
    composition, hook for hook (attention.py:462-574); it still lands on the new attention,
    LayerNorm and GEGLU kernels through the modules below.
  * fast path -- no patches: residual add + LayerNorm fused (cgs_add_layernorm_fwd), q/k/v of the
    self-attention from one fused projection whose column slices feed the attention kernel
    without a copy, cross-attention K/V cached per context tensor (they depend only on the text
    embedding, so they are computed once per schedule instead of once per step).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from . import _native
from .ops import b200_ops

ops = b200_ops


def exists(val):
    return val is not None


def default(val, d):
    return val if exists(val) else d


def _ver(t) -> int:
    """Version counter of a tensor; inference tensors do not track one (treated as immutable)."""
    try:
        return t._version
    except RuntimeError:
        return -1


# ----------------------------------------------------------------------------------------------
# attention function (module attribute, resolved by global name at call time like the reference)
# ----------------------------------------------------------------------------------------------
def attention_b200(q, k, v, heads, mask=None):
    """Drop-in for attention_pytorch (attention.py:337-349): q [B, Nq, heads*d], k/v [B, Nk, heads*d]
    -> [B, Nq, heads*d]; inputs are not modified. Runs cgs_attention_fwd (cgs_attention_masked_fwd when a mask is
    given: boolean keep-mask or additive float, attention.py:111-123); no SDPA, no fallback."""
    if q.dtype != torch.bfloat16:
        raise RuntimeError(f"attention_b200: bf16 CUDA tensors required, got {q.dtype} on {q.device}")
    return _native.attention(q, k, v, heads, mask)


optimized_attention = attention_b200
optimized_attention_masked = attention_b200


def optimized_attention_for_device(device, mask=False, small_input=False):
    """attention.py:370-383. There is one implementation here; CPU tensors are an error."""
    if torch.device(device).type != "cuda":
        raise RuntimeError("optimized_attention_for_device: the B200 path has no CPU implementation")
    return optimized_attention_masked if mask else optimized_attention


# ----------------------------------------------------------------------------------------------
# feed-forward
# ----------------------------------------------------------------------------------------------
class GEGLU(nn.Module):
    """attention.py:56-63. forward = one tcgen05 GEMM with the gate in its epilogue."""
    def __init__(self, dim_in, dim_out, dtype=None, device=None, operations=ops):
        super().__init__()
        self.proj = operations.Linear(dim_in, dim_out * 2, dtype=dtype, device=device)

    def forward(self, x):
        proj = self.proj
        if getattr(proj, "comfy_cast_weights", False):
            from .ops import cast_bias_weight
            weight, bias = cast_bias_weight(proj, x)
        else:
            weight, bias = proj.weight, proj.bias
        return _native.geglu(x, weight, bias)


class FeedForward(nn.Module):
    """attention.py:66-83."""
    def __init__(self, dim, dim_out=None, mult=4, glu=False, dropout=0., dtype=None, device=None, operations=ops):
        super().__init__()
        inner_dim = int(dim * mult)
        dim_out = default(dim_out, dim)
        project_in = nn.Sequential(
            operations.Linear(dim, inner_dim, dtype=dtype, device=device), nn.GELU()
        ) if not glu else GEGLU(dim, inner_dim, dtype=dtype, device=device, operations=operations)
        self.net = nn.Sequential(project_in, nn.Dropout(dropout),
                                 operations.Linear(inner_dim, dim_out, dtype=dtype, device=device))

    def forward(self, x):
        return self.net(x)


# ----------------------------------------------------------------------------------------------
# attention module
# ----------------------------------------------------------------------------------------------
class CrossAttention(nn.Module):
    """attention.py:386-415. to_q/to_k/to_v have no bias, to_out[0] has one."""
    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0., dtype=None, device=None,
                 operations=ops):
        super().__init__()
        inner_dim = dim_head * heads
        context_dim = default(context_dim, query_dim)
        self.heads = heads
        self.dim_head = dim_head
        self.to_q = operations.Linear(query_dim, inner_dim, bias=False, dtype=dtype, device=device)
        self.to_k = operations.Linear(context_dim, inner_dim, bias=False, dtype=dtype, device=device)
        self.to_v = operations.Linear(context_dim, inner_dim, bias=False, dtype=dtype, device=device)
        self.to_out = nn.Sequential(operations.Linear(inner_dim, query_dim, dtype=dtype, device=device),
                                    nn.Dropout(dropout))
        self._fused_w = None       # cached [to_q; to_k; to_v] weight for the self-attention fast path
        self._fused_key = None
        self._kv_cache = None      # (key, k, v) for cross-attention on an unchanged context tensor

    def forward(self, x, context=None, value=None, mask=None):
        q = self.to_q(x)
        context = default(context, x)
        k = self.to_k(context)
        if value is not None:
            v = self.to_v(value)
            del value
        else:
            v = self.to_v(context)
        if mask is None:
            out = optimized_attention(q, k, v, self.heads)
        else:
            out = optimized_attention_masked(q, k, v, self.heads, mask)
        return self.to_out(out)

    # -- fast path helpers ---------------------------------------------------------------------
    def _plain_weights(self) -> bool:
        for m in (self.to_q, self.to_k, self.to_v):
            if m.comfy_cast_weights or m.weight_function is not None or m.weight.dtype != torch.bfloat16:
                return False
        return True

    def _qkv_weight(self):
        ws = (self.to_q.weight, self.to_k.weight, self.to_v.weight)
        key = tuple((w.data_ptr(), _ver(w)) for w in ws)
        if self._fused_key != key:
            self._fused_w = torch.cat([w.detach() for w in ws], dim=0)
            self._fused_key = key
        return self._fused_w

    def forward_self_fused(self, x):
        """Self-attention: one [3C, C] projection; q/k/v are column slices (no copies)."""
        if not self._plain_weights() or self.to_k.weight.shape != self.to_q.weight.shape:
            return self.forward(x)
        inner = self.heads * self.dim_head
        qkv = torch.nn.functional.linear(x, self._qkv_weight())
        out = optimized_attention(qkv[..., :inner], qkv[..., inner:2 * inner], qkv[..., 2 * inner:], self.heads)
        return self.to_out(out)

    def _kv_key(self, context):
        return (context.data_ptr(), _ver(context), tuple(context.shape), self.to_k.weight.data_ptr(),
                _ver(self.to_k.weight), self.to_v.weight.data_ptr(), _ver(self.to_v.weight))

    def set_kv_cache(self, context, k, v):
        """Install K/V computed by the caller for `context` (the step-graph code keeps them at fixed addresses)."""
        self._kv_cache = (self._kv_key(context), k, v, context)

    def forward_cross_cached(self, x, context):
        """Cross-attention: K/V depend only on the context; reuse them while it is unchanged."""
        if not self._plain_weights():
            return self.forward(x, context)
        key = self._kv_key(context)
        if self._kv_cache is None or self._kv_cache[0] != key:
            self._kv_cache = (key, self.to_k(context), self.to_v(context), context)  # keeps context alive
        _, k, v, _ = self._kv_cache
        out = optimized_attention(self.to_q(x), k, v, self.heads)
        return self.to_out(out)


# ----------------------------------------------------------------------------------------------
# transformer block
# ----------------------------------------------------------------------------------------------
class BasicTransformerBlock(nn.Module):
    """attention.py:418-574 (ff_in / temporal variants are video-only and out of scope)."""
    def __init__(self, dim, n_heads, d_head, dropout=0., context_dim=None, gated_ff=True, checkpoint=True,
                 ff_in=False, inner_dim=None, disable_self_attn=False, disable_temporal_crossattention=False,
                 switch_temporal_ca_to_sa=False, dtype=None, device=None, operations=ops):
        super().__init__()
        if ff_in or inner_dim is not None or disable_temporal_crossattention or switch_temporal_ca_to_sa:
            raise NotImplementedError("video transformer options (ff_in / temporal) are out of scope")
        inner_dim = dim
        self.is_res = True
        self.ff_in = False
        self.disable_self_attn = disable_self_attn
        self.attn1 = CrossAttention(query_dim=inner_dim, heads=n_heads, dim_head=d_head, dropout=dropout,
                                    context_dim=context_dim if self.disable_self_attn else None, dtype=dtype,
                                    device=device, operations=operations)
        self.ff = FeedForward(inner_dim, dim_out=dim, dropout=dropout, glu=gated_ff, dtype=dtype, device=device,
                              operations=operations)
        self.attn2 = CrossAttention(query_dim=inner_dim, context_dim=context_dim, heads=n_heads, dim_head=d_head,
                                    dropout=dropout, dtype=dtype, device=device, operations=operations)
        self.norm2 = operations.LayerNorm(inner_dim, dtype=dtype, device=device)
        self.norm1 = operations.LayerNorm(inner_dim, dtype=dtype, device=device)
        self.norm3 = operations.LayerNorm(inner_dim, dtype=dtype, device=device)
        self.checkpoint = checkpoint
        self.n_heads = n_heads
        self.d_head = d_head
        self.switch_temporal_ca_to_sa = False

    def forward(self, x, context=None, transformer_options={}):
        return self._forward(x, context, transformer_options)

    def _fast_path_ok(self, x, transformer_options) -> bool:
        if transformer_options.get("patches") or transformer_options.get("patches_replace"):
            return False
        if self.disable_self_attn or not x.is_contiguous():
            return False
        for n in (self.norm2, self.norm3):
            if n.comfy_cast_weights or n.weight is None or n.weight.dtype != torch.bfloat16:
                return False
        return x.dtype == torch.bfloat16 and x.is_cuda

    def _forward_fast(self, x, context):
        n = self.norm1(x)
        a = self.attn1.forward_self_fused(n)
        n = _native.add_layer_norm_(x, a, self.norm2.weight, self.norm2.bias, self.norm2.eps)   # x += a
        a = self.attn2.forward_cross_cached(n, context) if context is not None else self.attn2(n)
        n = _native.add_layer_norm_(x, a, self.norm3.weight, self.norm3.bias, self.norm3.eps)   # x += a
        ff_out = self.ff.net[2]
        if isinstance(self.ff.net[0], GEGLU) and not ff_out.comfy_cast_weights and ff_out.weight_function is None \
                and ff_out.bias_function is None and ff_out.weight.dtype == torch.bfloat16 \
                and ff_out.weight.shape[0] % 32 == 0:
            # x = ff.net.2(geglu(n)) + x  with the residual add in the GEMM epilogue, written in place
            return _native.linear(self.ff.net[0](n), ff_out.weight, ff_out.bias, residual=x, out=x)
        h = self.ff(n)
        h += x
        return h

    def _forward(self, x, context=None, transformer_options={}):
        if self._fast_path_ok(x, transformer_options):
            return self._forward_fast(x, context)

        extra_options = {}
        block = transformer_options.get("block", None)
        block_index = transformer_options.get("block_index", 0)
        transformer_patches = {}
        transformer_patches_replace = {}
        for k in transformer_options:
            if k == "patches":
                transformer_patches = transformer_options[k]
            elif k == "patches_replace":
                transformer_patches_replace = transformer_options[k]
            else:
                extra_options[k] = transformer_options[k]
        extra_options["n_heads"] = self.n_heads
        extra_options["dim_head"] = self.d_head

        n = self.norm1(x)
        context_attn1 = context if self.disable_self_attn else None
        value_attn1 = None
        if "attn1_patch" in transformer_patches:
            if context_attn1 is None:
                context_attn1 = n
            value_attn1 = context_attn1
            for p in transformer_patches["attn1_patch"]:
                n, context_attn1, value_attn1 = p(n, context_attn1, value_attn1, extra_options)

        transformer_block = (block[0], block[1], block_index) if block is not None else None
        attn1_replace_patch = transformer_patches_replace.get("attn1", {})
        block_attn1 = transformer_block
        if block_attn1 not in attn1_replace_patch:
            block_attn1 = block
        if block_attn1 in attn1_replace_patch:
            if context_attn1 is None:
                context_attn1 = n
                value_attn1 = n
            n = self.attn1.to_q(n)
            context_attn1 = self.attn1.to_k(context_attn1)
            value_attn1 = self.attn1.to_v(value_attn1)
            n = attn1_replace_patch[block_attn1](n, context_attn1, value_attn1, extra_options)
            n = self.attn1.to_out(n)
        else:
            n = self.attn1(n, context=context_attn1, value=value_attn1)
        for p in transformer_patches.get("attn1_output_patch", []):
            n = p(n, extra_options)

        x += n
        for p in transformer_patches.get("middle_patch", []):
            x = p(x, extra_options)

        n = self.norm2(x)
        context_attn2 = context
        value_attn2 = None
        if "attn2_patch" in transformer_patches:
            value_attn2 = context_attn2
            for p in transformer_patches["attn2_patch"]:
                n, context_attn2, value_attn2 = p(n, context_attn2, value_attn2, extra_options)
        attn2_replace_patch = transformer_patches_replace.get("attn2", {})
        block_attn2 = transformer_block
        if block_attn2 not in attn2_replace_patch:
            block_attn2 = block
        if block_attn2 in attn2_replace_patch:
            if value_attn2 is None:
                value_attn2 = context_attn2
            n = self.attn2.to_q(n)
            context_attn2 = self.attn2.to_k(context_attn2)
            value_attn2 = self.attn2.to_v(value_attn2)
            n = attn2_replace_patch[block_attn2](n, context_attn2, value_attn2, extra_options)
            n = self.attn2.to_out(n)
        else:
            n = self.attn2(n, context=context_attn2, value=value_attn2)
        for p in transformer_patches.get("attn2_output_patch", []):
            n = p(n, extra_options)

        x += n
        x_skip = x
        x = self.ff(self.norm3(x))
        x += x_skip
        return x


# ----------------------------------------------------------------------------------------------
# spatial transformer
# ----------------------------------------------------------------------------------------------
def _fast_layout(x) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16 and x.is_contiguous()


def _is_channels_last(x) -> bool:
    return x.dim() == 4 and not x.is_contiguous() and x.is_contiguous(memory_format=torch.channels_last)


def _to_tokens(x):
    """'b c h w -> b (h w) c' (attention.py:628). Free for channels_last activations: NHWC *is* token-major."""
    b, c, h, w = x.shape
    if _is_channels_last(x):
        return x.permute(0, 2, 3, 1).reshape(b, h * w, c)
    if _fast_layout(x) and c % 8 == 0 and (h * w) % 8 == 0:
        return _native.transpose_add(x.view(b, c, h * w))
    return x.permute(0, 2, 3, 1).reshape(b, h * w, c).contiguous()


def _to_image(x, h, w, residual):
    """'b (h w) c -> b c h w' (attention.py:636), plus the `+ x_in` of attention.py:639 when `residual` is given."""
    b, n, c = x.shape
    if residual is not None and _is_channels_last(residual) and x.is_contiguous():
        # channels_last pipeline: the token tensor already is the NHWC image; the add stays a plain vectorised add
        return _native.residual_add(x.view(b, h, w, c).permute(0, 3, 1, 2), residual)
    if _fast_layout(x) and c % 8 == 0 and n % 8 == 0 and (residual is None or _fast_layout(residual)):
        return _native.transpose_add(x, residual).view(b, c, h, w)
    x = x.reshape(b, h, w, c).permute(0, 3, 1, 2).contiguous()
    return x if residual is None else x + residual
class SpatialTransformer(nn.Module):
    """attention.py:577-639: GroupNorm(32, eps 1e-6) -> proj_in -> tokens -> blocks -> proj_out -> image, + x_in."""
    def __init__(self, in_channels, n_heads, d_head, depth=1, dropout=0., context_dim=None, disable_self_attn=False,
                 use_linear=False, use_checkpoint=True, dtype=None, device=None, operations=ops):
        super().__init__()
        if exists(context_dim) and not isinstance(context_dim, list):
            context_dim = [context_dim] * depth
        self.in_channels = in_channels
        inner_dim = n_heads * d_head
        self.norm = operations.GroupNorm(num_groups=32, num_channels=in_channels, eps=1e-6, affine=True, dtype=dtype,
                                         device=device)
        if not use_linear:
            self.proj_in = operations.Conv2d(in_channels, inner_dim, kernel_size=1, stride=1, padding=0, dtype=dtype,
                                             device=device)
        else:
            self.proj_in = operations.Linear(in_channels, inner_dim, dtype=dtype, device=device)
        self.transformer_blocks = nn.ModuleList([
            BasicTransformerBlock(inner_dim, n_heads, d_head, dropout=dropout, context_dim=context_dim[d],
                                  disable_self_attn=disable_self_attn, checkpoint=use_checkpoint, dtype=dtype,
                                  device=device, operations=operations) for d in range(depth)])
        if not use_linear:
            self.proj_out = operations.Conv2d(inner_dim, in_channels, kernel_size=1, stride=1, padding=0,
                                              dtype=dtype, device=device)
        else:
            self.proj_out = operations.Linear(in_channels, inner_dim, dtype=dtype, device=device)
        self.use_linear = use_linear

    def forward(self, x, context=None, transformer_options={}):
        if not isinstance(context, list):
            context = [context] * len(self.transformer_blocks)
        b, c, h, w = x.shape
        x_in = x
        x = self.norm(x)
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
            return _to_image(x, h, w, x_in)            # layout change and `+ x_in` in one kernel
        if _is_channels_last(x_in) and x.is_contiguous():
            x = x.view(b, h, w, -1).permute(0, 3, 1, 2)         # NHWC view, no copy
        else:
            x = _to_image(x, h, w, None)
        x = self.proj_out(x)
        return _native.residual_add(x, x_in)
