"""Operator-level entry points of the B200 path, written as functions over the reference's OWN module classes
(comfy/ops.py, comfy/ldm/modules/attention.py): integration.install() binds them as the `forward` of those classes
and as the module-level attention function. They never subclass or re-declare a reference class.

Routing rule, the same everywhere: a bf16 CUDA tensor runs the hand-written kernel (through the C ABI,
_native.py); anything else -- the CLIP text encoder on the CPU, an fp32 VAE, a manual-cast module -- is handed to the
function that was bound before install(), i.e. stays exactly what the reference does. The attention function is the
exception: `optimized_attention` is the UNet's hot call and has no other implementation here, so it raises for
tensors the kernel cannot take (the reference's CPU / small-input callers go through
optimized_attention_for_device, attention.py:370-383, which install() leaves alone).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _native

# the functions that were bound before install() (filled in by integration.install)
ORIGINAL: dict = {}


# What the kernels take. (Module attributes rather than literals so that the CPU test-suite can run this host code
# against a torch stand-in for _native: tests/fake_native.py. Nothing in the product changes them.)
ACT_DTYPE = torch.bfloat16
DEVICE_TYPE = "cuda"


def on_fast_device(t: torch.Tensor) -> bool:
    """A tensor of the UNet's activation type on the device the kernels run on."""
    return t.device.type == DEVICE_TYPE and t.dtype == ACT_DTYPE


def is_latent(t: torch.Tensor) -> bool:
    """A sampler-side tensor the tail kernels take: contiguous fp32 on the device."""
    return t.device.type == DEVICE_TYPE and t.dtype == torch.float32 and t.is_contiguous()


def _cast(module, x):
    """weight / bias of a reference op module for input x: the reference's per-call cast (comfy/ops.py:22-32) when
    the module asks for it (manual_cast classes, low-vram weight functions), the parameters themselves otherwise."""
    if module.comfy_cast_weights:
        import comfy.ops
        return comfy.ops.cast_bias_weight(module, x)
    return module.weight, module.bias


# ----------------------------------------------------------------------------------------------
# attention  (comfy/ldm/modules/attention.py:337-368)
# ----------------------------------------------------------------------------------------------
def attention_b200(q, k, v, heads, mask=None):
    """Drop-in for attention_pytorch: q [B, Nq, heads*d], k/v [B, Nk, heads*d] -> [B, Nq, heads*d]; inputs are not
    modified. cgs_attention_fwd, or cgs_attention_masked_fwd when a mask is given (boolean keep-mask or additive
    float, attention.py:111-123). No SDPA, no fallback."""
    if not on_fast_device(q):
        raise RuntimeError(f"attention_b200: bf16 CUDA tensors required, got {q.dtype} on {q.device} "
                           "(run the UNet in bf16: install() sets --bf16-unet; CPU callers use "
                           "optimized_attention_for_device)")
    return _native.attention(q, k, v, heads, mask)


# ----------------------------------------------------------------------------------------------
# ops.LayerNorm / ops.GroupNorm  (comfy/ops.py:82-113)
# ----------------------------------------------------------------------------------------------
def layernorm_forward(self, input):
    if on_fast_device(input) and self.weight is not None and self.bias is not None and len(self.normalized_shape) == 1:
        weight, bias = _cast(self, input)
        if weight.dtype == ACT_DTYPE:
            return _native.layer_norm(input, weight, bias, self.eps)
    return ORIGINAL["LayerNorm.forward"](self, input)


def groupnorm_kernel_ok(self, input) -> bool:
    if not (on_fast_device(input) and input.dim() == 4 and self.affine):
        return False
    c = input.shape[1]
    if input.is_contiguous():                                       # NCHW kernel
        return (input.shape[2] * input.shape[3]) % 8 == 0
    if input.is_contiguous(memory_format=torch.channels_last):      # NHWC kernels
        cpg = c // self.num_groups
        return c % 8 == 0 and c % self.num_groups == 0 and (cpg >= 8 or cpg == 4) and self.num_groups <= 32 and c <= 2560
    return False


def groupnorm_apply(self, input, silu: bool = False, add_nc: Optional[torch.Tensor] = None):
    """GroupNorm with the SiLU that follows it in a ResBlock and the per-(sample, channel) addend that precedes it
    (timestep embedding + the previous convolution's bias) in one kernel; the reference's three library calls
    when the tensor is not one the kernels take."""
    weight, bias = _cast(self, input)
    if groupnorm_kernel_ok(self, input) and weight.dtype == ACT_DTYPE:
        return _native.group_norm(input, weight, bias, self.num_groups, self.eps, silu, add_nc)
    if add_nc is not None:
        input = input + add_nc[..., None, None]
    out = torch.nn.functional.group_norm(input, self.num_groups, weight, bias, self.eps)
    return torch.nn.functional.silu(out) if silu else out


def groupnorm_forward(self, input):
    if groupnorm_kernel_ok(self, input):
        return groupnorm_apply(self, input)
    return ORIGINAL["GroupNorm.forward"](self, input)


# ----------------------------------------------------------------------------------------------
# GEGLU  (comfy/ldm/modules/attention.py:56-63)
# ----------------------------------------------------------------------------------------------
def geglu_forward(self, x):
    """Linear -> chunk -> gelu -> mul as ONE GEMM with the gate in its epilogue; the [M, 2N] tensor never exists."""
    proj = self.proj
    if on_fast_device(x) and proj.weight.shape[0] % 256 == 0 and proj.weight.shape[1] % 8 == 0:
        weight, bias = _cast(proj, x)
        if weight.dtype == ACT_DTYPE:
            return _native.geglu(x, weight, bias)
    return ORIGINAL["GEGLU.forward"](self, x)


# ----------------------------------------------------------------------------------------------
# convolutions
# ----------------------------------------------------------------------------------------------
def conv_plain(conv) -> bool:
    """A resident bf16 Conv2d whose bias may be applied by the consumer of its output."""
    return (isinstance(conv, torch.nn.Conv2d) and not conv.comfy_cast_weights and conv.weight_function is None
            and conv.bias_function is None and conv.bias is not None and conv.bias.dtype == ACT_DTYPE
            and conv.weight.dtype == ACT_DTYPE)


def conv_nobias(conv, x):
    """The convolution without its bias: ATen adds a cuDNN convolution's bias in a separate full pass over the
    output, so the caller hands the bias to the kernel that consumes this output (GroupNorm's addend, the residual
    add, the skip concatenation)."""
    return conv._conv_forward(x, conv.weight, None)


def tensor_version(t) -> int:
    """Version counter of a tensor; inference tensors do not track one (treated as immutable: the reference's
    executor loads checkpoints under torch.inference_mode(), execution.py:537)."""
    try:
        return t._version
    except RuntimeError:
        return -1


def conv3x3_native_ok(conv, x) -> bool:
    """The convolution can run on the library's implicit GEMM (cgs_conv3x3_nhwc_fwd): 3x3, stride 1, padding 1,
    channels-last bf16 input and weight, and an image whose pixel tiles are mostly real pixels."""
    if not _native.FEATURES["native_conv"]:
        return False
    if not (conv.kernel_size == (3, 3) and conv.stride == (1, 1) and conv.padding == (1, 1) and conv.dilation == (1, 1)
            and conv.groups == 1 and conv.padding_mode == "zeros"):
        return False
    cout, cin = conv.weight.shape[:2]
    if cin % 8 or cout % 32 or x.dim() != 4 or x.shape[1] != cin:
        return False
    cl = torch.channels_last
    if not (x.is_contiguous(memory_format=cl) and conv.weight.is_contiguous(memory_format=cl)):
        return False
    if x.shape[1] == 1 or x.stride(1) != 1:          # (size-1 dims make is_contiguous ambiguous)
        return False
    return _native.conv3x3_tile_efficiency(x.shape[2], x.shape[3]) >= 0.75


def conv3x3(conv, x, bias=None, residual=None):
    """conv(x) without its own bias, + bias + residual: one launch on the library's implicit GEMM when it can run
    there, otherwise the bias-free cuDNN convolution followed by the add kernel."""
    if conv3x3_native_ok(conv, x):
        if residual is not None:
            residual = residual.contiguous(memory_format=torch.channels_last)
        return _native.conv3x3(x, conv.weight, bias, residual)
    y = conv_nobias(conv, x)
    if residual is not None:
        return _native.residual_add(residual, y, bias)
    if bias is not None:
        y = y + bias.view(1, -1, 1, 1)
    return y


def upsample2x_conv3x3(conv, x, bias=None):
    """conv(interpolate(x, 2x nearest)) (+ bias, not the conv's own), or None when it cannot run on the library's kernel.
    The pre-summed phase weights are derived once per weight tensor and kept on the module."""
    if not (_native.FEATURES["native_upconv"] and conv3x3_native_ok(conv, x)):
        return None
    key = (conv.weight.data_ptr(), tensor_version(conv.weight))
    cached = conv.__dict__.get("_cgs_upconv")
    if cached is None or cached[0] != key:
        cached = (key, _native.upsample2x_conv_weights(conv.weight))
        conv.__dict__["_cgs_upconv"] = cached
    return _native.upsample2x_conv3x3(x, cached[1], bias)


def linear_plain(lin) -> bool:
    return (not lin.comfy_cast_weights and lin.weight_function is None and lin.bias_function is None
            and lin.weight.dtype == ACT_DTYPE)
