"""Operator classes with the reference's `operations=` contract (comfy/ops.py:34-143).

`b200_ops` is a drop-in for `comfy.ops.disable_weight_init`: the same attribute names
(Linear, Conv2d, Conv3d, GroupNorm, LayerNorm, ConvTranspose2d, conv_nd), each a subclass of the
matching torch.nn class with `comfy_cast_weights`, `weight_function`, `bias_function`, a no-op
`reset_parameters` and `forward_comfy_cast_weights`; state-dict keys stay `weight` / `bias`.

What changes is where the math runs for bf16 CUDA tensors:
  LayerNorm.forward -> cgs_layernorm_fwd   (hand-written, HBM-bound)
  Linear.forward    -> cuBLASLt via F.linear by default (the reference's own path, K4 in SURVEY 2b),
                       or cgs_linear_fwd (tcgen05) when `b200_ops.native_linear` is set.
Conv / GroupNorm stay on torch (cuDNN / ATen), as SURVEY.md section 8 scopes them.
"""
from __future__ import annotations

import torch

from . import _native


def cast_bias_weight(s, input):
    """comfy/ops.py:22-32: per-call cast of weight/bias to the input's device and dtype."""
    bias = None
    if s.bias is not None:
        bias = s.bias.to(device=input.device, dtype=input.dtype, non_blocking=True)
        if s.bias_function is not None:
            bias = s.bias_function(bias)
    weight = s.weight.to(device=input.device, dtype=input.dtype, non_blocking=True)
    if s.weight_function is not None:
        weight = s.weight_function(weight)
    return weight, bias


class CastWeightBiasOp:
    comfy_cast_weights = False
    weight_function = None
    bias_function = None


def _native_ok(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16


class b200_ops:
    native_linear = False  # route Linear through the tcgen05 GEMM instead of cuBLASLt

    class Linear(torch.nn.Linear, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def _apply_linear(self, input, weight, bias):
            if b200_ops.native_linear and _native_ok(input) and weight.dtype == torch.bfloat16 \
                    and weight.shape[0] % 32 == 0 and weight.shape[1] % 8 == 0:
                return _native.linear(input, weight, bias)
            return torch.nn.functional.linear(input, weight, bias)

        def forward_comfy_cast_weights(self, input):
            weight, bias = cast_bias_weight(self, input)
            return self._apply_linear(input, weight, bias)

        def forward(self, input):
            if self.comfy_cast_weights:
                return self.forward_comfy_cast_weights(input)
            return self._apply_linear(input, self.weight, self.bias)

    class LayerNorm(torch.nn.LayerNorm, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def _apply_ln(self, input, weight, bias):
            # No CPU / PyTorch fallback: this class is the CUDA path or an error.
            if not _native_ok(input):
                raise RuntimeError(f"b200_ops.LayerNorm needs a CUDA bf16 tensor (got {input.device}, "
                                   f"{input.dtype}); there is no fallback on this path")
            if weight is None or bias is None or len(self.normalized_shape) != 1:
                raise RuntimeError("b200_ops.LayerNorm: only affine LayerNorm over the last dim is supported")
            return _native.layer_norm(input, weight, bias, self.eps)

        def forward_comfy_cast_weights(self, input):
            weight, bias = (None, None)
            if self.weight is not None:
                weight, bias = cast_bias_weight(self, input)
            return self._apply_ln(input, weight, bias)

        def forward(self, input):
            if self.comfy_cast_weights:
                return self.forward_comfy_cast_weights(input)
            return self._apply_ln(input, self.weight, self.bias)

    class Conv2d(torch.nn.Conv2d, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def forward_comfy_cast_weights(self, input):
            weight, bias = cast_bias_weight(self, input)
            return self._conv_forward(input, weight, bias)

        def forward(self, *args, **kwargs):
            if self.comfy_cast_weights:
                return self.forward_comfy_cast_weights(*args, **kwargs)
            return super().forward(*args, **kwargs)

        def forward_nobias(self, input):
            """The convolution without its bias: ATen adds a cuDNN convolution's bias in a separate full pass, so
            the ResBlock hands the bias to the kernel that consumes this output instead (GroupNorm's add_nc, the
            residual add)."""
            return self._conv_forward(input, self.weight, None)

    class Conv3d(torch.nn.Conv3d, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def forward_comfy_cast_weights(self, input):
            weight, bias = cast_bias_weight(self, input)
            return self._conv_forward(input, weight, bias)

        def forward(self, *args, **kwargs):
            if self.comfy_cast_weights:
                return self.forward_comfy_cast_weights(*args, **kwargs)
            return super().forward(*args, **kwargs)

    class GroupNorm(torch.nn.GroupNorm, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def native_ok(self, input):
            if not (_native_ok(input) and input.dim() == 4 and self.affine):
                return False
            c = input.shape[1]
            if input.is_contiguous():                                       # NCHW kernel
                return (input.shape[2] * input.shape[3]) % 8 == 0
            if input.is_contiguous(memory_format=torch.channels_last):      # NHWC kernels
                return c % 8 == 0 and c // self.num_groups >= 8 and self.num_groups <= 32 and c <= 2560
            return False

        def forward_fused(self, input, silu=False, add_nc=None):
            """GroupNorm with the SiLU that follows it in a ResBlock fused in (cgs_groupnorm_fwd) when the tensor
            is a contiguous NCHW bf16 CUDA tensor; otherwise the reference's F.group_norm (+ F.silu) library path,
            which is what the reference runs for this op."""
            if self.comfy_cast_weights:
                weight, bias = cast_bias_weight(self, input)
            else:
                weight, bias = self.weight, self.bias
            if self.native_ok(input) and weight.dtype == torch.bfloat16:
                return _native.group_norm(input, weight, bias, self.num_groups, self.eps, silu, add_nc)
            if add_nc is not None:
                input = input + add_nc[..., None, None]
            out = torch.nn.functional.group_norm(input, self.num_groups, weight, bias, self.eps)
            return torch.nn.functional.silu(out) if silu else out

        def forward_comfy_cast_weights(self, input):
            return self.forward_fused(input, False)

        def forward(self, input):
            return self.forward_fused(input, False)

    class ConvTranspose2d(torch.nn.ConvTranspose2d, CastWeightBiasOp):
        def reset_parameters(self):
            return None

        def forward_comfy_cast_weights(self, input, output_size=None):
            num_spatial_dims = 2
            output_padding = self._output_padding(input, output_size, self.stride, self.padding,
                                                  self.kernel_size, num_spatial_dims, self.dilation)
            weight, bias = cast_bias_weight(self, input)
            return torch.nn.functional.conv_transpose2d(input, weight, bias, self.stride, self.padding,
                                                        output_padding, self.groups, self.dilation)

        def forward(self, *args, **kwargs):
            if self.comfy_cast_weights:
                return self.forward_comfy_cast_weights(*args, **kwargs)
            return super().forward(*args, **kwargs)

    @classmethod
    def conv_nd(s, dims, *args, **kwargs):
        if dims == 2:
            return s.Conv2d(*args, **kwargs)
        if dims == 3:
            return s.Conv3d(*args, **kwargs)
        raise ValueError(f"unsupported dimensions: {dims}")


class manual_cast(b200_ops):
    """comfy/ops.py:145-163: same classes with comfy_cast_weights = True."""
    class Linear(b200_ops.Linear):
        comfy_cast_weights = True

    class Conv2d(b200_ops.Conv2d):
        comfy_cast_weights = True

    class Conv3d(b200_ops.Conv3d):
        comfy_cast_weights = True

    class GroupNorm(b200_ops.GroupNorm):
        comfy_cast_weights = True

    class LayerNorm(b200_ops.LayerNorm):
        comfy_cast_weights = True

    class ConvTranspose2d(b200_ops.ConvTranspose2d):
        comfy_cast_weights = True
