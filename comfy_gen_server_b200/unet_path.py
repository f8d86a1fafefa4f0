"""The UNet around the transformer blocks on the B200 path: ResBlock._forward, Upsample.forward and UNetModel.forward
(comfy/ldm/modules/diffusionmodules/openaimodel.py:60-96, 235-264, 819-890) as functions bound onto the reference's
classes by integration.install(). SURVEY.md section 8 rows f2 / f3.

Every function takes its fast path only for the plain case (2-D image UNet, resident bf16 weights, bf16 CUDA
activations, no ControlNet residuals, no block patches, no video arguments) and hands anything else to the
reference's own method.

What the fast path changes:
  * activations are channels-last end to end: cuDNN runs its NHWC kernels without layout conversions, GroupNorm reads
    16-byte channel vectors, and a SpatialTransformer's token matrix is a view of the image;
  * GroupNorm -> SiLU is one kernel, and it also applies `h + emb_out` (the timestep embedding);
  * convolutions run bias-free (ATen adds a cuDNN convolution's bias in a separate full pass); the biases ride along
    in the kernels that consume the outputs: conv1's in GroupNorm's addend, conv2's (+ the skip conv's) in the
    residual add, an Upsample conv's in the next skip concatenation;
  * skip concatenation and nearest 2x up-sampling are 16-byte-per-thread channels-last kernels.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import _native
from .layers import (ORIGINAL, on_fast_device, conv_plain, conv_nobias, conv3x3, upsample2x_conv3x3,
                     groupnorm_apply)


# ----------------------------------------------------------------------------------------------
# ResBlock._forward
# ----------------------------------------------------------------------------------------------
def _resblock_fast_ok(blk, x) -> bool:
    if blk.updown or blk.use_scale_shift_norm or blk.skip_t_emb or blk.exchange_temb_dims:
        return False
    if not (on_fast_device(x) and x.dim() == 4):
        return False
    if len(blk.in_layers) != 3 or len(blk.out_layers) != 4:
        return False
    convs = [blk.in_layers[2], blk.out_layers[3]]
    if not isinstance(blk.skip_connection, nn.Identity):
        convs.append(blk.skip_connection)
    return all(conv_plain(c) for c in convs)


def resblock_forward(self, x, emb):
    if not _resblock_fast_ok(self, x):
        return ORIGINAL["ResBlock._forward"](self, x, emb)
    conv1, conv2, skip = self.in_layers[2], self.out_layers[3], self.skip_connection
    h = conv3x3(conv1, groupnorm_apply(self.in_layers[0], x, silu=True))
    emb_out = self.emb_layers(emb).type(h.dtype) + conv1.bias            # [N, C]: conv1's bias joins the embedding
    h = groupnorm_apply(self.out_layers[0], h, silu=True, add_nc=emb_out.contiguous())
    h = self.out_layers[2](h)                                             # Dropout (identity at p = 0)
    # skip_connection(x) + conv2(h): the biases and the skip image ride in conv2's epilogue (or in the add kernel)
    if isinstance(skip, nn.Identity):
        return conv3x3(conv2, h, conv2.bias, x)
    return conv3x3(conv2, h, conv2.bias + skip.bias, _skip_conv(skip, x))


def _skip_conv(skip, x):
    """The ResBlock's 1x1 channel-changing skip convolution, bias-free: on channels-last pixels it is a plain GEMM."""
    cout, cin = skip.weight.shape[:2]
    if (_native.FEATURES["native_skip"] and skip.kernel_size == (1, 1) and x.is_contiguous(memory_format=torch.channels_last)
            and x.stride(1) == 1 and cin % 8 == 0 and cout % 32 == 0):
        n, _, hh, ww = x.shape
        y = _native.linear(x.permute(0, 2, 3, 1).reshape(n * hh * ww, cin), skip.weight.reshape(cout, cin))
        return y.view(n, hh, ww, cout).permute(0, 3, 1, 2)
    return conv_nobias(skip, x)


# ----------------------------------------------------------------------------------------------
# Upsample.forward
# ----------------------------------------------------------------------------------------------
def upsample_forward(self, x, output_shape=None, bias_out=None):
    """bias_out (a list): when given and the conv can run bias-free, its bias is appended there instead of being
    added here -- the caller folds it into the skip concatenation that consumes this output."""
    if self.dims == 3 or not (on_fast_device(x) and x.dim() == 4):
        return ORIGINAL["Upsample.forward"](self, x, output_shape)
    shape = [x.shape[2] * 2, x.shape[3] * 2]
    if output_shape is not None:
        shape = [output_shape[2], output_shape[3]]
    if shape == [x.shape[2] * 2, x.shape[3] * 2]:
        if self.use_conv and bias_out is not None and conv_plain(self.conv):
            y = upsample2x_conv3x3(self.conv, x)          # interpolate + conv in one go, on the small image
            if y is not None:
                bias_out.append(self.conv.bias)
                return y
        x = _native.upsample_nearest2x(x)
    else:
        x = torch.nn.functional.interpolate(x, size=shape, mode="nearest")
    if self.use_conv:
        if bias_out is not None and conv_plain(self.conv):
            bias_out.append(self.conv.bias)
            return conv3x3(self.conv, x)
        x = self.conv(x)
    return x


# ----------------------------------------------------------------------------------------------
# UNetModel.forward
# ----------------------------------------------------------------------------------------------
def _unet_fast_ok(unet, x, y, control, transformer_options, kwargs) -> bool:
    if control is not None or kwargs or unet.predict_codebook_ids or unet.default_num_video_frames is not None:
        return False
    patches = transformer_options.get("patches", {})
    if any(k in patches for k in ("input_block_patch", "input_block_patch_after_skip", "output_block_patch")):
        return False
    return on_fast_device(x) and x.dim() == 4 and (y is not None) == (unet.num_classes is not None)


def _wants_channels_last(unet) -> bool:
    """Channels-last needs >= 8 channels per GroupNorm group for the 16-byte vector kernels (model_channels >= 256:
    every SD / SDXL UNet). Decided once per module and remembered on it."""
    flag = unet.__dict__.get("_cgs_nhwc")
    if flag is None:
        flag = unet.model_channels // 32 >= 8
        unet.__dict__["_cgs_nhwc"] = flag
    return flag


def _ensure_channels_last_weights(unet) -> None:
    """Convolution weights in channels-last, converted in place the first time (and again for any tensor a weight
    patch replaced since): cuDNN then needs no per-call weight transform."""
    for p in unet.parameters():
        if p.dim() == 4 and not p.is_contiguous(memory_format=torch.channels_last):
            with torch.inference_mode(p.is_inference()):    # an inference tensor stays one (the server loads checkpoints
                # under torch.inference_mode()), a normal tensor stays usable outside inference mode
                p.data = p.data.contiguous(memory_format=torch.channels_last)


def _run_block(seq, x, emb, context, transformer_options, output_shape, up_bias, ref):
    for layer in seq:
        if isinstance(layer, ref.TimestepBlock):
            x = layer(x, emb)
        elif isinstance(layer, ref.SpatialTransformer):
            x = layer(x, context, transformer_options)
            if "transformer_index" in transformer_options:
                transformer_options["transformer_index"] += 1
        elif isinstance(layer, ref.Upsample):
            x = upsample_forward(layer, x, output_shape, up_bias)
        else:
            x = layer(x)
    return x


def unet_forward(self, x, timesteps=None, context=None, y=None, control=None, transformer_options={}, **kwargs):
    if not _unet_fast_ok(self, x, y, control, transformer_options, kwargs):
        return ORIGINAL["UNetModel.forward"](self, x, timesteps, context, y, control, transformer_options, **kwargs)
    import comfy.ldm.modules.diffusionmodules.openaimodel as ref
    transformer_options["original_shape"] = list(x.shape)
    transformer_options["transformer_index"] = 0
    if _wants_channels_last(self):
        _ensure_channels_last_weights(self)
        x = x.contiguous(memory_format=torch.channels_last)

    emb = self.time_embed(ref.timestep_embedding(timesteps, self.model_channels, repeat_only=False).to(x.dtype))
    if self.num_classes is not None:
        emb = emb + self.label_emb(y)

    hs = []
    h = x
    for i, module in enumerate(self.input_blocks):
        transformer_options["block"] = ("input", i)
        h = _run_block(module, h, emb, context, transformer_options, None, None, ref)
        hs.append(h)
    transformer_options["block"] = ("middle", 0)
    if self.middle_block is not None:
        h = _run_block(self.middle_block, h, emb, context, transformer_options, None, None, ref)

    pending = []                 # bias of the Upsample conv that produced h: added by the next skip concatenation
    last = len(self.output_blocks) - 1
    for i, module in enumerate(self.output_blocks):
        transformer_options["block"] = ("output", i)
        h = _native.cat_channels(h, hs.pop(), pending.pop() if pending else None)
        h = _run_block(module, h, emb, context, transformer_options, hs[-1].shape if hs else None,
                       None if i == last else pending, ref)
    h = h.type(x.dtype)
    out_norm, out_conv = self.out[0], self.out[2]
    return out_conv(groupnorm_apply(out_norm, h, silu=True))
