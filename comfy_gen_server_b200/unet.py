"""UNet host for the hot path: the reference's UNetModel interface and state-dict layout
(comfy/ldm/modules/diffusionmodules/openaimodel.py:819-890), with every SpatialTransformer being the
B200 implementation from .attention. ResBlock / Up/Downsample / GroupNorm / conv math stays on
torch (cuDNN / ATen) -- SURVEY.md section 8 scopes it out of the hand-written kernels.

Supported constructor surface: what the SD1.x / SD2.x / SDXL literal configs use
(comfy/model_detection.py:270-304). Video / temporal / codebook options raise NotImplementedError.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _native
from .attention import SpatialTransformer
from .ops import b200_ops

ops = b200_ops


def timestep_embedding(timesteps, dim, max_period=10000):
    """Sinusoidal embedding, cos first then sin (diffusionmodules/util.py:229-249)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(start=0, end=half, dtype=torch.float32,
                                                            device=timesteps.device) / half)
    args = timesteps[:, None].float() * freqs[None]
    embedding = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
    if dim % 2:
        embedding = torch.cat([embedding, torch.zeros_like(embedding[:, :1])], dim=-1)
    return embedding


class TimestepBlock(nn.Module):
    pass


def _norm_silu(norm, x, add_nc=None):
    """GroupNorm -> SiLU (the first two layers of ResBlock.in_layers / out_layers and of UNetModel.out), fused
    into one kernel when the norm is the B200 operator class. `add_nc` [N, C]: the timestep embedding that the
    ResBlock adds to h right before out_layers, folded into the same kernel."""
    fused = getattr(norm, "forward_fused", None)
    if fused is not None:
        return fused(x, True, add_nc)
    if add_nc is not None:
        x = x + add_nc[..., None, None]
    return F.silu(norm(x))


class Upsample(nn.Module):
    """openaimodel.py:60-96 (2-D, nearest, optional 3x3 conv)."""
    def __init__(self, channels, use_conv, dims=2, out_channels=None, padding=1, dtype=None, device=None,
                 operations=ops):
        super().__init__()
        self.channels = channels
        self.out_channels = out_channels or channels
        self.use_conv = use_conv
        if use_conv:
            self.conv = operations.conv_nd(dims, self.channels, self.out_channels, 3, padding=padding, dtype=dtype,
                                           device=device)

    def forward(self, x, output_shape=None, bias_out=None):
        """bias_out (a list): when given and the conv can run bias-free, its bias is appended there instead of being
        added here -- the caller folds it into the skip concatenation that consumes this output."""
        shape = [x.shape[2] * 2, x.shape[3] * 2]
        if output_shape is not None:
            shape = [output_shape[2], output_shape[3]]
        if shape == [x.shape[2] * 2, x.shape[3] * 2] and x.is_cuda:
            x = _native.upsample_nearest2x(x)
        else:
            x = F.interpolate(x, size=shape, mode="nearest")
        if self.use_conv:
            conv = self.conv
            if (bias_out is not None and _native.FEATURES["fold_conv_bias"] and x.is_cuda
                    and x.dtype == torch.bfloat16 and hasattr(conv, "forward_nobias")
                    and not conv.comfy_cast_weights and conv.bias is not None and conv.bias.dtype == torch.bfloat16):
                bias_out.append(conv.bias)
                return conv.forward_nobias(x)
            x = conv(x)
        return x


class Downsample(nn.Module):
    """openaimodel.py:98-126 (stride-2 3x3 conv)."""
    def __init__(self, channels, use_conv, dims=2, out_channels=None, padding=1, dtype=None, device=None,
                 operations=ops):
        super().__init__()
        self.channels = channels
        self.out_channels = out_channels or channels
        if not use_conv:
            raise NotImplementedError("avg-pool Downsample is not used by the SD/SDXL configs")
        self.op = operations.conv_nd(dims, self.channels, self.out_channels, 3, stride=2, padding=padding,
                                     dtype=dtype, device=device)

    def forward(self, x):
        return self.op(x)


class ResBlock(TimestepBlock):
    """openaimodel.py:129-264 without up/down/scale-shift (unused by the SD/SDXL configs)."""
    def __init__(self, channels, emb_channels, dropout, out_channels=None, dims=2, dtype=None, device=None,
                 operations=ops):
        super().__init__()
        self.channels = channels
        self.out_channels = out_channels or channels
        self.in_layers = nn.Sequential(
            operations.GroupNorm(32, channels, dtype=dtype, device=device), nn.SiLU(),
            operations.conv_nd(dims, channels, self.out_channels, 3, padding=1, dtype=dtype, device=device))
        self.emb_layers = nn.Sequential(
            nn.SiLU(), operations.Linear(emb_channels, self.out_channels, dtype=dtype, device=device))
        self.out_layers = nn.Sequential(
            operations.GroupNorm(32, self.out_channels, dtype=dtype, device=device), nn.SiLU(),
            nn.Dropout(p=dropout),
            operations.conv_nd(dims, self.out_channels, self.out_channels, 3, padding=1, dtype=dtype, device=device))
        if self.out_channels == channels:
            self.skip_connection = nn.Identity()
        else:
            self.skip_connection = operations.conv_nd(dims, channels, self.out_channels, 1, dtype=dtype, device=device)

    def _fold_bias(self, x):
        """Conv biases ride along in the consumers of the conv outputs when everything is resident bf16 on CUDA."""
        convs = [self.in_layers[2], self.out_layers[3]]
        if not isinstance(self.skip_connection, nn.Identity):
            convs.append(self.skip_connection)
        return (_native.FEATURES["fold_conv_bias"] and x.is_cuda and x.dtype == torch.bfloat16
                and hasattr(self.out_layers[0], "forward_fused")
                and all(hasattr(c, "forward_nobias") and not c.comfy_cast_weights and c.bias is not None
                        and c.bias.dtype == torch.bfloat16 for c in convs))

    def forward(self, x, emb):
        if self._fold_bias(x):
            conv1, conv2, skip = self.in_layers[2], self.out_layers[3], self.skip_connection
            h = conv1.forward_nobias(_norm_silu(self.in_layers[0], x))
            emb_out = self.emb_layers(emb).type(h.dtype) + conv1.bias          # [N, C]: conv1's bias joins the embedding
            h = conv2.forward_nobias(self.out_layers[2](_norm_silu(self.out_layers[0], h, emb_out.contiguous())))
            if isinstance(skip, nn.Identity):
                return _native.residual_add(x, h, conv2.bias)
            return _native.residual_add(skip.forward_nobias(x), h, conv2.bias + skip.bias)
        h = self.in_layers[2](_norm_silu(self.in_layers[0], x))
        emb_out = self.emb_layers(emb).type(h.dtype)
        h = self.out_layers[3](self.out_layers[2](_norm_silu(self.out_layers[0], h, emb_out.contiguous())))
        return _native.residual_add(self.skip_connection(x), h)


def forward_timestep_embed(ts, x, emb, context=None, transformer_options={}, output_shape=None, up_bias=None):
    """openaimodel.py:33-51. up_bias: see Upsample.forward."""
    for layer in ts:
        if isinstance(layer, TimestepBlock):
            x = layer(x, emb)
        elif isinstance(layer, SpatialTransformer):
            x = layer(x, context, transformer_options)
            if "transformer_index" in transformer_options:
                transformer_options["transformer_index"] += 1
        elif isinstance(layer, Upsample):
            x = layer(x, output_shape=output_shape, bias_out=up_bias)
        else:
            x = layer(x)
    return x


class TimestepEmbedSequential(nn.Sequential, TimestepBlock):
    def forward(self, *args, **kwargs):
        return forward_timestep_embed(self, *args, **kwargs)


def apply_control(h, control, name):
    if control is not None and name in control and len(control[name]) > 0:
        ctrl = control[name].pop()
        if ctrl is not None:
            h = h + ctrl
    return h


class UNetModel(nn.Module):
    def __init__(self, image_size, in_channels, model_channels, out_channels, num_res_blocks, dropout=0,
                 channel_mult=(1, 2, 4, 8), conv_resample=True, dims=2, num_classes=None, use_checkpoint=False,
                 dtype=torch.float32, num_heads=-1, num_head_channels=-1, num_heads_upsample=-1,
                 use_scale_shift_norm=False, resblock_updown=False, use_new_attention_order=False,
                 use_spatial_transformer=False, transformer_depth=1, context_dim=None, n_embed=None, legacy=True,
                 disable_self_attentions=None, num_attention_blocks=None, disable_middle_self_attn=False,
                 use_linear_in_transformer=False, adm_in_channels=None, transformer_depth_middle=None,
                 transformer_depth_output=None, use_temporal_resblock=False, use_temporal_attention=False,
                 device=None, operations=ops, **unused):
        super().__init__()
        if use_temporal_resblock or use_temporal_attention or n_embed is not None or use_scale_shift_norm \
                or resblock_updown or dims != 2:
            raise NotImplementedError("UNet option outside the SD1.x / SD2.x / SDXL image configs")
        if not use_spatial_transformer:
            raise NotImplementedError("UNets without SpatialTransformer blocks are out of scope")
        if num_heads == -1 and num_head_channels == -1:
            raise ValueError("Either num_heads or num_head_channels has to be set")
        self.in_channels = in_channels
        self.model_channels = model_channels
        self.out_channels = out_channels
        self.num_res_blocks = len(channel_mult) * [num_res_blocks] if isinstance(num_res_blocks, int) \
            else list(num_res_blocks)
        transformer_depth = list(transformer_depth)
        transformer_depth_output = list(transformer_depth_output)
        self.num_classes = num_classes
        self.dtype = dtype

        def heads_for(ch):
            if num_head_channels == -1:
                return num_heads, ch // num_heads
            return ch // num_head_channels, num_head_channels

        def attn(ch, depth, level):
            nh, dh = heads_for(ch)
            dsa = disable_self_attentions[level] if disable_self_attentions is not None else False
            return SpatialTransformer(ch, nh, dh, depth=depth, context_dim=context_dim, disable_self_attn=dsa,
                                      use_linear=use_linear_in_transformer, use_checkpoint=use_checkpoint,
                                      dtype=self.dtype, device=device, operations=operations)

        def res(ch, out_ch):
            return ResBlock(ch, time_embed_dim, dropout, out_channels=out_ch, dims=dims, dtype=self.dtype,
                            device=device, operations=operations)

        time_embed_dim = model_channels * 4
        self.time_embed = nn.Sequential(
            operations.Linear(model_channels, time_embed_dim, dtype=self.dtype, device=device), nn.SiLU(),
            operations.Linear(time_embed_dim, time_embed_dim, dtype=self.dtype, device=device))
        if self.num_classes is not None:
            if self.num_classes != "sequential":
                raise NotImplementedError("only the 'sequential' (ADM vector) class embedding is supported")
            assert adm_in_channels is not None
            self.label_emb = nn.Sequential(nn.Sequential(
                operations.Linear(adm_in_channels, time_embed_dim, dtype=self.dtype, device=device), nn.SiLU(),
                operations.Linear(time_embed_dim, time_embed_dim, dtype=self.dtype, device=device)))

        self.input_blocks = nn.ModuleList([TimestepEmbedSequential(
            operations.conv_nd(dims, in_channels, model_channels, 3, padding=1, dtype=self.dtype, device=device))])
        input_block_chans = [model_channels]
        ch = model_channels
        for level, mult in enumerate(channel_mult):
            for nr in range(self.num_res_blocks[level]):
                layers = [res(ch, mult * model_channels)]
                ch = mult * model_channels
                depth = transformer_depth.pop(0)
                if depth > 0 and (num_attention_blocks is None or nr < num_attention_blocks[level]):
                    layers.append(attn(ch, depth, level))
                self.input_blocks.append(TimestepEmbedSequential(*layers))
                input_block_chans.append(ch)
            if level != len(channel_mult) - 1:
                self.input_blocks.append(TimestepEmbedSequential(
                    Downsample(ch, conv_resample, dims=dims, out_channels=ch, dtype=self.dtype, device=device,
                               operations=operations)))
                input_block_chans.append(ch)

        mid = [res(ch, None)]
        self.middle_block = None
        if transformer_depth_middle >= -1:
            if transformer_depth_middle >= 0:
                nh, dh = heads_for(ch)
                mid += [SpatialTransformer(ch, nh, dh, depth=transformer_depth_middle, context_dim=context_dim,
                                           disable_self_attn=disable_middle_self_attn,
                                           use_linear=use_linear_in_transformer, use_checkpoint=use_checkpoint,
                                           dtype=self.dtype, device=device, operations=operations),
                        res(ch, None)]
            self.middle_block = TimestepEmbedSequential(*mid)

        self.output_blocks = nn.ModuleList([])
        for level, mult in list(enumerate(channel_mult))[::-1]:
            for i in range(self.num_res_blocks[level] + 1):
                ich = input_block_chans.pop()
                layers = [res(ch + ich, model_channels * mult)]
                ch = model_channels * mult
                depth = transformer_depth_output.pop()
                if depth > 0 and (num_attention_blocks is None or i < num_attention_blocks[level]):
                    layers.append(attn(ch, depth, level))
                if level and i == self.num_res_blocks[level]:
                    layers.append(Upsample(ch, conv_resample, dims=dims, out_channels=ch, dtype=self.dtype,
                                           device=device, operations=operations))
                self.output_blocks.append(TimestepEmbedSequential(*layers))

        self.out = nn.Sequential(
            operations.GroupNorm(32, ch, dtype=self.dtype, device=device), nn.SiLU(),
            operations.conv_nd(dims, model_channels, out_channels, 3, padding=1, dtype=self.dtype, device=device))

    def forward(self, x, timesteps=None, context=None, y=None, control=None, transformer_options={}, **kwargs):
        """openaimodel.py:819-890: same hook points (block ids, input/output block patches, control)."""
        transformer_options["original_shape"] = list(x.shape)
        transformer_options["transformer_index"] = 0
        transformer_patches = transformer_options.get("patches", {})
        assert (y is not None) == (self.num_classes is not None), \
            "must specify y if and only if the model is class-conditional"
        hs = []
        t_emb = timestep_embedding(timesteps, self.model_channels).to(x.dtype)
        emb = self.time_embed(t_emb)
        if self.num_classes is not None:
            assert y.shape[0] == x.shape[0]
            emb = emb + self.label_emb(y)

        h = x
        for id, module in enumerate(self.input_blocks):
            transformer_options["block"] = ("input", id)
            h = forward_timestep_embed(module, h, emb, context, transformer_options)
            h = apply_control(h, control, 'input')
            for p in transformer_patches.get("input_block_patch", []):
                h = p(h, transformer_options)
            hs.append(h)
            for p in transformer_patches.get("input_block_patch_after_skip", []):
                h = p(h, transformer_options)

        transformer_options["block"] = ("middle", 0)
        if self.middle_block is not None:
            h = forward_timestep_embed(self.middle_block, h, emb, context, transformer_options)
        h = apply_control(h, control, 'middle')

        # an Upsample conv's bias is added by the next skip concatenation (nothing reads h in between unless a control
        # net or an output_block_patch is installed)
        defer = [] if control is None and not transformer_patches.get("output_block_patch") else None
        for id, module in enumerate(self.output_blocks):
            transformer_options["block"] = ("output", id)
            hsp = hs.pop()
            hsp = apply_control(hsp, control, 'output')
            for p in transformer_patches.get("output_block_patch", []):
                h, hsp = p(h, hsp, transformer_options)
            pending = defer.pop() if defer else None
            h = _native.cat_channels(h, hsp, pending) if h.is_cuda else torch.cat([h, hsp], dim=1)
            del hsp
            output_shape = hs[-1].shape if len(hs) > 0 else None
            last = id == len(self.output_blocks) - 1
            h = forward_timestep_embed(module, h, emb, context, transformer_options, output_shape,
                                       None if last else defer)
        h = h.type(x.dtype)
        return self.out[2](_norm_silu(self.out[0], h))
