"""UNet configurations of the BASELINE.json model families, the block plan derived from them, and
seeded random-init state dicts (there are no checkpoints: bench and tests use random weights).

The literal configs are the reference's (comfy/model_detection.py:270-304); the plan follows the
reference constructor's traversal (comfy/ldm/modules/diffusionmodules/openaimodel.py:609-810) so
that state-dict key names are identical to the reference's and checkpoints / LoRA key maps stay valid.
"""
from __future__ import annotations

import math
import zlib
from typing import Dict, List, Tuple

import torch

SD15 = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True, 'legacy': False,
        'adm_in_channels': None, 'in_channels': 4, 'model_channels': 320, 'num_res_blocks': [2, 2, 2, 2],
        'transformer_depth': [1, 1, 1, 1, 1, 1, 0, 0], 'channel_mult': [1, 2, 4, 4], 'transformer_depth_middle': 1,
        'use_linear_in_transformer': False, 'context_dim': 768, 'num_heads': 8,
        'transformer_depth_output': [1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0],
        'use_temporal_attention': False, 'use_temporal_resblock': False}

SD21 = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True, 'legacy': False,
        'adm_in_channels': None, 'in_channels': 4, 'model_channels': 320, 'num_res_blocks': [2, 2, 2, 2],
        'transformer_depth': [1, 1, 1, 1, 1, 1, 0, 0], 'channel_mult': [1, 2, 4, 4], 'transformer_depth_middle': 1,
        'use_linear_in_transformer': True, 'context_dim': 1024, 'num_head_channels': 64,
        'transformer_depth_output': [1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0],
        'use_temporal_attention': False, 'use_temporal_resblock': False}

SDXL = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True, 'legacy': False,
        'num_classes': 'sequential', 'adm_in_channels': 2816, 'in_channels': 4, 'model_channels': 320,
        'num_res_blocks': [2, 2, 2], 'transformer_depth': [0, 0, 2, 2, 10, 10], 'channel_mult': [1, 2, 4],
        'transformer_depth_middle': 10, 'use_linear_in_transformer': True, 'context_dim': 2048,
        'num_head_channels': 64, 'transformer_depth_output': [0, 0, 0, 2, 2, 2, 10, 10, 10],
        'use_temporal_attention': False, 'use_temporal_resblock': False}

# Small configs with the same structure, for tests and golden fixtures (seconds on CPU).
TINY_SD15 = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True,
             'legacy': False, 'adm_in_channels': None, 'in_channels': 4, 'model_channels': 64,
             'num_res_blocks': [1, 1], 'transformer_depth': [1, 1], 'channel_mult': [1, 2],
             'transformer_depth_middle': 1, 'use_linear_in_transformer': False, 'context_dim': 96, 'num_heads': 8,
             'transformer_depth_output': [1, 1, 1, 1], 'use_temporal_attention': False,
             'use_temporal_resblock': False}

TINY_SD21 = dict(TINY_SD15, use_linear_in_transformer=True, num_head_channels=64, context_dim=128,
                 model_channels=64, channel_mult=[1, 2])
TINY_SD21.pop('num_heads')

TINY_SDXL = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True,
             'legacy': False, 'num_classes': 'sequential', 'adm_in_channels': 2816, 'in_channels': 4,
             'model_channels': 64, 'num_res_blocks': [1, 1], 'transformer_depth': [0, 2], 'channel_mult': [1, 2],
             'transformer_depth_middle': 2, 'use_linear_in_transformer': True, 'context_dim': 128,
             'num_head_channels': 64, 'transformer_depth_output': [0, 0, 2, 2],
             'use_temporal_attention': False, 'use_temporal_resblock': False}

# family -> (unet config, latent scale factor (comfy/latent_formats.py:15,28), prediction type, pooled/ADM)
FAMILIES = {
    "sd15": dict(unet=SD15, scale_factor=0.18215, prediction="eps", adm=False),
    "sd21v": dict(unet=SD21, scale_factor=0.18215, prediction="v", adm=False),
    "sdxl": dict(unet=SDXL, scale_factor=0.13025, prediction="eps", adm=True),
    "tiny_sd15": dict(unet=TINY_SD15, scale_factor=0.18215, prediction="eps", adm=False),
    "tiny_sd21v": dict(unet=TINY_SD21, scale_factor=0.18215, prediction="v", adm=False),
    "tiny_sdxl": dict(unet=TINY_SDXL, scale_factor=0.13025, prediction="eps", adm=True),
}


def _heads(cfg: dict, ch: int) -> Tuple[int, int]:
    """(num_heads, dim_head) for a transformer at `ch` channels (openaimodel.py:632-641)."""
    nhc = cfg.get('num_head_channels', -1)
    nh = cfg.get('num_heads', -1)
    if nhc == -1:
        return nh, ch // nh
    return ch // nhc, nhc


def unet_plan(cfg: dict) -> dict:
    """Block plan: lists of ('conv'|'res'|'attn'|'down'|'up', prefix, ...) per TimestepEmbedSequential."""
    mc = cfg['model_channels']
    mult = cfg['channel_mult']
    nrb = cfg['num_res_blocks']
    if isinstance(nrb, int):
        nrb = [nrb] * len(mult)
    td = list(cfg['transformer_depth'])
    tdo = list(cfg['transformer_depth_output'])
    ctx = cfg['context_dim']
    plan = {'input': [], 'middle': [], 'output': [], 'time_embed_dim': mc * 4, 'model_channels': mc,
            'adm_in_channels': cfg.get('adm_in_channels'), 'use_linear': cfg.get('use_linear_in_transformer', False),
            'context_dim': ctx, 'in_channels': cfg['in_channels'], 'out_channels': cfg['out_channels']}
    plan['input'].append([('conv', 'input_blocks.0.0', cfg['in_channels'], mc)])
    chans = [mc]
    ch = mc
    idx = 1
    for level, m in enumerate(mult):
        for _ in range(nrb[level]):
            layers = [('res', f'input_blocks.{idx}.0', ch, m * mc)]
            ch = m * mc
            depth = td.pop(0)
            if depth > 0:
                nh, dh = _heads(cfg, ch)
                layers.append(('attn', f'input_blocks.{idx}.1', ch, nh, dh, depth))
            plan['input'].append(layers)
            chans.append(ch)
            idx += 1
        if level != len(mult) - 1:
            plan['input'].append([('down', f'input_blocks.{idx}.0', ch)])
            chans.append(ch)
            idx += 1
    nh, dh = _heads(cfg, ch)
    tdm = cfg['transformer_depth_middle']
    mid = [('res', 'middle_block.0', ch, ch)]
    if tdm is not None and tdm >= 0:
        mid += [('attn', 'middle_block.1', ch, nh, dh, tdm), ('res', 'middle_block.2', ch, ch)]
    plan['middle'] = mid
    idx = 0
    for level, m in list(enumerate(mult))[::-1]:
        for i in range(nrb[level] + 1):
            ich = chans.pop()
            layers = [('res', f'output_blocks.{idx}.0', ch + ich, mc * m)]
            ch = mc * m
            depth = tdo.pop()
            if depth > 0:
                nh, dh = _heads(cfg, ch)
                layers.append(('attn', f'output_blocks.{idx}.{len(layers)}', ch, nh, dh, depth))
            if level and i == nrb[level]:
                layers.append(('up', f'output_blocks.{idx}.{len(layers)}', ch))
            plan['output'].append(layers)
            idx += 1
    plan['final_channels'] = ch
    return plan


def param_shapes(cfg: dict) -> Dict[str, Tuple[int, ...]]:
    """name -> shape for every parameter of the UNet, with the reference's state-dict key names."""
    plan = unet_plan(cfg)
    mc, ted, ctx = plan['model_channels'], plan['time_embed_dim'], plan['context_dim']
    s: Dict[str, Tuple[int, ...]] = {}

    def lin(p, i, o, bias=True):
        s[p + '.weight'] = (o, i)
        if bias:
            s[p + '.bias'] = (o,)

    def conv(p, i, o, k):
        s[p + '.weight'] = (o, i, k, k)
        s[p + '.bias'] = (o,)

    def norm(p, c):
        s[p + '.weight'] = (c,)
        s[p + '.bias'] = (c,)

    lin('time_embed.0', mc, ted)
    lin('time_embed.2', ted, ted)
    if plan['adm_in_channels']:
        lin('label_emb.0.0', plan['adm_in_channels'], ted)
        lin('label_emb.0.2', ted, ted)

    def block(layer):
        kind, p = layer[0], layer[1]
        if kind == 'conv':
            conv(p, layer[2], layer[3], 3)
        elif kind == 'res':
            cin, cout = layer[2], layer[3]
            norm(p + '.in_layers.0', cin)
            conv(p + '.in_layers.2', cin, cout, 3)
            lin(p + '.emb_layers.1', ted, cout)
            norm(p + '.out_layers.0', cout)
            conv(p + '.out_layers.3', cout, cout, 3)
            if cin != cout:
                conv(p + '.skip_connection', cin, cout, 1)
        elif kind == 'attn':
            _, _, c, nh, dh, depth = layer
            inner = nh * dh
            norm(p + '.norm', c)
            if plan['use_linear']:
                lin(p + '.proj_in', c, inner)
                lin(p + '.proj_out', inner, c)
            else:
                conv(p + '.proj_in', c, inner, 1)
                conv(p + '.proj_out', inner, c, 1)
            for d in range(depth):
                b = f'{p}.transformer_blocks.{d}'
                for a, cdim in (('attn1', inner), ('attn2', ctx)):
                    lin(f'{b}.{a}.to_q', inner, inner, bias=False)
                    lin(f'{b}.{a}.to_k', cdim, inner, bias=False)
                    lin(f'{b}.{a}.to_v', cdim, inner, bias=False)
                    lin(f'{b}.{a}.to_out.0', inner, inner)
                lin(f'{b}.ff.net.0.proj', inner, inner * 8)
                lin(f'{b}.ff.net.2', inner * 4, inner)
                for n in ('norm1', 'norm2', 'norm3'):
                    norm(f'{b}.{n}', inner)
        elif kind == 'down':
            conv(p + '.op', layer[2], layer[2], 3)
        elif kind == 'up':
            conv(p + '.conv', layer[2], layer[2], 3)

    for seq in plan['input']:
        for layer in seq:
            block(layer)
    for layer in plan['middle']:
        block(layer)
    for seq in plan['output']:
        for layer in seq:
            block(layer)
    norm('out.0', plan['final_channels'])
    conv('out.2', mc, plan['out_channels'], 3)
    return s


_RESIDUAL_OUT = ('.to_out.0.weight', '.ff.net.2.weight', '.proj_out.weight', '.out_layers.3.weight')


def seeded_state_dict(shapes: Dict[str, Tuple[int, ...]], seed: int, dtype=torch.float32, gain: float = 1.0,
                      residual_damp: float = 0.25, names=None) -> Dict[str, torch.Tensor]:
    """Deterministic random-init weights, independent of iteration order (each tensor has its own CPU
    generator seeded from (seed, crc32(name))). Fan-in-scaled normal for matrices/convs (residual
    output projections damped so activations stay O(1) through 16-70 transformer blocks), ones for
    norm scales, small normal for biases. The reference's modules leave parameters uninitialised
    (comfy/ops.py:41-42), so both sides load this dict."""
    out = {}
    for name in (names if names is not None else shapes):
        shape = shapes[name]
        g = torch.Generator(device='cpu').manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 63))
        if name.endswith('.weight') and len(shape) >= 2:
            fan_in = 1
            for d in shape[1:]:
                fan_in *= d
            std = gain / math.sqrt(fan_in)
            if name.endswith(_RESIDUAL_OUT):
                std *= residual_damp
            t = torch.randn(shape, generator=g) * std
        elif name.endswith('.weight'):
            t = torch.ones(shape) + 0.05 * torch.randn(shape, generator=g)
        else:
            t = 0.02 * torch.randn(shape, generator=g)
        out[name] = t.to(dtype)
    return out
