"""ORACLE -- test infrastructure, not product code.

A CPU restatement (plain torch fp32 / float64, functional, no nn.Module) of the reference's
algorithm for the per-step UNet denoise hot path. Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this package; the product
(comfy_gen_server_b200/) never does and has no CPU path of its own.

Pinning: the reference's own tests hold NO golden vectors for this path,
so the oracle is pinned against outputs of the reference itself, generated in the build
container by oracle/make_golden.py (imports /root/reference) and committed under tests/golden/.
tests/test_oracle_golden.py checks every function below against those fixtures.

Every function cites the reference file:line it restates (paths relative to the reference tree).
"""
from __future__ import annotations

import math
from typing import Callable, Dict, List, Optional

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# ----------------------------------------------------------------------------------------------
# attention / transformer block   (comfy/ldm/modules/attention.py)
# ----------------------------------------------------------------------------------------------
def attention(q: Tensor, k: Tensor, v: Tensor, heads: int, mask: Optional[Tensor] = None) -> Tensor:
    """attention_basic / attention_pytorch (attention.py:88-140, 337-349):
    q [B,Nq,H*d], k/v [B,Nk,H*d] -> softmax(q k^T / sqrt(d) (+ mask)) v, heads interleaved in the last dim.
    mask (attention.py:111-123): bool keep-mask -> masked_fill(~mask, -finfo.max); float -> added to the scaled scores;
    2-D [Nq,Nk], 3-D [bs,Nq,Nk] or broadcastable to [B,H,Nq,Nk]."""
    b, nq, inner = q.shape
    d = inner // heads
    qh = q.float().reshape(b, nq, heads, d).permute(0, 2, 1, 3)
    kh = k.float().reshape(b, -1, heads, d).permute(0, 2, 1, 3)
    vh = v.float().reshape(b, -1, heads, d).permute(0, 2, 1, 3)
    sim = torch.matmul(qh, kh.transpose(-1, -2)) * (d ** -0.5)
    if mask is not None:
        m = mask
        if m.dim() == 2:
            m = m[None, None]
        elif m.dim() == 3:
            m = m[:, None]
        if m.dtype == torch.bool:
            sim = sim.masked_fill(~m.expand_as(sim), -torch.finfo(sim.dtype).max)
        else:
            sim = sim + m.float()
    out = torch.matmul(sim.softmax(dim=-1), vh)
    return out.permute(0, 2, 1, 3).reshape(b, nq, inner)


def layer_norm(x: Tensor, w: Tensor, b: Tensor, eps: float = 1e-5) -> Tensor:
    """ops.LayerNorm -> F.layer_norm, eps = torch default 1e-5 (comfy/ops.py:97-113)."""
    x = x.float()
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * w.float() + b.float()


def gelu_erf(x: Tensor) -> Tensor:
    """F.gelu default (exact erf), as used by GEGLU (attention.py:63)."""
    return 0.5 * x * (1.0 + torch.erf(x / math.sqrt(2.0)))


def linear(x: Tensor, w: Tensor, b: Optional[Tensor] = None) -> Tensor:
    """ops.Linear -> F.linear (comfy/ops.py:40-52)."""
    y = x.float() @ w.float().t()
    return y if b is None else y + b.float()


def conv3x3(x: Tensor, w: Tensor, b: Optional[Tensor] = None, residual: Optional[Tensor] = None) -> Tensor:
    """ops.Conv2d -> F.conv2d with kernel 3, stride 1, padding 1 (comfy/ops.py:67-80), as the ResBlock / Upsample use
    it (openaimodel.py:83-96, 177-232), plus ResBlock._forward's `skip_connection(x) + h` (:264)."""
    y = F.conv2d(x.float(), w.float(), None if b is None else b.float(), padding=1)
    return y if residual is None else y + residual.float()


def upsample2x_conv3x3(x: Tensor, w: Tensor, b: Optional[Tensor] = None) -> Tensor:
    """Upsample.forward with use_conv (openaimodel.py:83-96): nearest 2x interpolation, then the 3x3 convolution."""
    return conv3x3(F.interpolate(x.float(), scale_factor=2, mode="nearest"), w, b)


def geglu(x: Tensor, w: Tensor, b: Tensor) -> Tensor:
    """GEGLU.forward: proj -> chunk(2) -> first half * gelu(second half) (attention.py:61-63)."""
    h = linear(x, w, b)
    val, gate = h.chunk(2, dim=-1)
    return val * gelu_erf(gate)


def feed_forward(x: Tensor, sd: Dict[str, Tensor], p: str) -> Tensor:
    """FeedForward with glu=True: GEGLU -> Dropout(0) -> Linear (attention.py:66-83)."""
    h = geglu(x, sd[p + '.net.0.proj.weight'], sd[p + '.net.0.proj.bias'])
    return linear(h, sd[p + '.net.2.weight'], sd[p + '.net.2.bias'])


def cross_attention(x: Tensor, context: Optional[Tensor], sd: Dict[str, Tensor], p: str, heads: int) -> Tensor:
    """CrossAttention.forward: to_q/to_k/to_v (no bias) -> attention -> to_out[0] (attention.py:401-415)."""
    ctx = x if context is None else context
    q = linear(x, sd[p + '.to_q.weight'])
    k = linear(ctx, sd[p + '.to_k.weight'])
    v = linear(ctx, sd[p + '.to_v.weight'])
    return linear(attention(q, k, v, heads), sd[p + '.to_out.0.weight'], sd[p + '.to_out.0.bias'])


def basic_transformer_block(x: Tensor, context: Tensor, sd: Dict[str, Tensor], p: str, heads: int) -> Tensor:
    """BasicTransformerBlock._forward without patches (attention.py:462-574):
    x += attn1(norm1(x)); x += attn2(norm2(x), ctx); x = ff(norm3(x)) + x."""
    n = layer_norm(x, sd[p + '.norm1.weight'], sd[p + '.norm1.bias'])
    x = x + cross_attention(n, None, sd, p + '.attn1', heads)
    n = layer_norm(x, sd[p + '.norm2.weight'], sd[p + '.norm2.bias'])
    x = x + cross_attention(n, context, sd, p + '.attn2', heads)
    n = layer_norm(x, sd[p + '.norm3.weight'], sd[p + '.norm3.bias'])
    return feed_forward(n, sd, p + '.ff') + x


def spatial_transformer(x: Tensor, context: Tensor, sd: Dict[str, Tensor], p: str, heads: int, depth: int,
                        use_linear: bool) -> Tensor:
    """SpatialTransformer.forward (attention.py:619-639): GroupNorm(32, eps 1e-6) -> proj_in ->
    b c h w -> b (h w) c -> blocks -> proj_out -> back -> + x_in. proj are 1x1 convs when not use_linear."""
    b, c, h, w = x.shape
    x_in = x
    x = F.group_norm(x.float(), 32, sd[p + '.norm.weight'].float(), sd[p + '.norm.bias'].float(), eps=1e-6)
    if not use_linear:
        x = F.conv2d(x, sd[p + '.proj_in.weight'].float(), sd[p + '.proj_in.bias'].float())
    x = x.permute(0, 2, 3, 1).reshape(b, h * w, -1)
    if use_linear:
        x = linear(x, sd[p + '.proj_in.weight'], sd[p + '.proj_in.bias'])
    for d in range(depth):
        x = basic_transformer_block(x, context, sd, f'{p}.transformer_blocks.{d}', heads)
    if use_linear:
        x = linear(x, sd[p + '.proj_out.weight'], sd[p + '.proj_out.bias'])
    x = x.reshape(b, h, w, -1).permute(0, 3, 1, 2)
    if not use_linear:
        x = F.conv2d(x, sd[p + '.proj_out.weight'].float(), sd[p + '.proj_out.bias'].float())
    return x + x_in


# ----------------------------------------------------------------------------------------------
# UNet host   (comfy/ldm/modules/diffusionmodules/openaimodel.py, util.py)
# ----------------------------------------------------------------------------------------------
def timestep_embedding(t: Tensor, dim: int, max_period: int = 10000) -> Tensor:
    """util.py:229-249."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(0, half, dtype=torch.float32) / half)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def res_block(x: Tensor, emb: Tensor, sd: Dict[str, Tensor], p: str) -> Tensor:
    """ResBlock._forward, no up/down, no scale-shift (openaimodel.py:235-264)."""
    f = lambda k: sd[p + k].float()
    h = F.group_norm(x, 32, f('.in_layers.0.weight'), f('.in_layers.0.bias'), eps=1e-5)
    h = F.conv2d(F.silu(h), f('.in_layers.2.weight'), f('.in_layers.2.bias'), padding=1)
    e = linear(F.silu(emb), sd[p + '.emb_layers.1.weight'], sd[p + '.emb_layers.1.bias'])
    h = h + e[:, :, None, None]
    h = F.group_norm(h, 32, f('.out_layers.0.weight'), f('.out_layers.0.bias'), eps=1e-5)
    h = F.conv2d(F.silu(h), f('.out_layers.3.weight'), f('.out_layers.3.bias'), padding=1)
    if (p + '.skip_connection.weight') in sd:
        x = F.conv2d(x, f('.skip_connection.weight'), f('.skip_connection.bias'))
    return x + h


def _unet_walk(cfg: dict):
    """The constructor's traversal (openaimodel.py:609-810) as a list of layer descriptors."""
    mc = cfg['model_channels']
    mult = cfg['channel_mult']
    nrb = cfg['num_res_blocks']
    nrb = [nrb] * len(mult) if isinstance(nrb, int) else list(nrb)
    td = list(cfg['transformer_depth'])
    tdo = list(cfg['transformer_depth_output'])

    def heads_of(ch):
        if cfg.get('num_head_channels', -1) == -1:
            return cfg['num_heads']
        return ch // cfg['num_head_channels']

    inp, out = [], []
    ch = mc
    idx = 1
    for level, m in enumerate(mult):
        for _ in range(nrb[level]):
            ch = m * mc
            depth = td.pop(0)
            inp.append([('res', f'input_blocks.{idx}.0')] +
                       ([('attn', f'input_blocks.{idx}.1', heads_of(ch), depth)] if depth > 0 else []))
            idx += 1
        if level != len(mult) - 1:
            inp.append([('down', f'input_blocks.{idx}.0')])
            idx += 1
    mid = [('res', 'middle_block.0')]
    if cfg['transformer_depth_middle'] is not None and cfg['transformer_depth_middle'] >= 0:
        mid += [('attn', 'middle_block.1', heads_of(ch), cfg['transformer_depth_middle']), ('res', 'middle_block.2')]
    idx = 0
    for level, m in list(enumerate(mult))[::-1]:
        for i in range(nrb[level] + 1):
            ch = mc * m
            layers = [('res', f'output_blocks.{idx}.0')]
            depth = tdo.pop()
            if depth > 0:
                layers.append(('attn', f'output_blocks.{idx}.{len(layers)}', heads_of(ch), depth))
            if level and i == nrb[level]:
                layers.append(('up', f'output_blocks.{idx}.{len(layers)}'))
            out.append(layers)
            idx += 1
    return inp, mid, out


def unet_forward(cfg: dict, sd: Dict[str, Tensor], x: Tensor, timesteps: Tensor, context: Tensor,
                 y: Optional[Tensor] = None) -> Tensor:
    """UNetModel.forward (openaimodel.py:819-890), no control / patches."""
    use_linear = cfg.get('use_linear_in_transformer', False)
    inp, mid, out = _unet_walk(cfg)
    emb = timestep_embedding(timesteps, cfg['model_channels'])
    emb = linear(F.silu(linear(emb, sd['time_embed.0.weight'], sd['time_embed.0.bias'])),
                 sd['time_embed.2.weight'], sd['time_embed.2.bias'])
    if cfg.get('adm_in_channels'):
        e = linear(F.silu(linear(y, sd['label_emb.0.0.weight'], sd['label_emb.0.0.bias'])),
                   sd['label_emb.0.2.weight'], sd['label_emb.0.2.bias'])
        emb = emb + e

    def run(layers, h, output_shape=None):
        for layer in layers:
            kind, p = layer[0], layer[1]
            if kind == 'res':
                h = res_block(h, emb, sd, p)
            elif kind == 'attn':
                h = spatial_transformer(h, context.float(), sd, p, layer[2], layer[3], use_linear)
            elif kind == 'down':
                h = F.conv2d(h, sd[p + '.op.weight'].float(), sd[p + '.op.bias'].float(), stride=2, padding=1)
            elif kind == 'up':
                size = [h.shape[2] * 2, h.shape[3] * 2] if output_shape is None else list(output_shape[2:])
                h = F.interpolate(h, size=size, mode='nearest')
                h = F.conv2d(h, sd[p + '.conv.weight'].float(), sd[p + '.conv.bias'].float(), padding=1)
        return h

    h = x.float()
    hs = []
    h = F.conv2d(h, sd['input_blocks.0.0.weight'].float(), sd['input_blocks.0.0.bias'].float(), padding=1)
    hs.append(h)
    for layers in inp:
        h = run(layers, h)
        hs.append(h)
    h = run(mid, h)
    for layers in out:
        h = torch.cat([h, hs.pop()], dim=1)
        h = run(layers, h, hs[-1].shape if hs else None)
    h = F.group_norm(h, 32, sd['out.0.weight'].float(), sd['out.0.bias'].float(), eps=1e-5)
    return F.conv2d(F.silu(h), sd['out.2.weight'].float(), sd['out.2.bias'].float(), padding=1)


# ----------------------------------------------------------------------------------------------
# model sampling   (comfy/model_sampling.py)
# ----------------------------------------------------------------------------------------------
def _sigmas_f64(linear_start: float, linear_end: float, n: int) -> Tensor:
    betas = torch.linspace(linear_start ** 0.5, linear_end ** 0.5, n, dtype=torch.float64) ** 2
    ac = torch.cumprod(1.0 - betas, dim=0)
    return ((1 - ac) / ac) ** 0.5


def discrete_sigmas(linear_start: float = 0.00085, linear_end: float = 0.012, n: int = 1000) -> Tensor:
    """ModelSamplingDiscrete._register_schedule, 'linear' betas in float64 -> float32 sigmas
    (model_sampling.py:53-76; util.py:89-93)."""
    return _sigmas_f64(linear_start, linear_end, n).float()


def discrete_log_sigmas(linear_start: float = 0.00085, linear_end: float = 0.012, n: int = 1000) -> Tensor:
    """set_sigmas: the log table is taken in float64 and then rounded (model_sampling.py:74-76)."""
    return _sigmas_f64(linear_start, linear_end, n).log().float()


def timestep_of_sigma(sigma: Tensor, log_sigmas: Tensor) -> Tensor:
    """ModelSamplingDiscrete.timestep: argmin |log sigma - log sigma_j| (model_sampling.py:86-89)."""
    dists = sigma.log() - log_sigmas[:, None]
    return dists.abs().argmin(dim=0).view(sigma.shape)


def sigma_of_timestep(t: Tensor, log_sigmas: Tensor) -> Tensor:
    """ModelSamplingDiscrete.sigma: log-linear interpolation (model_sampling.py:91-97)."""
    t = torch.clamp(t.float(), min=0, max=len(log_sigmas) - 1)
    lo, hi, w = t.floor().long(), t.ceil().long(), t.frac()
    return ((1 - w) * log_sigmas[lo] + w * log_sigmas[hi]).exp()


def calculate_input(sigma: Tensor, x: Tensor, sigma_data: float = 1.0) -> Tensor:
    """EPS.calculate_input (model_sampling.py:6-8)."""
    s = sigma.view(sigma.shape[:1] + (1,) * (x.ndim - 1))
    return x / (s ** 2 + sigma_data ** 2) ** 0.5


def calculate_denoised(sigma: Tensor, model_output: Tensor, model_input: Tensor, prediction: str,
                       sigma_data: float = 1.0) -> Tensor:
    """EPS / V_PREDICTION.calculate_denoised (model_sampling.py:10-12, 27-29)."""
    s = sigma.view(sigma.shape[:1] + (1,) * (model_output.ndim - 1))
    if prediction == 'eps':
        return model_input - model_output * s
    return model_input * sigma_data ** 2 / (s ** 2 + sigma_data ** 2) \
        - model_output * s * sigma_data / (s ** 2 + sigma_data ** 2) ** 0.5


def noise_scaling(sigma: Tensor, noise: Tensor, latent: Tensor, max_denoise: bool) -> Tensor:
    """EPS.noise_scaling (model_sampling.py:14-21)."""
    noise = noise * torch.sqrt(1.0 + sigma ** 2.0) if max_denoise else noise * sigma
    return noise + latent


# ----------------------------------------------------------------------------------------------
# schedules   (comfy/k_diffusion/sampling.py:16-22, comfy/samplers.py:277-313, 663-678)
# ----------------------------------------------------------------------------------------------
def get_sigmas_karras(n: int, sigma_min: float, sigma_max: float, rho: float = 7.0) -> Tensor:
    ramp = torch.linspace(0, 1, n)
    min_inv_rho = sigma_min ** (1 / rho)
    max_inv_rho = sigma_max ** (1 / rho)
    sigmas = (max_inv_rho + ramp * (min_inv_rho - max_inv_rho)) ** rho
    return torch.cat([sigmas, sigmas.new_zeros([1])])


def normal_scheduler(sigmas_table: Tensor, log_sigmas: Tensor, steps: int, sgm: bool = False) -> Tensor:
    start = timestep_of_sigma(sigmas_table[-1], log_sigmas)
    end = timestep_of_sigma(sigmas_table[0], log_sigmas)
    ts = torch.linspace(start, end, steps + 1)[:-1] if sgm else torch.linspace(start, end, steps)
    sigs = [float(sigma_of_timestep(t, log_sigmas)) for t in ts]
    return torch.FloatTensor(sigs + [0.0])


def simple_scheduler(sigmas_table: Tensor, steps: int) -> Tensor:
    ss = len(sigmas_table) / steps
    sigs = [float(sigmas_table[-(1 + int(x * ss))]) for x in range(steps)]
    return torch.FloatTensor(sigs + [0.0])


def calculate_sigmas(sigmas_table: Tensor, scheduler: str, steps: int, log_sigmas: Optional[Tensor] = None) -> Tensor:
    log_sigmas = discrete_log_sigmas() if log_sigmas is None else log_sigmas
    if scheduler == 'karras':
        return get_sigmas_karras(steps, float(sigmas_table[0]), float(sigmas_table[-1]))
    if scheduler == 'normal':
        return normal_scheduler(sigmas_table, log_sigmas, steps)
    if scheduler == 'sgm_uniform':
        return normal_scheduler(sigmas_table, log_sigmas, steps, sgm=True)
    if scheduler == 'simple':
        return simple_scheduler(sigmas_table, steps)
    raise ValueError(scheduler)


# ----------------------------------------------------------------------------------------------
# CFG + samplers   (comfy/samplers.py, comfy/k_diffusion/sampling.py)
# ----------------------------------------------------------------------------------------------
def cfg_function(cond_pred: Tensor, uncond_pred: Tensor, cond_scale: float) -> Tensor:
    """cfg_function default branch (samplers.py:240)."""
    return uncond_pred + (cond_pred - uncond_pred) * cond_scale


def lcm_pad_context(a: Tensor, b: Tensor):
    """CONDCrossAttn.concat: repeat-pad text lengths to their LCM (comfy/conds.py:50-63)."""
    n = abs(a.shape[1] * b.shape[1]) // math.gcd(a.shape[1], b.shape[1])
    return a.repeat(1, n // a.shape[1], 1), b.repeat(1, n // b.shape[1], 1)


def sampling_function(apply_model: Callable, x: Tensor, sigma: Tensor, uncond_ctx: Tensor, cond_ctx: Tensor,
                      cond_scale: float, extra_u: Optional[dict] = None, extra_c: Optional[dict] = None) -> Tensor:
    """sampling_function + calc_cond_batch for plain full-frame conds (samplers.py:131-259): one batched
    model call on [uncond x B, cond x B] (order cond_or_uncond == [1, 0]), then cfg_function."""
    b = x.shape[0]
    rep = lambda c: c if c.shape[0] == b else c.repeat(b, *([1] * (c.ndim - 1)))
    if math.isclose(cond_scale, 1.0):
        return apply_model(x, sigma, rep(cond_ctx), **{k: rep(v) for k, v in (extra_c or {}).items()})
    cu, cc = lcm_pad_context(rep(uncond_ctx), rep(cond_ctx))
    extras = {k: torch.cat([rep((extra_u or {})[k]), rep(v)]) for k, v in (extra_c or {}).items()}
    out = apply_model(torch.cat([x, x]), torch.cat([sigma, sigma]), torch.cat([cu, cc]), **extras)
    uncond_pred, cond_pred = out.chunk(2)
    return cfg_function(cond_pred, uncond_pred, cond_scale)


def sample_euler(model: Callable, x: Tensor, sigmas: Tensor, callback=None) -> Tensor:
    """sample_euler with s_churn = 0 (k_diffusion/sampling.py:126-144)."""
    s_in = x.new_ones([x.shape[0]])
    for i in range(len(sigmas) - 1):
        sigma_hat = sigmas[i]
        denoised = model(x, sigma_hat * s_in)
        d = (x - denoised) / sigma_hat
        if callback is not None:
            callback({'x': x, 'i': i, 'sigma': sigmas[i], 'sigma_hat': sigma_hat, 'denoised': denoised})
        x = x + d * (sigmas[i + 1] - sigma_hat)
    return x


def sample_dpmpp_2m(model: Callable, x: Tensor, sigmas: Tensor, callback=None) -> Tensor:
    """sample_dpmpp_2m (k_diffusion/sampling.py:570-593)."""
    s_in = x.new_ones([x.shape[0]])
    sigma_fn = lambda t: t.neg().exp()
    t_fn = lambda sigma: sigma.log().neg()
    old_denoised = None
    for i in range(len(sigmas) - 1):
        denoised = model(x, sigmas[i] * s_in)
        if callback is not None:
            callback({'x': x, 'i': i, 'sigma': sigmas[i], 'sigma_hat': sigmas[i], 'denoised': denoised})
        t, t_next = t_fn(sigmas[i]), t_fn(sigmas[i + 1])
        h = t_next - t
        if old_denoised is None or sigmas[i + 1] == 0:
            x = (sigma_fn(t_next) / sigma_fn(t)) * x - (-h).expm1() * denoised
        else:
            h_last = t - t_fn(sigmas[i - 1])
            r = h_last / h
            denoised_d = (1 + 1 / (2 * r)) * denoised - (1 / (2 * r)) * old_denoised
            x = (sigma_fn(t_next) / sigma_fn(t)) * x - (-h).expm1() * denoised_d
        old_denoised = denoised
    return x


SAMPLERS = {'euler': sample_euler, 'dpmpp_2m': sample_dpmpp_2m}


# ----------------------------------------------------------------------------------------------
# end to end   (comfy/sample.py, comfy/samplers.py:517-536, 615-650, comfy/model_base.py:74-98, 317-334)
# ----------------------------------------------------------------------------------------------
def prepare_noise(latent: Tensor, seed: int) -> Tensor:
    """comfy/sample.py:8-15: CPU generator, whole batch at once."""
    g = torch.manual_seed(seed)
    return torch.randn(latent.size(), dtype=latent.dtype, generator=g, device='cpu')


def sdxl_adm(pooled: Tensor, height: int, width: int) -> Tensor:
    """SDXL.encode_adm: pooled (1280) + 6 x Timestep(256) of (h, w, crop_h=0, crop_w=0, target_h, target_w)
    (model_base.py:317-334)."""
    emb = [timestep_embedding(torch.Tensor([v]), 256) for v in (height, width, 0, 0, height, width)]
    flat = torch.flatten(torch.cat(emb)).unsqueeze(0).repeat(pooled.shape[0], 1)
    return torch.cat((pooled, flat), dim=1)


def sample(cfg: dict, sd: Dict[str, Tensor], prediction: str, scale_factor: float, noise: Tensor, steps: int,
           cond_scale: float, sampler: str, scheduler: str, positive: Tensor, negative: Tensor,
           latent: Tensor, pooled_pos: Optional[Tensor] = None, pooled_neg: Optional[Tensor] = None,
           callback=None, max_steps: Optional[int] = None, denoise: Optional[float] = None,
           start_step: Optional[int] = None, last_step: Optional[int] = None, force_full_denoise: bool = False,
           noise_mask: Optional[Tensor] = None) -> Tensor:
    """comfy.sample.sample -> KSampler -> CFGGuider -> KSAMPLER -> sample_<name> with plain conds.
    denoise / start_step / last_step / force_full_denoise: KSampler.set_steps and KSampler.sample
    (comfy/samplers.py:716-755); noise_mask: KSamplerX0Inpaint (comfy/samplers.py:262-275)."""
    table = discrete_sigmas()
    log_table = discrete_log_sigmas()
    if denoise is None or denoise > 0.9999:
        sigmas = calculate_sigmas(table, scheduler, steps, log_table)
    elif denoise <= 0.0:
        return latent
    else:
        sigmas = calculate_sigmas(table, scheduler, int(steps / denoise), log_table)[-(steps + 1):]
    if max_steps is not None:
        sigmas = sigmas[:max_steps + 1]
    if last_step is not None and last_step < len(sigmas) - 1:
        sigmas = sigmas[:last_step + 1].clone()
        if force_full_denoise:
            sigmas[-1] = 0
    if start_step is not None:
        if start_step < len(sigmas) - 1:
            sigmas = sigmas[start_step:]
        else:
            return latent
    extra_c = extra_u = None
    if cfg.get('adm_in_channels'):
        h8, w8 = noise.shape[2] * 8, noise.shape[3] * 8
        extra_c = {'y': sdxl_adm(pooled_pos, h8, w8)}
        extra_u = {'y': sdxl_adm(pooled_neg, h8, w8)}

    def apply_model(x, sigma, context, y=None):
        xc = calculate_input(sigma, x)
        t = timestep_of_sigma(sigma, log_table).float()
        out = unet_forward(cfg, sd, xc, t, context, y)
        return calculate_denoised(sigma, out, x, prediction)

    def model(x, sigma):
        return sampling_function(apply_model, x, sigma, negative, positive, cond_scale, extra_u, extra_c)

    if torch.count_nonzero(latent) > 0:
        latent = latent * scale_factor
    max_denoise = math.isclose(float(table[-1]), float(sigmas[0]), rel_tol=1e-5) or float(sigmas[0]) > float(table[-1])
    x = noise_scaling(sigmas[0], noise, latent, max_denoise)
    denoiser = model
    if noise_mask is not None:
        keep = 1.0 - noise_mask

        def denoiser(x_in, sigma):       # KSamplerX0Inpaint: the kept region follows the noised input image
            shaped = sigma.reshape([sigma.shape[0]] + [1] * (noise.dim() - 1))
            x_in = x_in * noise_mask + noise_scaling(shaped, noise, latent, False) * keep
            return model(x_in, sigma) * noise_mask + latent * keep
    x = SAMPLERS[sampler](denoiser, x, sigmas, callback)
    return x / scale_factor
