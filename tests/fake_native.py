"""TEST INFRASTRUCTURE: torch (CPU, fp32) stand-ins for the functions of comfy_gen_server_b200._native, so that the
host side of the product -- the rebinding done by integration.install(), the fast-path predicates, the fused step's
conditioning batch / tables / CFG-half bookkeeping, the placement code -- can be exercised end to end on a machine
without a GPU and compared with the reference's golden outputs at fp32 accuracy.

Nothing here is shipped or measured: the product's _native has no CPU path and raises for CPU tensors
(tests/test_cpu_host.py::test_product_refuses_cpu_tensors). `patched()` swaps the stand-ins in for the duration of a
test and tells layers.py that "the device" is the CPU and the activation type fp32.
"""
from __future__ import annotations

import contextlib

import torch
import torch.nn.functional as F

CALLS: dict = {}


def _count(name):
    CALLS[name] = CALLS.get(name, 0) + 1


def attention(q, k, v, heads, mask=None):
    _count("attention")
    b, nq, inner = q.shape
    d = inner // heads
    qh, kh, vh = (t.reshape(b, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    m = mask
    if m is not None:
        if m.dim() == 2:
            m = m[None, None]
        elif m.dim() == 3:
            m = m[:, None]
    return F.scaled_dot_product_attention(qh, kh, vh, attn_mask=m).transpose(1, 2).reshape(b, nq, inner)


def layer_norm(x, weight, bias, eps=1e-5):
    _count("layer_norm")
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def add_layer_norm_(x, residual, weight, bias, eps=1e-5):
    _count("add_layer_norm_")
    x += residual
    return F.layer_norm(x, (x.shape[-1],), weight, bias, eps)


def group_norm(x, weight, bias, groups, eps, silu=False, add_nc=None):
    _count("group_norm")
    if add_nc is not None:
        x = x + add_nc[..., None, None]
    y = F.group_norm(x, groups, weight, bias, eps)
    return F.silu(y) if silu else y


def residual_add(a, b, bias=None):
    _count("residual_add")
    out = a + b
    return out if bias is None else out + bias.view(1, -1, *([1] * (out.dim() - 2)))


def cat_channels(a, b, bias_a=None):
    _count("cat_channels")
    if bias_a is not None:
        a = a + bias_a.view(1, -1, 1, 1)
    return torch.cat([a, b], dim=1)


def upsample_nearest2x(x):
    _count("upsample_nearest2x")
    return F.interpolate(x, scale_factor=2, mode="nearest")


def transpose_add(x, residual=None):
    _count("transpose_add")
    out = x.transpose(1, 2).contiguous()
    return out if residual is None else out + residual.reshape(out.shape)


def geglu(x, weight, bias):
    _count("geglu")
    a, g = F.linear(x, weight, bias).chunk(2, dim=-1)
    return a * F.gelu(g)


def linear(x, weight, bias=None, residual=None, out=None):
    _count("linear")
    y = F.linear(x, weight, bias)
    if residual is not None:
        y = y + residual
    if out is not None:
        out.copy_(y)
        return out
    return y


def conv3x3(x, weight, bias=None, residual=None):
    _count("conv3x3")
    y = F.conv2d(x, weight, bias, padding=1)
    return y if residual is None else y + residual


def upsample2x_conv_weights(weight):
    return weight


def upsample2x_conv3x3(x, w4, bias=None):
    _count("upsample2x_conv3x3")
    return F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), w4, bias, padding=1)


def softmax_rows_(x, scale=1.0):
    _count("softmax_rows")
    x.copy_(torch.softmax(x.float() * scale, dim=-1))
    return x


def unet_input(x, sigma, copies, sigma_data=1.0):
    _count("unet_input")
    s = sigma.view(-1, *([1] * (x.dim() - 1)))
    xc = x / (s ** 2 + sigma_data ** 2) ** 0.5
    return xc.repeat(copies, *([1] * (x.dim() - 1)))


def _denoised(x, out, sigma, pred_type, sigma_data):
    s = sigma.view(-1, *([1] * (x.dim() - 1)))
    out = out.float()
    if pred_type == 0:
        return x - out * s
    return x * sigma_data ** 2 / (s ** 2 + sigma_data ** 2) - out * s * sigma_data / (s ** 2 + sigma_data ** 2) ** 0.5


def cfg_denoise(x, model_out_uncond, model_out_cond, sigma, cond_scale, pred_type, sigma_data=1.0):
    _count("cfg_denoise")
    dc = _denoised(x, model_out_cond, sigma, pred_type, sigma_data)
    if model_out_uncond is None:
        return dc
    du = _denoised(x, model_out_uncond, sigma, pred_type, sigma_data)
    return du + (dc - du) * cond_scale


def sampler_step_euler(x, denoised, sigma_hat, dt, out=None):
    _count("sampler_step")
    return x + (x - denoised) / sigma_hat * dt


def sampler_step_dpmpp2m(x, denoised, old_denoised, c_x, c_d, c_cur, c_old, out=None):
    _count("sampler_step")
    f = lambda v: torch.tensor(v, dtype=torch.float32)
    d = denoised if old_denoised is None else f(c_cur) * denoised - f(c_old) * old_denoised
    return f(c_x) * x - f(c_d) * d


def cfg_sampler_tail(x, model_out_uncond, model_out_cond, sigma, cond_scale, pred_type, old_denoised, sampler,
                     c0, c1, c2, c3, sigma_data=1.0, x_out=None):
    _count("cfg_sampler_tail")
    den = cfg_denoise(x, model_out_uncond, model_out_cond, sigma, cond_scale, pred_type, sigma_data)
    if sampler == 1:
        xn = sampler_step_euler(x, den, c0, c1)
    else:
        xn = sampler_step_dpmpp2m(x, den, old_denoised if sampler == 3 else None, c0, c1, c2, c3)
    return den, xn


@contextlib.contextmanager
def patched():
    """Swap the stand-ins into _native and point layers.py at (cpu, fp32)."""
    from comfy_gen_server_b200 import _native, layers
    names = ["attention", "layer_norm", "add_layer_norm_", "group_norm", "residual_add", "cat_channels",
             "upsample_nearest2x", "transpose_add", "geglu", "linear", "conv3x3", "upsample2x_conv_weights", "upsample2x_conv3x3", "softmax_rows_", "unet_input", "cfg_denoise",
             "sampler_step_euler", "sampler_step_dpmpp2m", "cfg_sampler_tail"]
    saved = {n: getattr(_native, n) for n in names}
    saved_layers = (layers.ACT_DTYPE, layers.DEVICE_TYPE)
    saved_load, saved_reg = _native.load_library, _native.register_torch_ops
    try:
        for n in names:
            setattr(_native, n, globals()[n])
        layers.ACT_DTYPE, layers.DEVICE_TYPE = torch.float32, "cpu"
        _native.register_torch_ops = lambda: None
        CALLS.clear()
        yield CALLS
    finally:
        for n, f in saved.items():
            setattr(_native, n, f)
        layers.ACT_DTYPE, layers.DEVICE_TYPE = saved_layers
        _native.load_library, _native.register_torch_ops = saved_load, saved_reg
