"""ORACLE (test infrastructure) -- the reference's GPU op sequence in plain PyTorch.

`torch_backend()` temporarily replaces the functions of comfy_gen_server_b200._native with the PyTorch
library calls the reference itself makes on a GPU (F.scaled_dot_product_attention, F.layer_norm,
F.linear -> chunk -> F.gelu -> mul, eager elementwise CFG / sampler arithmetic; SURVEY.md section 2b
rows K1-K9). Used only by tests and tools/bench to answer two questions the CPU oracle cannot:
  * how far is the reference's own bf16 GPU path from the fp32 reference (the inherent bf16 error)?
  * how fast is that library path on this B200 (the "bar to beat")?
Never imported by the product.
"""
from __future__ import annotations

import contextlib

import torch
import torch.nn.functional as F


def _attention(q, k, v, heads, mask=None):
    b, _, inner = q.shape
    d = inner // heads
    qh, kh, vh = (t.reshape(b, -1, heads, d).transpose(1, 2) for t in (q, k, v))
    if mask is not None and mask.dtype != torch.bool:
        mask = mask.to(q.dtype)
    out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=mask, dropout_p=0.0, is_causal=False)
    return out.transpose(1, 2).reshape(b, -1, inner)


def _layer_norm(x, w, b, eps=1e-5):
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def _add_layer_norm_(x, r, w, b, eps=1e-5):
    x += r
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def _geglu(x, w, b):
    h, gate = F.linear(x, w, b).chunk(2, dim=-1)
    return h * F.gelu(gate)


def _linear(x, w, b=None, residual=None, out=None):
    y = F.linear(x, w, b)
    if residual is not None:
        y = y + residual
    if out is not None:
        out.copy_(y)
        return out
    return y


def _unet_input(x, sigma, copies, sigma_data=1.0):
    s = sigma.view(-1, *([1] * (x.ndim - 1)))
    xc = (x / (s ** 2 + sigma_data ** 2) ** 0.5).to(torch.bfloat16)
    return torch.cat([xc] * copies)


def _denoise(x, out, sigma, pred_type, sigma_data):
    s = sigma.view(-1, *([1] * (x.ndim - 1)))
    out = out.float()
    if pred_type == 0:
        return x - out * s
    return x * sigma_data ** 2 / (s ** 2 + sigma_data ** 2) - out * s * sigma_data / (s ** 2 + sigma_data ** 2) ** 0.5


def _cfg_denoise(x, out_u, out_c, sigma, cond_scale, pred_type, sigma_data=1.0):
    dc = _denoise(x, out_c, sigma, pred_type, sigma_data)
    if out_u is None:
        return dc
    du = _denoise(x, out_u, sigma, pred_type, sigma_data)
    return du + (dc - du) * cond_scale


def _step_euler(x, denoised, sigma_hat, dt, out=None):
    return x + (x - denoised) / sigma_hat * dt


def _step_dpmpp2m(x, denoised, old, c_x, c_d, c_cur, c_old, out=None):
    d = denoised if old is None else c_cur * denoised - c_old * old
    return c_x * x - c_d * d


def _tail(x, out_u, out_c, sigma, cond_scale, pred_type, old, sampler, c0, c1, c2, c3, sigma_data=1.0, x_out=None):
    d = _cfg_denoise(x, out_u, out_c, sigma, cond_scale, pred_type, sigma_data)
    if sampler == 1:
        return d, _step_euler(x, d, c0, c1)
    return d, _step_dpmpp2m(x, d, old if sampler == 3 else None, c0, c1, c2, c3)


_PATCH = {"attention": _attention, "layer_norm": _layer_norm, "add_layer_norm_": _add_layer_norm_, "geglu": _geglu,
          "linear": _linear, "unet_input": _unet_input, "cfg_denoise": _cfg_denoise,
          "sampler_step_euler": _step_euler, "sampler_step_dpmpp2m": _step_dpmpp2m, "cfg_sampler_tail": _tail}


@contextlib.contextmanager
def torch_backend():
    from comfy_gen_server_b200 import _native
    saved = {k: getattr(_native, k) for k in _PATCH}
    try:
        for k, fn in _PATCH.items():
            setattr(_native, k, fn)
        yield
    finally:
        for k, fn in saved.items():
            setattr(_native, k, fn)
