"""sample_euler / sample_dpmpp_2m for the B200 path, with the reference's sampler-function contract
(comfy/k_diffusion/sampling.py:126-144, 570-593); integration.install() binds them as
comfy.k_diffusion.sampling.sample_euler / sample_dpmpp_2m, where comfy.samplers.ksampler() finds them by name
(samplers.py:556).

    sample_<name>(model, x, sigmas, extra_args=None, callback=None, disable=None, **opts) -> x
    model(x, sigma_vec, **extra_args) -> denoised
    callback({'x', 'i', 'sigma', 'sigma_hat', 'denoised'}) every step

Differences in mechanism, not in results: the per-step scalar coefficients are computed on the host from one host
copy of `sigmas` (the reference's `sigmas[i + 1] == 0` / `s_tmin <= sigmas[i]` tests force a device->host sync every
step), and the update runs as one fused CUDA kernel (csrc/sampler_tail.cu). When the guider behind `model` is a stock
CFG guider on plain conds (guider.FusedDenoise) the CFG combine and the update are a single launch after the UNet.
Latents that are not contiguous fp32 CUDA tensors are handed to the reference's own sampler function.
"""
from __future__ import annotations

import torch

from . import _native
from .guider import FusedDenoise
from .layers import ORIGINAL, is_latent as _native_ok


def _host(sigmas):
    """One host copy of the schedule (fp32), so the loop never syncs on device scalars."""
    return sigmas.detach().to("cpu", torch.float32)


def dpmpp_2m_coefficients(sig, i, have_old):
    """(mode, c_x, c_d, c_cur, c_old) for step i, from the reference's expressions evaluated in fp32 on the
    host (k_diffusion/sampling.py:574-592). mode 2: x' = c_x x - c_d D;  mode 3: x' = c_x x - c_d (c_cur D - c_old D_old)."""
    sigma_fn = lambda t: t.neg().exp()
    t_fn = lambda sigma: sigma.log().neg()
    t, t_next = t_fn(sig[i]), t_fn(sig[i + 1])
    h = t_next - t
    c_x = sigma_fn(t_next) / sigma_fn(t)
    c_d = (-h).expm1()
    if not have_old or sig[i + 1] == 0:
        return 2, float(c_x), float(c_d), 0.0, 0.0
    h_last = t - t_fn(sig[i - 1])
    r = h_last / h
    return 3, float(c_x), float(c_d), float(1 + 1 / (2 * r)), float(1 / (2 * r))


@torch.no_grad()
def sample_euler(model, x, sigmas, extra_args=None, callback=None, disable=None, s_churn=0., s_tmin=0.,
                 s_tmax=float('inf'), s_noise=1.):
    """Algorithm 2 (Euler steps) of Karras et al. (2022)."""
    if not _native_ok(x):
        return ORIGINAL["sample_euler"](model, x, sigmas, extra_args=extra_args, callback=callback, disable=disable,
                                        s_churn=s_churn, s_tmin=s_tmin, s_tmax=s_tmax, s_noise=s_noise)
    extra_args = {} if extra_args is None else extra_args
    sig = _host(sigmas)
    n = len(sig) - 1
    s_in = x.new_ones([x.shape[0]])
    fused = FusedDenoise.create(model, x, extra_args, sig) if s_churn == 0 else None
    for i in range(n):
        gamma = min(s_churn / n, 2 ** 0.5 - 1) if s_tmin <= float(sig[i]) <= s_tmax else 0.
        sigma_hat = sig[i] * (gamma + 1)
        if gamma > 0:
            eps = torch.randn_like(x) * s_noise
            x = x + eps * float((sigma_hat ** 2 - sig[i] ** 2) ** 0.5)
        dt = sig[i + 1] - sigma_hat
        if fused is not None:
            denoised, x_next = fused.step(x, i, sigma_hat, 1, float(sigma_hat), float(dt), 0.0, 0.0, None)
        else:
            denoised = model(x, float(sigma_hat) * s_in, **extra_args)
            if not _native_ok(denoised):
                raise RuntimeError("sample_euler: the model must return a contiguous fp32 CUDA tensor")
            x_next = _native.sampler_step_euler(x, denoised, float(sigma_hat), float(dt))
        if callback is not None:
            callback({'x': x, 'i': i, 'sigma': sig[i], 'sigma_hat': sigma_hat, 'denoised': denoised})
        x = x_next
    return x


@torch.no_grad()
def sample_dpmpp_2m(model, x, sigmas, extra_args=None, callback=None, disable=None):
    """DPM-Solver++(2M)."""
    if not _native_ok(x):
        return ORIGINAL["sample_dpmpp_2m"](model, x, sigmas, extra_args=extra_args, callback=callback, disable=disable)
    extra_args = {} if extra_args is None else extra_args
    sig = _host(sigmas)
    s_in = x.new_ones([x.shape[0]])
    old_denoised = None
    fused = FusedDenoise.create(model, x, extra_args, sig)
    for i in range(len(sig) - 1):
        mode, c_x, c_d, c_cur, c_old = dpmpp_2m_coefficients(sig, i, old_denoised is not None)
        if fused is not None:
            denoised, x_next = fused.step(x, i, sig[i], mode, c_x, c_d, c_cur, c_old,
                                          old_denoised if mode == 3 else None)
        else:
            denoised = model(x, float(sig[i]) * s_in, **extra_args)
            if not _native_ok(denoised):
                raise RuntimeError("sample_dpmpp_2m: the model must return a contiguous fp32 CUDA tensor")
            x_next = _native.sampler_step_dpmpp2m(x, denoised, old_denoised if mode == 3 else None,
                                                  c_x, c_d, c_cur, c_old)
        if callback is not None:
            callback({'x': x, 'i': i, 'sigma': sig[i], 'sigma_hat': sig[i], 'denoised': denoised})
        x = x_next
        old_denoised = denoised
    return x
