"""The two sampler nodes every workflow JSON reaches this path through, with the reference's call contract
(nodes.py:1420-1470, 1753-1785): LATENT = {"samples": [B,4,h,w], optional "noise_mask", "batch_index"},
CONDITIONING = [[cond [1,77k,ctx], {"pooled_output": ...}], ...], returns a 1-tuple holding a LATENT dict.

Only the sampling nodes live here; loaders, VAE, CLIP, image and preview nodes are outside this path.
"""
from __future__ import annotations

import torch

from . import sample as _sample
from . import samplers as _samplers

MAX_SEED = 0xffffffffffffffff


def common_ksampler(model, seed, steps, cfg, sampler_name, scheduler, positive, negative, latent, denoise=1.0,
                    disable_noise=False, start_step=None, last_step=None, force_full_denoise=False, callback=None,
                    disable_pbar=True):
    """nodes.py:1420-1439. Noise comes from the CPU generator for the whole batch (or per `batch_index`), zeros when
    `disable_noise`; `noise_mask` selects the region to denoise; the input dict is copied, not modified."""
    latent_image = latent["samples"]
    if disable_noise:
        noise = torch.zeros(latent_image.size(), dtype=latent_image.dtype, layout=latent_image.layout, device="cpu")
    else:
        noise = _sample.prepare_noise(latent_image, seed, latent.get("batch_index"))
    samples = _sample.sample(model, noise, steps, cfg, sampler_name, scheduler, positive, negative, latent_image,
                             denoise=denoise, disable_noise=disable_noise, start_step=start_step, last_step=last_step,
                             force_full_denoise=force_full_denoise, noise_mask=latent.get("noise_mask"),
                             callback=callback, disable_pbar=disable_pbar, seed=seed)
    out = latent.copy()
    out["samples"] = samples
    return (out,)


def _sampler_inputs(seed_name, extra):
    spec = {"model": ("MODEL",)}
    spec.update(extra.get("before_seed", {}))
    spec[seed_name] = ("INT", {"default": 0, "min": 0, "max": MAX_SEED})
    spec["steps"] = ("INT", {"default": 20, "min": 1, "max": 10000})
    spec["cfg"] = ("FLOAT", {"default": 8.0, "min": 0.0, "max": 100.0, "step": 0.1, "round": 0.01})
    spec["sampler_name"] = (_samplers.KSampler.SAMPLERS,)
    spec["scheduler"] = (_samplers.KSampler.SCHEDULERS,)
    spec["positive"] = ("CONDITIONING",)
    spec["negative"] = ("CONDITIONING",)
    spec["latent_image"] = ("LATENT",)
    spec.update(extra.get("after", {}))
    return {"required": spec}


class KSampler:
    """nodes.py:1441-1465."""
    RETURN_TYPES = ("LATENT",)
    FUNCTION = "sample"
    CATEGORY = "sampling"

    @classmethod
    def INPUT_TYPES(cls):
        return _sampler_inputs("seed", {"after": {"denoise": ("FLOAT", {"default": 1.0, "min": 0.0, "max": 1.0,
                                                                         "step": 0.01})}})

    def sample(self, model, seed, steps, cfg, sampler_name, scheduler, positive, negative, latent_image, denoise=1.0):
        return common_ksampler(model, seed, steps, cfg, sampler_name, scheduler, positive, negative, latent_image,
                               denoise=denoise)


class KSamplerAdvanced:
    """nodes.py:1753-1785: a window [start_at_step, end_at_step) of the schedule, optionally without adding noise
    and optionally leaving the remaining noise in (for chained samplers)."""
    RETURN_TYPES = ("LATENT",)
    FUNCTION = "sample"
    CATEGORY = "sampling"

    @classmethod
    def INPUT_TYPES(cls):
        return _sampler_inputs("noise_seed", {
            "before_seed": {"add_noise": (["enable", "disable"],)},
            "after": {"start_at_step": ("INT", {"default": 0, "min": 0, "max": 10000}),
                      "end_at_step": ("INT", {"default": 10000, "min": 0, "max": 10000}),
                      "return_with_leftover_noise": (["disable", "enable"],)}})

    def sample(self, model, add_noise, noise_seed, steps, cfg, sampler_name, scheduler, positive, negative,
               latent_image, start_at_step, end_at_step, return_with_leftover_noise, denoise=1.0):
        return common_ksampler(model, noise_seed, steps, cfg, sampler_name, scheduler, positive, negative,
                               latent_image, denoise=denoise, disable_noise=(add_noise == "disable"),
                               start_step=start_at_step, last_step=end_at_step,
                               force_full_denoise=(return_with_leftover_noise != "enable"))


NODE_CLASS_MAPPINGS = {"KSampler": KSampler, "KSamplerAdvanced": KSamplerAdvanced}
