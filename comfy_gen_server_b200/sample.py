"""Top-level sampling entry points with the reference's signatures (comfy/sample.py:8-44)."""
from __future__ import annotations

import numpy as np
import torch

from . import model_management, samplers


def prepare_noise(latent_image, seed, noise_inds=None):
    """Noise for the whole batch from the CPU generator seeded with `seed` (comfy/sample.py:8-25)."""
    generator = torch.manual_seed(seed)
    if noise_inds is None:
        return torch.randn(latent_image.size(), dtype=latent_image.dtype, layout=latent_image.layout,
                           generator=generator, device="cpu")
    unique_inds, inverse = np.unique(noise_inds, return_inverse=True)
    noises = []
    for i in range(unique_inds[-1] + 1):
        noise = torch.randn([1] + list(latent_image.size())[1:], dtype=latent_image.dtype,
                            layout=latent_image.layout, generator=generator, device="cpu")
        if i in unique_inds:
            noises.append(noise)
    noises = [noises[i] for i in inverse]
    return torch.cat(noises, axis=0)


def sample(model, noise, steps, cfg, sampler_name, scheduler, positive, negative, latent_image, denoise=1.0,
           disable_noise=False, start_step=None, last_step=None, force_full_denoise=False, noise_mask=None,
           sigmas=None, callback=None, disable_pbar=False, seed=None):
    sampler = samplers.KSampler(model, steps=steps, device=model.load_device, sampler=sampler_name,
                                scheduler=scheduler, denoise=denoise, model_options=model.model_options)
    out = sampler.sample(noise, positive, negative, cfg=cfg, latent_image=latent_image, start_step=start_step,
                         last_step=last_step, force_full_denoise=force_full_denoise, denoise_mask=noise_mask,
                         sigmas=sigmas, callback=callback, disable_pbar=disable_pbar, seed=seed)
    return out.to(model_management.intermediate_device())


def sample_custom(model, noise, cfg, sampler, sigmas, positive, negative, latent_image, noise_mask=None,
                  callback=None, disable_pbar=False, seed=None):
    out = samplers.sample(model, noise, positive, negative, cfg, model.load_device, sampler, sigmas,
                          model_options=model.model_options, latent_image=latent_image, denoise_mask=noise_mask,
                          callback=callback, disable_pbar=disable_pbar, seed=seed)
    return out.to(model_management.intermediate_device())


def sample_distributed(model, noise, steps, cfg, sampler_name, scheduler, positive, negative, latent_image,
                       cfg_split=None, group=None, callback=None, disable_pbar=True, seed=None):
    """Multi-GPU form of `sample` (one process per GPU, torch.distributed initialised): this rank denoises its
    slice of the batch -- and, in cfg_split placement, its half of the cond/uncond pair -- and every rank
    returns the full batch of final latents after one all-gather. `noise` and `latent_image` are FULL-batch
    tensors, identical on every rank. Without an initialised process group this is `sample`."""
    import torch.distributed as dist

    def run(noise_slice, latent_slice, placement):
        ks = samplers.KSampler(model, steps=steps, device=model.load_device, sampler=sampler_name,
                               scheduler=scheduler, model_options=model.model_options)
        if placement.mode == "cfg_split":
            guider = samplers.SplitCFGGuider(model, placement, model_management.pair_group(placement))
        else:
            guider = samplers.CFGGuider(model)
        guider.set_conds(positive, negative)
        guider.set_cfg(cfg)
        cb = callback if placement.rank == 0 else None
        return guider.sample(noise_slice, latent_slice, samplers.sampler_object(sampler_name), ks.sigmas,
                             callback=cb, disable_pbar=disable_pbar, seed=seed)

    out = model_management.distributed_sample(run, noise, latent_image, cfg_split=cfg_split, group=group)
    return out.to(model_management.intermediate_device())
