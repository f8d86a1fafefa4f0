"""Prediction types and the discrete sigma table, with the reference's interface
(comfy/model_sampling.py:5-101). Host-side, tiny tensors; the per-element math of
calculate_input / calculate_denoised runs fused in csrc/sampler_tail.cu on the fast path, and the
methods here are the generic composition used when a hook forces the reference's step-by-step path.
"""
from __future__ import annotations

import torch


class EPS:
    prediction = "eps"
    pred_type = 0

    def calculate_input(self, sigma, noise):
        sigma = sigma.view(sigma.shape[:1] + (1,) * (noise.ndim - 1))
        return noise / (sigma ** 2 + self.sigma_data ** 2) ** 0.5

    def calculate_denoised(self, sigma, model_output, model_input):
        sigma = sigma.view(sigma.shape[:1] + (1,) * (model_output.ndim - 1))
        return model_input - model_output * sigma

    def noise_scaling(self, sigma, noise, latent_image, max_denoise=False):
        if max_denoise:
            noise = noise * torch.sqrt(1.0 + sigma ** 2.0)
        else:
            noise = noise * sigma
        noise += latent_image
        return noise

    def inverse_noise_scaling(self, sigma, latent):
        return latent


class V_PREDICTION(EPS):
    prediction = "v"
    pred_type = 1

    def calculate_denoised(self, sigma, model_output, model_input):
        sigma = sigma.view(sigma.shape[:1] + (1,) * (model_output.ndim - 1))
        return model_input * self.sigma_data ** 2 / (sigma ** 2 + self.sigma_data ** 2) \
            - model_output * sigma * self.sigma_data / (sigma ** 2 + self.sigma_data ** 2) ** 0.5


class ModelSamplingDiscrete(torch.nn.Module):
    """1000-entry sigma table from the linear-beta schedule, built in float64 and rounded once, exactly
    as the reference does (model_sampling.py:53-76) so schedules match bit for bit."""
    def __init__(self, sampling_settings=None):
        super().__init__()
        s = sampling_settings or {}
        linear_start = s.get("linear_start", 0.00085)
        linear_end = s.get("linear_end", 0.012)
        betas = torch.linspace(linear_start ** 0.5, linear_end ** 0.5, 1000, dtype=torch.float64) ** 2
        alphas_cumprod = torch.cumprod(1.0 - betas, dim=0)
        sigmas = ((1 - alphas_cumprod) / alphas_cumprod) ** 0.5
        self.num_timesteps = 1000
        self.linear_start = linear_start
        self.linear_end = linear_end
        self.sigma_data = 1.0
        self.set_sigmas(sigmas)

    def set_sigmas(self, sigmas):
        self.register_buffer('sigmas', sigmas.float())
        self.register_buffer('log_sigmas', sigmas.log().float())

    @property
    def sigma_min(self):
        return self.sigmas[0]

    @property
    def sigma_max(self):
        return self.sigmas[-1]

    def timestep(self, sigma):
        log_sigma = sigma.log()
        dists = log_sigma.to(self.log_sigmas.device) - self.log_sigmas[:, None]
        return dists.abs().argmin(dim=0).view(sigma.shape).to(sigma.device)

    def sigma(self, timestep):
        t = torch.clamp(timestep.float().to(self.log_sigmas.device), min=0, max=(len(self.sigmas) - 1))
        low_idx = t.floor().long()
        high_idx = t.ceil().long()
        w = t.frac()
        log_sigma = (1 - w) * self.log_sigmas[low_idx] + w * self.log_sigmas[high_idx]
        return log_sigma.exp().to(timestep.device)

    def percent_to_sigma(self, percent):
        if percent <= 0.0:
            return 999999999.9
        if percent >= 1.0:
            return 0.0
        percent = 1.0 - percent
        return self.sigma(torch.tensor(percent * 999.0)).item()


def model_sampling(prediction: str, sampling_settings=None):
    """comfy/model_base.py:30-45: a class combining the discrete schedule with the prediction type."""
    c = {"eps": EPS, "v": V_PREDICTION}[prediction]

    class ModelSampling(ModelSamplingDiscrete, c):
        pass

    return ModelSampling(sampling_settings)
