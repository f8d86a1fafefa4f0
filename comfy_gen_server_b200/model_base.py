"""Model wrapper with the reference's interface (comfy/model_base.py:47-98, 227-238, 317-334):
BaseModel.apply_model(x, t, c_concat, c_crossattn, control, transformer_options, **kwargs) -> denoised.

apply_model keeps the reference's semantics for any caller (hooks, wrappers, custom nodes); the fused
sampler fast path in .samplers calls `unet_raw` and finishes with one cgs_cfg_sampler_tail_fwd launch.
"""
from __future__ import annotations

import torch

from . import _native
from .conds import CONDCrossAttn, CONDRegular
from .model_sampling import model_sampling
from .ops import b200_ops, manual_cast
from .unet import UNetModel, timestep_embedding
from . import unet_config as uc


class LatentFormat:
    def __init__(self, scale_factor):
        self.scale_factor = scale_factor

    def process_in(self, latent):
        return latent * self.scale_factor

    def process_out(self, latent):
        return latent / self.scale_factor


class BaseModel(torch.nn.Module):
    def __init__(self, unet_config, prediction="eps", scale_factor=0.18215, dtype=torch.bfloat16, device=None,
                 manual_cast_dtype=None, unet_model=UNetModel):
        super().__init__()
        self.unet_config = dict(unet_config)
        self.latent_format = LatentFormat(scale_factor)
        self.manual_cast_dtype = manual_cast_dtype
        operations = manual_cast if manual_cast_dtype is not None else b200_ops
        cfg = {k: v for k, v in self.unet_config.items() if k != "dtype"}
        self.diffusion_model = unet_model(**cfg, dtype=dtype, device=device, operations=operations)
        self.prediction = prediction
        self.model_sampling = model_sampling(prediction)
        self.adm_channels = self.unet_config.get("adm_in_channels", None) or 0
        self.concat_keys = ()
        self.channels_last = False    # set by use_channels_last(): activations NHWC end to end

    def use_channels_last(self, enable=True):
        """Keep the UNet's activations (and conv weights) in torch.channels_last: cuDNN then runs its NHWC kernels
        without per-convolution layout conversions, GroupNorm uses the NHWC kernels and the SpatialTransformer's
        image <-> token rearranges become views. Requires every GroupNorm to have >= 8 channels per group."""
        self.channels_last = bool(enable)
        fmt = torch.channels_last if enable else torch.contiguous_format
        for p in self.diffusion_model.parameters():
            if p.dim() == 4:
                p.data = p.data.contiguous(memory_format=fmt)
        return self

    # -- reference semantics -------------------------------------------------------------------
    def apply_model(self, x, t, c_concat=None, c_crossattn=None, control=None, transformer_options={}, **kwargs):
        sigma = t
        dtype = self.get_dtype() if self.manual_cast_dtype is None else self.manual_cast_dtype
        if c_concat is None and x.is_cuda and x.dtype == torch.float32 and dtype == torch.bfloat16 \
                and x.is_contiguous() and (x.numel() // max(x.shape[0], 1)) % 4 == 0:
            xc = _native.unet_input(x, sigma.float().contiguous(), 1, float(self.model_sampling.sigma_data))
        else:
            xc = self.model_sampling.calculate_input(sigma, x)
            if c_concat is not None:
                xc = torch.cat([xc] + [c_concat], dim=1)
            xc = xc.to(dtype)
        if self.channels_last:
            xc = xc.contiguous(memory_format=torch.channels_last)
        t = self.model_sampling.timestep(t).float()
        context = c_crossattn.to(dtype)
        extra_conds = {}
        for o in kwargs:
            extra = kwargs[o]
            if hasattr(extra, "dtype") and extra.dtype != torch.int and extra.dtype != torch.long:
                extra = extra.to(dtype)
            extra_conds[o] = extra
        model_output = self.diffusion_model(xc, t, context=context, control=control,
                                            transformer_options=transformer_options, **extra_conds)
        if model_output.is_cuda and x.dtype == torch.float32 and x.is_contiguous() \
                and model_output.dtype in (torch.bfloat16, torch.float32) \
                and (x.numel() // max(x.shape[0], 1)) % 4 == 0:
            return _native.cfg_denoise(x, None, model_output.contiguous(), sigma.float().contiguous(), 1.0,
                                       self.model_sampling.pred_type, float(self.model_sampling.sigma_data))
        return self.model_sampling.calculate_denoised(sigma, model_output.float(), x)

    # -- fast path -----------------------------------------------------------------------------
    def unet_raw(self, xc, timesteps, context, y=None, transformer_options=None):
        """Raw UNet output (eps or v) for an already scaled / cast input batch."""
        kw = {} if y is None else {"y": y}
        if self.channels_last:
            xc = xc.contiguous(memory_format=torch.channels_last)
        return self.diffusion_model(xc, timesteps, context=context,
                                    transformer_options={} if transformer_options is None else transformer_options,
                                    **kw)

    def get_dtype(self):
        return self.diffusion_model.dtype

    def is_adm(self):
        return self.adm_channels > 0

    def encode_adm(self, **kwargs):
        return None

    def extra_conds(self, **kwargs):
        """comfy/model_base.py:107-160 for the text-to-image case: cross-attention context and ADM vector."""
        out = {}
        adm = self.encode_adm(**kwargs)
        if adm is not None:
            out['y'] = CONDRegular(adm)
        cross_attn = kwargs.get("cross_attn", None)
        if cross_attn is not None:
            out['c_crossattn'] = CONDCrossAttn(cross_attn)
        return out

    def process_latent_in(self, latent):
        return self.latent_format.process_in(latent)

    def process_latent_out(self, latent):
        return self.latent_format.process_out(latent)

    def memory_required(self, input_shape):
        """comfy/model_base.py:227-238 (flash-attention branch: linear in the token count)."""
        dtype = self.manual_cast_dtype if self.manual_cast_dtype is not None else self.get_dtype()
        area = input_shape[0] * input_shape[2] * input_shape[3]
        return (area * torch.finfo(dtype).bits // 8 / 50) * (1024 * 1024)


class SDXL(BaseModel):
    """comfy/model_base.py:311-334: ADM vector = pooled text (1280) + 6 sinusoidal size embeddings (256 each)."""
    def encode_adm(self, **kwargs):
        clip_pooled = kwargs["pooled_output"]
        width = kwargs.get("width", 768)
        height = kwargs.get("height", 768)
        crop_w = kwargs.get("crop_w", 0)
        crop_h = kwargs.get("crop_h", 0)
        target_width = kwargs.get("target_width", width)
        target_height = kwargs.get("target_height", height)
        out = [timestep_embedding(torch.Tensor([v]), 256)
               for v in (height, width, crop_h, crop_w, target_height, target_width)]
        flat = torch.flatten(torch.cat(out)).unsqueeze(dim=0).repeat(clip_pooled.shape[0], 1)
        return torch.cat((clip_pooled.to(flat.device), flat), dim=1)


def build_model(family: str, dtype=torch.bfloat16, device=None, state_dict=None, seed=None, channels_last=None):
    """Construct one of the BASELINE.json model families ('sd15', 'sd21v', 'sdxl', 'tiny_*') and load either
    a given UNet state dict or seeded random weights (there are no checkpoints in this setting)."""
    fam = uc.FAMILIES[family]
    cls = SDXL if fam["adm"] else BaseModel
    model = cls(fam["unet"], prediction=fam["prediction"], scale_factor=fam["scale_factor"], dtype=dtype,
                device="meta")
    shapes = uc.param_shapes(fam["unet"])
    own = model.diffusion_model.state_dict()
    assert {k: tuple(v.shape) for k, v in own.items()} == shapes, "UNet layout diverged from unet_config.param_shapes"
    model.diffusion_model.to_empty(device=device if device is not None else "cpu")
    if state_dict is None:
        if seed is None:
            raise ValueError("build_model needs a state_dict or a seed")
        # generate + upload tensor by tensor so a 2.6 B-parameter model never needs two fp32 copies on the host
        with torch.no_grad():
            for name, p in model.diffusion_model.state_dict().items():
                t = uc.seeded_state_dict(shapes, seed, names=[name])[name]
                p.copy_(t.to(dtype))
    else:
        model.diffusion_model.load_state_dict({k: v.to(dtype) for k, v in state_dict.items()})
    model.model_sampling.to("cpu")
    if channels_last is None:      # auto: every GroupNorm needs >= 8 channels per group (model_channels >= 256)
        channels_last = fam["unet"]["model_channels"] // 32 >= 8 and torch.device(device or "cpu").type == "cuda"
    if channels_last:
        model.use_channels_last(True)
    return model.eval()
