"""Random-init models of the BASELINE.json families built from the REFERENCE's own classes, for bench.py, smoke()
and the tests (there are no checkpoints in this setting): comfy.model_detection picks the model config from the
literal UNet config, comfy.model_base builds the BaseModel / SDXL wrapper, and the UNet is filled with the seeded
weights of unet_config.seeded_state_dict (the reference leaves parameters uninitialised, comfy/ops.py:41-42).
Nothing here is on the product path: a server loads its checkpoints through the reference as always.
"""
from __future__ import annotations

import torch

from . import unet_config as uc
from .reference import ensure_importable


def reference_model_config(family: str, dtype):
    ensure_importable()
    import comfy.model_detection as md
    import comfy.supported_models as sm
    fam = uc.FAMILIES[family]
    cfg = dict(fam["unet"])
    cfg["dtype"] = dtype
    mc = md.model_config_from_unet_config(dict(cfg))
    if mc is None:          # the tiny test configs are not in the reference's supported list: wrap the closest family
        base = {"tiny_sd15": sm.SD15, "tiny_sd21v": sm.SD20, "tiny_sdxl": sm.SDXL}[family]
        mc = base(dict(cfg))
    mc.set_inference_dtype(dtype, None)
    return mc


def build_model(family: str, device, seed=None, dtype=torch.bfloat16, state_dict=None):
    """-> (comfy.model_patcher.ModelPatcher, comfy.model_base.BaseModel) on `device`."""
    ensure_importable()
    import comfy.model_base as mb
    import comfy.model_patcher
    fam = uc.FAMILIES[family]
    mc = reference_model_config(family, dtype)
    dev = torch.device(device)
    if fam["prediction"] == "v":        # SD20.model_type() infers v-prediction from checkpoint statistics: state it
        model = mb.BaseModel(mc, model_type=mb.ModelType.V_PREDICTION, device=dev)
    elif fam["adm"]:
        model = mb.SDXL(mc, model_type=mb.ModelType.EPS, device=dev)
    else:
        model = mb.BaseModel(mc, model_type=mb.ModelType.EPS, device=dev)
    shapes = uc.param_shapes(fam["unet"])
    own = model.diffusion_model.state_dict()
    assert {k: tuple(v.shape) for k, v in own.items()} == shapes, "unet_config.param_shapes diverged from the reference UNet"
    with torch.no_grad():
        if state_dict is None:
            if seed is None:
                raise ValueError("build_model needs a state_dict or a seed")
            # generate + upload tensor by tensor so a 2.6 B-parameter model never needs two fp32 copies on the host
            for name, p in own.items():
                p.copy_(uc.seeded_state_dict(shapes, seed, names=[name])[name].to(dtype))
        else:
            model.diffusion_model.load_state_dict({k: v.to(dtype) for k, v in state_dict.items()})
    model.eval()
    patcher = comfy.model_patcher.ModelPatcher(model, load_device=dev, offload_device=dev)
    if dev.type == "cuda":
        # Bring the model into the state every sample() call after the first one sees: the reference's loader moves
        # the WHOLE BaseModel to the device, model_sampling's sigma table included, and from then on schedules are
        # computed with the device's log / exp (an ulp away from the host's). Doing it here keeps the first call
        # comparable with later ones.
        import comfy.model_management
        comfy.model_management.load_models_gpu([patcher])
    return patcher, model
