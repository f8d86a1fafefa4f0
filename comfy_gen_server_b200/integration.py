"""Install the B200 denoise path into the reference tree (the drop-in of INTEGRATION.md).

    import comfy_gen_server_b200.integration as b200
    b200.install()        # returns the list of rebound attributes
    ... comfy.sample.sample(...) / the server's KSampler nodes run on the sm_100a kernels ...
    b200.uninstall()      # restores the reference's own functions

Only module attributes and class methods that the reference resolves at call time are rebound; no reference file is
modified and no reference class is re-declared. Each replacement takes its fast path for the plain case and hands
everything else (hooks, patches, ControlNet, CPU / fp32 callers such as the text encoder and the VAE) to the function
it replaced, which is kept in layers.ORIGINAL.

The kernels compute in bf16. The reference's own default for the UNet on this class of GPU is fp16
(comfy/model_management.py:505-518); install() therefore switches on its `--bf16-unet` option, which the reference
reads at model-load time (model_management.py:508), so every UNet loaded afterwards is bf16. A UNet that is already
resident in fp16 keeps running on the reference's own path except for attention, which raises (there is no fp16
instantiation of the attention kernel).
"""
from __future__ import annotations

from typing import List, Tuple

from .reference import ensure_importable

_saved: List[Tuple[object, str, object]] = []
_added: List[Tuple[object, str]] = []
_prev_bf16 = None


def _rebind(obj, name, value, log, key=None):
    from .layers import ORIGINAL
    old = getattr(obj, name)
    _saved.append((obj, name, old))
    if key is not None:
        ORIGINAL[key] = old
    setattr(obj, name, value)
    log.append(f"{getattr(obj, '__name__', obj)}.{name}")


def installed() -> bool:
    return bool(_saved)


def install(bf16_unet: bool = True) -> List[str]:
    global _prev_bf16
    if _saved:
        return []
    ensure_importable()
    from . import _native
    _native.load_library()              # raises when the CUDA library is not built: there is no fallback
    _native.register_torch_ops()

    import comfy.cli_args
    import comfy.k_diffusion.sampling as ref_kd
    import comfy.ldm.modules.attention as ref_attn
    import comfy.ldm.modules.diffusionmodules.openaimodel as ref_unet
    import comfy.model_base as ref_mb
    import comfy.model_management as ref_mm
    import comfy.ops as ref_ops

    from . import block, guider, layers, placement, sampling, unet_path

    log: List[str] = []
    if bf16_unet:
        _prev_bf16 = comfy.cli_args.args.bf16_unet
        comfy.cli_args.args.bf16_unet = True
        log.append("comfy.cli_args.args.bf16_unet")
    # attention back-end: resolved by global name at call time inside CrossAttention.forward (attention.py:412-414)
    _rebind(ref_attn, "optimized_attention", layers.attention_b200, log, "optimized_attention")
    _rebind(ref_attn, "optimized_attention_masked", layers.attention_b200, log, "optimized_attention_masked")
    # GEGLU: Linear -> chunk -> gelu -> mul  ==> one GEMM with the gate in its epilogue (attention.py:61-63)
    _rebind(ref_attn.GEGLU, "forward", layers.geglu_forward, log, "GEGLU.forward")
    # operator classes (ops.py:82-113); manual_cast's classes inherit these methods
    _rebind(ref_ops.disable_weight_init.LayerNorm, "forward", layers.layernorm_forward, log, "LayerNorm.forward")
    _rebind(ref_ops.disable_weight_init.GroupNorm, "forward", layers.groupnorm_forward, log, "GroupNorm.forward")
    # transformer blocks (attention.py:462-574, 619-639)
    _rebind(ref_attn.BasicTransformerBlock, "_forward", block.transformer_block_forward, log,
            "BasicTransformerBlock._forward")
    _rebind(ref_attn.SpatialTransformer, "forward", block.spatial_transformer_forward, log,
            "SpatialTransformer.forward")
    # the UNet around them (openaimodel.py:60-96, 235-264, 819-890)
    _rebind(ref_unet.ResBlock, "_forward", unet_path.resblock_forward, log, "ResBlock._forward")
    _rebind(ref_unet.Upsample, "forward", unet_path.upsample_forward, log, "Upsample.forward")
    _rebind(ref_unet.UNetModel, "forward", unet_path.unet_forward, log, "UNetModel.forward")
    # model wrapper: sigma scaling + cast and eps / v -> denoised as one kernel each (model_base.py:74-98)
    _rebind(ref_mb.BaseModel, "apply_model", guider.apply_model_forward, log, "BaseModel.apply_model")
    # VAE decode (diffusionmodules/model.py:48-75, 139-160, 281-294): GroupNorm + SiLU, the 3x3 convolutions, up-sampling
    import comfy.ldm.modules.diffusionmodules.model as ref_vae
    from . import vae_path
    _rebind(ref_vae.ResnetBlock, "forward", vae_path.resnet_forward, log, "vae.ResnetBlock.forward")
    _rebind(ref_vae.AttnBlock, "forward", vae_path.attn_forward, log, "vae.AttnBlock.forward")
    _rebind(ref_vae.Upsample, "forward", vae_path.upsample_forward, log, "vae.Upsample.forward")
    # sampler functions: looked up with getattr(k_diffusion_sampling, "sample_" + name) (samplers.py:556); they
    # reach the fused CFG + update step (guider.FusedDenoise) through the guider behind the model they are given
    _rebind(ref_kd, "sample_euler", sampling.sample_euler, log, "sample_euler")
    _rebind(ref_kd, "sample_dpmpp_2m", sampling.sample_dpmpp_2m, log, "sample_dpmpp_2m")
    # multi-GPU placement published next to the reference's single-device helpers (model_management.py:74-88)
    for name in ("Placement", "plan_placement", "shard_range", "gather_latents", "distributed_sample",
                 "sample_distributed"):
        if not hasattr(ref_mm, name):
            setattr(ref_mm, name, getattr(placement, name))
            _added.append((ref_mm, name))
            log.append(f"comfy.model_management.{name} (new)")
    return log


def uninstall() -> None:
    global _prev_bf16
    from .layers import ORIGINAL
    had = bool(_saved)
    while _saved:
        obj, name, value = _saved.pop()
        setattr(obj, name, value)
    while _added:
        obj, name = _added.pop()
        if name in vars(obj):
            delattr(obj, name)
    ORIGINAL.clear()
    if had and _prev_bf16 is not None:
        import comfy.cli_args
        comfy.cli_args.args.bf16_unet = _prev_bf16
        _prev_bf16 = None
