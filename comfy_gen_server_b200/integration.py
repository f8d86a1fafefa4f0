"""Install the B200 path into an importable reference tree (the drop-in of INTEGRATION.md section 2).

    import comfy_gen_server_b200.integration as b200
    b200.install()        # after `import comfy...`; returns the list of rebound attributes
    b200.uninstall()      # restores the reference's own functions

Only module attributes and class methods the reference already resolves at call time are rebound; no reference
file is modified. After install(), any UNet built by the reference runs its attention, LayerNorm, GroupNorm,
GEGLU and Euler / DPM++ 2M updates on the CUDA kernels (bf16 CUDA tensors required; there is no fallback).
"""
from __future__ import annotations

from typing import List, Tuple

_saved: List[Tuple[object, str, object]] = []
_added: List[Tuple[object, str]] = []


def _rebind(obj, name, value, log):
    _saved.append((obj, name, getattr(obj, name)))
    setattr(obj, name, value)
    log.append(f"{getattr(obj, '__name__', obj)}.{name}")


def install() -> List[str]:
    import comfy.k_diffusion.sampling as ref_kd
    import comfy.ldm.modules.attention as ref_attn
    import comfy.ops as ref_ops

    from . import _native, attention as A
    from .k_diffusion import sampling as kd
    from .ops import b200_ops

    if _saved:
        return []
    _native.load_library()
    _native.register_torch_ops()
    log: List[str] = []
    # attention back-end: resolved by global name at call time inside CrossAttention.forward (attention.py:412-414)
    _rebind(ref_attn, "optimized_attention", A.attention_b200, log)
    _rebind(ref_attn, "optimized_attention_masked", A.attention_b200, log)
    # GEGLU: Linear -> chunk -> gelu -> mul  ==> one GEMM with the gate in its epilogue (attention.py:61-63)
    _rebind(ref_attn.GEGLU, "forward", A.GEGLU.forward, log)
    # operator classes (ops.py:97-113, 66-80): LayerNorm / GroupNorm forward on the CUDA kernels
    for ns in (ref_ops.disable_weight_init, ref_ops.manual_cast):
        _rebind(ns.LayerNorm, "forward", b200_ops.LayerNorm.forward, log)
        _rebind(ns.LayerNorm, "forward_comfy_cast_weights", b200_ops.LayerNorm.forward_comfy_cast_weights, log)
        if not hasattr(ns.LayerNorm, "_apply_ln"):
            _added.append((ns.LayerNorm, "_apply_ln"))
            ns.LayerNorm._apply_ln = b200_ops.LayerNorm._apply_ln
    # sampler functions: looked up with getattr(k_diffusion_sampling, "sample_" + name) (samplers.py:556)
    _rebind(ref_kd, "sample_euler", kd.sample_euler, log)
    _rebind(ref_kd, "sample_dpmpp_2m", kd.sample_dpmpp_2m, log)
    return log


def uninstall() -> None:
    while _saved:
        obj, name, value = _saved.pop()
        setattr(obj, name, value)
    while _added:
        obj, name = _added.pop()
        if name in vars(obj):
            delattr(obj, name)
