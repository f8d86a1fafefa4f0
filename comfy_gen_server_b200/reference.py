"""Find and import the reference tree this package plugs into.

The B200 path is a plug-in: it does not carry its own copy of the sampler / cond / model-wrapper code, it rebinds
the hot functions of an importable `comfy` package (integration.install). This module puts that package on
sys.path. Search order:

    1. `comfy` already importable (the server's own process: nothing to do)
    2. $CGS_REFERENCE_ROOT
    3. <repo>/baseline/_ref      (a staged copy, git-ignored; `python -m comfy_gen_server_b200.reference --stage`)
    4. /root/reference           (the build container)

The reference's sampler module imports `torchsde` at module level (comfy/k_diffusion/sampling.py:6) although only
its SDE samplers use it; when the package is absent an empty stand-in module is registered so the import succeeds.
"""
from __future__ import annotations

import importlib.util
import os
import shutil
import sys
import types

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_PKG_DIR)
STAGED = os.path.join(REPO_ROOT, "baseline", "_ref")
_CANDIDATES = (os.environ.get("CGS_REFERENCE_ROOT"), STAGED, "/root/reference")


class ReferenceNotFound(RuntimeError):
    pass


def locate() -> str | None:
    for c in _CANDIDATES:
        if c and os.path.isfile(os.path.join(c, "comfy", "samplers.py")):
            return c
    return None


def ensure_importable(cpu: bool = False) -> str:
    """Make `import comfy` work and return where it came from. `cpu=True` selects the reference's --cpu mode
    (must be called before comfy.model_management is first imported; used by the CPU baseline and golden scripts)."""
    where = None
    if "comfy" in sys.modules:
        where = os.path.dirname(os.path.dirname(os.path.abspath(sys.modules["comfy"].__file__ or "")))
    elif importlib.util.find_spec("comfy") is not None:
        where = "sys.path"
    else:
        where = locate()
        if where is None:
            raise ReferenceNotFound(
                "the `comfy` package of the reference server is not importable: set CGS_REFERENCE_ROOT to its tree, or "
                "stage a copy with `python -m comfy_gen_server_b200.reference --stage` (this package is a plug-in for "
                "that tree, not a replacement of it)")
        sys.path.insert(0, where)
    if "torchsde" not in sys.modules and importlib.util.find_spec("torchsde") is None:
        sys.modules["torchsde"] = types.ModuleType("torchsde")
    if cpu:
        if "comfy.model_management" in sys.modules:
            import comfy.model_management as mm
            if mm.get_torch_device().type != "cpu":
                raise RuntimeError("comfy.model_management was already imported in GPU mode")
        import comfy.cli_args
        comfy.cli_args.args.cpu = True
    return where


def stage(src: str = "/root/reference", dst: str = STAGED) -> str:
    """Copy the Python packages of the reference that the denoise path imports (comfy/, ~2 MB) next to this repo so
    that they travel to machines where the reference is not checked out. The copy is git-ignored, never edited."""
    if not os.path.isfile(os.path.join(src, "comfy", "samplers.py")):
        raise ReferenceNotFound(f"no reference tree at {src}")
    os.makedirs(dst, exist_ok=True)
    for name in ("comfy",):
        target = os.path.join(dst, name)
        if os.path.isdir(target):
            shutil.rmtree(target)
        shutil.copytree(os.path.join(src, name), target,
                        ignore=shutil.ignore_patterns("__pycache__", "*.pyc", "*.safetensors", "*.ckpt"))
    for name in ("LICENSE.txt",):
        if os.path.isfile(os.path.join(src, name)):
            shutil.copy2(os.path.join(src, name), os.path.join(dst, name))
    return dst


if __name__ == "__main__":
    if "--stage" in sys.argv:
        print("staged at", stage())
    else:
        print(locate())
