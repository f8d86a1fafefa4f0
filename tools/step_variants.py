"""Step-time experiments on the bench workload (SDXL 1024^2, B=8): variants of host-side layout choices.
usage: python tools/step_variants.py [variant ...]   variants: base channels_last torch_baseline"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from comfy_gen_server_b200 import _native, samplers, unet_config as uc  # noqa: E402
from comfy_gen_server_b200.k_diffusion import sampling as ks  # noqa: E402
from comfy_gen_server_b200.model_base import build_model  # noqa: E402
from comfy_gen_server_b200.model_management import ModelPatcher  # noqa: E402


def make_stepper(model, patcher, dev, batch=8):
    fam = uc.FAMILIES["sdxl"]
    latent, noise, pos, neg = bench.synthetic_inputs(batch, fam["unet"]["context_dim"])
    guider = samplers.CFGGuider(patcher)
    guider.set_conds(pos, neg)
    guider.set_cfg(7.0)
    guider.conds = {k: [c.copy() for c in v] for k, v in guider.original_conds.items()}
    guider.inner_model = model
    guider.conds = samplers.process_conds(model, noise.to(dev), guider.conds, dev)
    sig = samplers.calculate_sigmas(model.model_sampling, "karras", 30)
    guider.prepare_schedule(sig)
    x = (noise * float(torch.sqrt(1.0 + sig[0] ** 2.0))).to(dev).contiguous()
    st = {"x": x, "old": None}

    def step(i):
        i = i % 30
        mode, a, b, c, d = ks.dpmpp_2m_coefficients(sig, i, st["old"] is not None and i > 0)
        res = guider.fused_step(st["x"], i, sig[i], mode, a, b, c, d, st["old"] if mode == 3 else None,
                                model_options=patcher.model_options)
        st["old"], xn = res
        st["x"] = x if i == 29 else xn
    return step


def timeit(step, warm=3, n=6):
    with torch.inference_mode():
        for i in range(warm):
            step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            step(warm + i)
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    variants = sys.argv[1:] or ["base", "channels_last", "torch_baseline"]
    dev = torch.device("cuda")
    model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
    patcher = ModelPatcher(model, dev)
    out = {}
    for v in variants:
        if v == "channels_last":
            model.diffusion_model.to(memory_format=torch.channels_last)
        step = make_stepper(model, patcher, dev)
        if v == "torch_baseline":
            from oracle.torch_baseline import torch_backend
            with torch_backend():
                ms = timeit(step)
        else:
            ms = timeit(step)
        out[v] = ms
        print(json.dumps({"variant": v, "ms_per_step": ms, "images_per_s": 8 / (30 * ms / 1e3)}), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "step_variants.json"), "w"))


if __name__ == "__main__":
    main()
