"""Run every BASELINE.json config through the public sample() API on one GPU (per-GPU share of the batch where
the config names 8 GPUs) and report images/s. usage: python tools/run_configs.py [cfg2 cfg3 cfg4 cfg5] [--steps N]"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from comfy_gen_server_b200 import sample as S, unet_config as uc  # noqa: E402
from comfy_gen_server_b200.model_base import build_model  # noqa: E402
from comfy_gen_server_b200.model_management import ModelPatcher  # noqa: E402

CONFIGS = {
    # name: (family, batch on this GPU, latent hw, steps, sampler, scheduler, note)
    "cfg2": ("sd15", 16, 64, 20, "euler", "normal", "SD1.5 512^2 batch 16, 20-step Euler CFG 7 (1 GPU)"),
    "cfg3": ("sdxl", 8, 128, 30, "dpmpp_2m", "karras", "SDXL 1024^2 batch 8, 30-step DPM++2M Karras CFG 7 (1 GPU)"),
    "cfg4": ("sd21v", 8, 96, 25, "euler", "normal", "SD2.1-v 768^2 batch 8, 25-step Euler CFG 7 (1 GPU)"),
    "cfg5": ("sdxl", 1, 256, 30, "dpmpp_2m", "karras", "SDXL 2048^2, 1 of 4 images (the 8-GPU share is half of this: CFG split)"),
}


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    steps_override = None
    if "--steps" in sys.argv:
        steps_override = int(sys.argv[sys.argv.index("--steps") + 1])
        args = [a for a in args if a != str(steps_override)]
    names = args or list(CONFIGS)
    dev = torch.device("cuda")
    out = []
    cache = {}
    for name in names:
        fam, batch, hw, steps, sampler, sched, note = CONFIGS[name]
        steps_run = steps_override or steps
        if fam not in cache:
            cache.clear()
            torch.cuda.empty_cache()
            cache[fam] = ModelPatcher(build_model(fam, dtype=torch.bfloat16, device=dev, seed=1), dev)
        patcher = cache[fam]
        ctx = uc.FAMILIES[fam]["unet"]["context_dim"]
        g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
        latent = torch.zeros(batch, 4, hw, hw)
        noise = S.prepare_noise(latent, 1)
        pos = [[g((1, 77, ctx), 1), {"pooled_output": g((1, 1280), 3)}]]
        neg = [[g((1, 77, ctx), 2), {"pooled_output": g((1, 1280), 4)}]]
        with torch.inference_mode():
            S.sample(patcher, noise, 2, 7.0, sampler, sched, pos, neg, latent, seed=1)      # warm-up
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            res = S.sample(patcher, noise, steps_run, 7.0, sampler, sched, pos, neg, latent, seed=1)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        r = {"config": name, "note": note, "steps_run": steps_run, "ms_per_step": dt / steps_run * 1e3,
             "images_per_s_at_full_schedule": batch / (dt / steps_run * steps), "finite": bool(torch.isfinite(res).all()),
             "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
        print(json.dumps(r), flush=True)
        out.append(r)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "configs.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
