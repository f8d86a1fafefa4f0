"""Accuracy diagnostics on the GPU: this path vs the reference's fp32 golden, next to the reference's own
bf16 GPU op sequence (oracle/torch_baseline.py) vs the same golden. Writes gpurun_out/e2e_diag.json."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden, cosine  # noqa: E402
from comfy_gen_server_b200 import sample as S  # noqa: E402
from comfy_gen_server_b200.model_base import build_model  # noqa: E402
from comfy_gen_server_b200.model_management import ModelPatcher  # noqa: E402
from oracle.torch_baseline import torch_backend  # noqa: E402

CASES = [("tiny_sd15", "e2e_tiny_sd15_euler", 2, 16, 20, "euler", "normal"),
         ("tiny_sd21v", "e2e_tiny_sd21v_euler", 2, 12, 25, "euler", "normal"),
         ("tiny_sdxl", "e2e_tiny_sdxl_dpmpp2m", 2, 16, 30, "dpmpp_2m", "karras"),
         ("sd15", "e2e_sd15_config1", 1, 64, 20, "euler", "normal")]


def run(patcher, g, b, hw, steps, sampler, sched):
    seed = int(g["seed"])
    latent = torch.zeros(b, 4, hw, hw)
    noise = S.prepare_noise(latent, seed)
    den = []
    with torch.inference_mode():
        out = S.sample(patcher, noise, steps, 7.0, sampler, sched, [[g["positive"], {"pooled_output": g["pooled_pos"]}]],
                       [[g["negative"], {"pooled_output": g["pooled_neg"]}]], latent, seed=seed,
                       callback=lambda i, d, x, t: den.append(d.float().cpu()))
    return out, den


def main():
    res = []
    for fam, name, b, hw, steps, sampler, sched in CASES:
        if len(sys.argv) > 1 and fam not in sys.argv[1:]:
            continue
        g = load_golden(name)
        patcher = ModelPatcher(build_model(fam, dtype=torch.bfloat16, device="cuda", seed=int(g["seed"])), "cuda")
        ours, dours = run(patcher, g, b, hw, steps, sampler, sched)
        with torch_backend():
            base, dbase = run(patcher, g, b, hw, steps, sampler, sched)
        r = {"case": name,
             "ours_vs_fp32ref_first": cosine(dours[0], g["denoised_first"]), "ours_vs_fp32ref_final": cosine(ours, g["final"]),
             "torchbf16_vs_fp32ref_first": cosine(dbase[0], g["denoised_first"]),
             "torchbf16_vs_fp32ref_final": cosine(base, g["final"]),
             "ours_vs_torchbf16_first": cosine(dours[0], dbase[0]), "ours_vs_torchbf16_final": cosine(ours, base)}
        print(json.dumps(r), flush=True)
        res.append(r)
        del patcher
        torch.cuda.empty_cache()
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "e2e_diag.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
