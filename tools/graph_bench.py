"""Small-batch sampling through the public sample() call with and without CUDA-graph replay of the denoise step.
usage: python tools/graph_bench.py"""
import json, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from comfy_gen_server_b200 import sample as S, unet_config as uc
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher

dev = torch.device("cuda")
res = []
for family, hw, batch, steps, sampler, sched in (("sd15", 64, 1, 20, "euler", "normal"), ("sdxl", 128, 1, 30, "dpmpp_2m", "karras"),
                                                 ("sdxl", 128, 8, 30, "dpmpp_2m", "karras")):
    model = build_model(family, dtype=torch.bfloat16, device=dev, seed=1234)
    patcher = ModelPatcher(model, dev)
    fam = uc.FAMILIES[family]
    bench.LATENT_HW = hw
    latent, noise, pos, neg = bench.synthetic_inputs(batch, fam["unet"]["context_dim"])
    if family == "sd15":
        pos = [[pos[0][0], {}]]; neg = [[neg[0][0], {}]]
    row = {"family": family, "latent": list(latent.shape), "steps": steps}
    for opt in (False, "auto", True):
        p2 = patcher.clone(); p2.model_options["cuda_graph"] = opt
        times = []
        with torch.inference_mode():
            for it in range(3):
                torch.cuda.synchronize(); t0 = time.perf_counter()
                out = S.sample(p2, noise, steps, 7.0, sampler, sched, pos, neg, latent, seed=1)
                torch.cuda.synchronize(); times.append(time.perf_counter() - t0)
        row[f"graph_{opt}_s"] = round(min(times[1:]), 4)
    print(json.dumps(row), flush=True); res.append(row)
    del model, patcher; torch.cuda.empty_cache()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "graph_bench.json"), "w"), indent=1)
