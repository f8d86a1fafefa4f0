"""Where the fixed cost of one sample() call goes (bench workload): cProfile of the host side of one call."""
import cProfile, io, os, pstats, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from comfy_gen_server_b200 import sample as S, unet_config as uc
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher

dev = torch.device("cuda")
model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
patcher = ModelPatcher(model, dev)
latent, noise, pos, neg = bench.synthetic_inputs(8, uc.FAMILIES["sdxl"]["unet"]["context_dim"])
run = lambda steps: S.sample(patcher, noise, steps, 7.0, "dpmpp_2m", "karras", pos, neg, latent, seed=1)
with torch.inference_mode():
    for _ in range(2):
        run(4)
    res = {}
    for steps in (2, 6, 10):
        torch.cuda.synchronize(); t0 = time.perf_counter(); run(steps); torch.cuda.synchronize()
        res[steps] = time.perf_counter() - t0
    print("wall seconds by step count:", {k: round(v, 4) for k, v in res.items()})
    per_step = (res[10] - res[2]) / 8
    print("per step %.1f ms, fixed cost per call %.1f ms" % (per_step * 1e3, (res[2] - 2 * per_step) * 1e3))
    pr = cProfile.Profile()
    pr.enable(); run(2); torch.cuda.synchronize(); pr.disable()
    s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue()[:6000])
