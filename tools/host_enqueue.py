"""Host time to enqueue one denoise step (GPU idle at the start, so the launch queue does not throttle the host)."""
import os, sys, time, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import step_variants as sv
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher

dev = torch.device("cuda")
model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
patcher = ModelPatcher(model, dev)
for batch in (8, 1):
    step = sv.make_stepper(model, patcher, dev, batch=batch)
    res = []
    with torch.inference_mode():
        for i in range(3):
            step(i)
        for i in range(4):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            step(3 + i)
            t1 = time.perf_counter()
            torch.cuda.synchronize()
            t2 = time.perf_counter()
            res.append((round((t1 - t0) * 1e3, 2), round((t2 - t0) * 1e3, 2)))
    print(json.dumps({"batch": batch, "host_enqueue_ms / step_wall_ms": res}))
