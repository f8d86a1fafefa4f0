"""A/B of cuDNN / cuBLAS library settings on the benchmark step (they are library code on this path; the question is
only which setting the host side should select). usage: python tools/cudnn_ab.py"""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import step_variants as sv
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher

dev = torch.device("cuda")
model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
patcher = ModelPatcher(model, dev)
patcher.model_options["cuda_graph"] = False
step = sv.make_stepper(model, patcher, dev)
sv.timeit(step, warm=3, n=3)
res = {}
for name, setter in (("default", lambda: None),
                     ("cudnn.benchmark", lambda: setattr(torch.backends.cudnn, "benchmark", True)),
                     ("bf16_reduced_precision_reduction_off", lambda: setattr(torch.backends.cuda.matmul, "allow_bf16_reduced_precision_reduction", False))):
    setter()
    sv.timeit(step, warm=3, n=2)
    res[name] = round(min(sv.timeit(step, warm=1, n=4) for _ in range(3)), 2)
    print(name, res[name], flush=True)
    torch.backends.cudnn.benchmark = False
    torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction = True
res["default_again"] = round(min(sv.timeit(step, warm=2, n=4) for _ in range(3)), 2)
print(json.dumps(res))
