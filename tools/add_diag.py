"""Which residual adds take the native kernel? Prints operand layouts of every residual_add in one UNet step."""
import os, sys, collections
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tools import step_variants as sv
from comfy_gen_server_b200 import _native
from comfy_gen_server_b200.model_base import build_model
from comfy_gen_server_b200.model_management import ModelPatcher

dev = torch.device("cuda")
model = build_model("sdxl", dtype=torch.bfloat16, device=dev, seed=1234)
patcher = ModelPatcher(model, dev)
step = sv.make_stepper(model, patcher, dev)
seen = collections.Counter()
orig = _native.residual_add
def spy(a, b):
    native = a.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and _native._same_dense_layout(a, b)
    seen[(tuple(a.shape), a.stride(), b.stride(), native)] += 1
    return orig(a, b)
_native.residual_add = spy
with torch.inference_mode():
    step(0)
torch.cuda.synchronize()
for k, v in sorted(seen.items(), key=lambda kv: -kv[1]):
    print(v, k)
