import os, sys
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R)
import torch
from comfy_gen_server_b200 import _native as N
rows, C = 16384, 1280
x = torch.randn(rows, C, device="cuda").bfloat16(); w = torch.randn(C, device="cuda").bfloat16(); b = torch.randn(C, device="cuda").bfloat16()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_(); y = N.layer_norm(x, w, b, 1e-5)
torch.cuda.synchronize()
r = torch.randn(rows, C, device="cuda").bfloat16()
for _ in range(2):
    flush.zero_(); y = N.add_layer_norm_(x, r, w, b, 1e-5)
torch.cuda.synchronize()
print("ok")
