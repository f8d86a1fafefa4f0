import sys, torch
sys.path.insert(0, '.')
from comfy_gen_server_b200 import _native as N
B, H, Nq, d = 16, 10, 4096, 64
q, k, v = (torch.randn(B, Nq, H * d, device='cuda').bfloat16() for _ in range(3))
for _ in range(3):
    o = N.attention(q, k, v, H)
torch.cuda.synchronize()
print("ok")
