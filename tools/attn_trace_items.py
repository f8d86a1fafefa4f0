"""Timeline of consecutive work items of one persistent CTA (one-key-tile attention, e.g. 77-key cross attention).
Needs the tracing build: CGS_BUILD_TRACE=1 python comfy_gen_server_b200/build.py"""
import os, sys, torch
sys.path.insert(0, '.')
buf = torch.zeros(4 * 64 * 8, dtype=torch.int64, device='cuda')
os.environ["CGS_ATTN_TRACE"] = hex(buf.data_ptr())
from comfy_gen_server_b200 import _native as N
B, H, Nq, Nk, d = 16, 20, 1024, 77, 64
q = torch.randn(B, Nq, H * d, device='cuda').bfloat16()
k, v = (torch.randn(B, Nk, H * d, device='cuda').bfloat16() for _ in range(2))
for _ in range(3):
    buf.zero_()
    o = N.attention(q, k, v, H)
torch.cuda.synchronize()
t = buf.cpu().view(4, 64, 8)
t0 = int(t[t > 0].min())
print("softmax warp (role 0): per item  start, S ready, S loaded, max done, exp done, P handed over")
for g in range(0, 9):
    print(g, [int(x) - t0 if x > 0 else -1 for x in t[0, g, :6]])
print("MMA issuer (role 2): per item  K ready, S issued, V ready, P ready, PV issued")
for g in range(0, 9):
    print(g, [int(x) - t0 if x > 0 else -1 for x in t[2, g, :5]])
print("item period (softmax start to start):", [int(t[0, g + 1, 0] - t[0, g, 0]) for g in range(0, 7)])
