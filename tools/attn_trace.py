"""Timeline of one CTA of the attention kernel (clock64 stamps per role / key tile).
Needs the tracing build: CGS_BUILD_TRACE=1 python comfy_gen_server_b200/build.py (the stamps cost ~10 % speed)."""
import os, sys, torch
sys.path.insert(0, '.')
buf = torch.zeros(4 * 64 * 8, dtype=torch.int64, device='cuda')
os.environ["CGS_ATTN_TRACE"] = hex(buf.data_ptr())
from comfy_gen_server_b200 import _native as N
B, H, Nq, d = 16, 10, int(os.environ.get("NQ", 4096)), 64
q, k, v = (torch.randn(B, Nq, H * d, device='cuda').bfloat16() for _ in range(3))
for _ in range(3):
    buf.zero_()
    o = N.attention(q, k, v, H)
torch.cuda.synchronize()
t = buf.cpu().view(4, 64, 8)
t0 = int(t[t > 0].min())
names = ["WG0", "WG1", "MMA0", "MMA1"]
for j in list(range(10, 14)):
    for r in range(4):
        row = [(int(x) - t0) if x > 0 else -1 for x in t[r, j]]
        print(f"j={j:2d} {names[r]:5s} " + " ".join(f"{x:7d}" for x in row[:7]))
print("per-iteration period (WG0 e1 deltas):", [int(t[0, j + 1, 1] - t[0, j, 1]) for j in range(8, 20)])
print("WG0 durations waitS/ld/max/exp/st+arrive:", [[int(t[0, j, e + 1] - t[0, j, e]) for e in range(5)] for j in (10, 11)])
print("WG1 durations waitS/ld/max/exp/st+arrive:", [[int(t[1, j, e + 1] - t[1, j, e]) for e in range(5)] for j in (10, 11)])
print("MMA0 waitSfree/gap/waitP/issuePV:", [[int(t[2, j, e + 1] - t[2, j, e]) for e in range(4)] for j in (10, 11)])
print("MMA1 waitSfree/gap/waitP/issuePV:", [[int(t[3, j, e + 1] - t[3, j, e]) for e in range(4)] for j in (10, 11)])
print("WG0 barwait / phase1+2 / phase3:", [[int(t[0, j, 6] - t[0, j, 3]), int(t[0, j, 7] - t[0, j, 6]), int(t[0, j, 4] - t[0, j, 7])] for j in (10, 11, 12)])
print("WG1 barwait / phase1+2 / phase3:", [[int(t[1, j, 6] - t[1, j, 3]), int(t[1, j, 7] - t[1, j, 6]), int(t[1, j, 4] - t[1, j, 7])] for j in (10, 11, 12)])
cta = [int(x) for x in t[3, 63, :5]]
nt = Nq // 128
print("CTA entry->setup, setup->first S ready, first S -> last P handed over, ->epilogue begin, epilogue, exit sync:",
      cta[1] - cta[0], int(t[0, 0, 1]) - cta[1], int(t[0, nt - 1, 5]) - int(t[0, 0, 1]), cta[2] - int(t[0, nt - 1, 5]),
      cta[3] - cta[2], cta[4] - cta[3], "total", cta[4] - cta[0])
