import json, os, subprocess, sys
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
code = r'''
import sys, json; sys.path.insert(0, %r)
import torch
from comfy_gen_server_b200 import _native as N
from tools import kernel_bench as kb
cases = [("L2_self_1024", 16, 20, 1024, 1024, 64), ("L1_self_4096", 16, 10, 4096, 4096, 64), ("L1_cross77", 16, 10, 4096, 77, 64), ("L2_cross77", 16, 20, 1024, 77, 64), ("sd15_d80_1024", 32, 8, 1024, 1024, 80), ("sd15_d40_4096", 32, 8, 4096, 4096, 40)]
row = {}
for label, B, H, Nq, Nk, d in cases:
    q, k, v = kb.rnd((B, Nq, H * d)), kb.rnd((B, Nk, H * d)), kb.rnd((B, Nk, H * d))
    ms = kb.timeit(lambda: N.attention(q, k, v, H))
    row[label] = [round(ms * 1e3, 1), round(4.0 * B * H * Nq * Nk * d / ms / 1e9)]
print(json.dumps(row))
''' % R
for env in ({}, {"CGS_ATTN_NO_BN96": "1"}, {"CGS_ATTN_MAX_TILES_1CTA": "4"}, {"CGS_ATTN_MAX_TILES_1CTA": "4", "CGS_ATTN_NO_3TILE": "1", "CGS_ATTN_2X64": "1"},
            {"CGS_ATTN_MAX_TILES_1CTA": "4", "CGS_ATTN_NO_3TILE": "1"}):
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=dict(os.environ, **env))
    print(env, p.stdout.strip().splitlines()[-1] if p.stdout.strip() else p.stderr[-500:], flush=True)
