"""Channels-last GroupNorm (+SiLU) at the UNet's shapes: GB/s of the stats + apply pair (3 passes over the tensor)."""
import sys, json, torch
sys.path.insert(0, '.')
from comfy_gen_server_b200 import _native as N
from tools import kernel_bench as kb
for (n, c, hw) in ((16, 320, 128), (16, 640, 128), (16, 960, 128), (16, 640, 64), (16, 1280, 64), (16, 1280, 32), (16, 2560, 32)):
    x = kb.rnd((n, c, hw, hw)).contiguous(memory_format=torch.channels_last)
    w, b = kb.rnd((c,)), kb.rnd((c,))
    ms = kb.timeit(lambda: N.group_norm(x, w, b, 32, 1e-5, True, None))
    print(json.dumps({"shape": [n, c, hw, hw], "ms": round(ms, 4), "gbs_3_passes": round(3 * x.numel() * 2 / ms / 1e6, 1)}))
