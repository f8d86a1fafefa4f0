import glob, json, os, sys
sys.path.insert(0, '.')
import torch
from comfy_gen_server_b200 import _native as N
from tools import kernel_bench as kb
cases = [("L1_cross77", 16, 10, 4096, 77, 64), ("L2_cross77", 16, 20, 1024, 77, 64), ("L2_self_1024", 16, 20, 1024, 1024, 64), ("L1_self_4096", 16, 10, 4096, 4096, 64)]
data = {c[0]: (kb.rnd((c[1], c[3], c[2]*c[5])), kb.rnd((c[1], c[4], c[2]*c[5])), kb.rnd((c[1], c[4], c[2]*c[5]))) for c in cases}
for lib in sorted(glob.glob("comfy_gen_server_b200/build/variants/libcgs_*.so")):
    N._lib = N.load_library(lib)
    row = {}
    for label, B, H, Nq, Nk, d in cases:
        q, k, v = data[label]
        row[label] = round(kb.timeit(lambda: N.attention(q, k, v, H)) * 1e3, 1)
    print(os.path.basename(lib), "us:", json.dumps(row))
