import json, os, sys
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R)
import torch
from comfy_gen_server_b200 import _native as N
from tools import kernel_bench as kb
shapes = [("linear", "L2_to_q", 16384, 1280, 1280), ("linear", "L2_ff_out", 16384, 5120, 1280), ("linear", "b1_ff_out", 2048, 5120, 1280), ("geglu", "L2", 16384, 1280, 5120)]
libs = [None] + sorted(os.path.join(R, "build_variants", f) for f in os.listdir(os.path.join(R, "build_variants")) if f.endswith(".so"))
for lib in libs:
    N._lib = None
    N._gemm_ws.clear()
    if lib: N._lib = N.load_library(lib)
    l = N.load_library()
    row = {}
    for kind, label, M, K, Nn in shapes:
        x = kb.rnd((M, K))
        if kind == "geglu":
            w, b = kb.rnd((2 * Nn, K), scale=K ** -0.5), kb.rnd((2 * Nn,), scale=0.1); fn = lambda: N.geglu(x, w, b); fl = 4.0 * M * K * Nn
        else:
            w, b = kb.rnd((Nn, K), scale=K ** -0.5), kb.rnd((Nn,), scale=0.1); fn = lambda: N.linear(x, w, b); fl = 2.0 * M * K * Nn
        for split in (False, True):
            N.SPLIT_TAIL = split
            for v in (1, 2):
                l.cgs_gemm_set_variant(v)
                row[f"{label}_v{v}_{'ws' if split else 'no'}"] = round(kb.timeit(fn) * 1e3, 1)
        l.cgs_gemm_set_variant(0)
    print(os.path.basename(lib) if lib else "product", json.dumps(row), flush=True)
