"""Bring-up probe: run each kernel in its own subprocess (a device trap must not kill the rest)
and report max-abs error against a plain torch fp32 computation on the same GPU.

usage: python tools/kernel_probe.py [case ...]     (no args = every case)
Writes gpurun_out/probe.json.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {}


def case(fn):
    CASES[fn.__name__] = fn
    return fn


def _rand(shape, seed, scale=1.0):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to("cuda", torch.bfloat16)


@case
def layernorm():
    import torch
    from comfy_gen_server_b200 import _native as N
    out = {}
    for C in (320, 640, 1280):
        x = _rand((3, 77, C), 1)
        w = _rand((C,), 2) + 1
        b = _rand((C,), 3)
        y = N.layer_norm(x, w, b, 1e-5)
        ref = torch.nn.functional.layer_norm(x.float(), (C,), w.float(), b.float(), 1e-5)
        out[f"C{C}"] = (y.float() - ref).abs().max().item()
        r = _rand((3, 77, C), 4)
        x2 = x.clone()
        y2 = N.add_layer_norm_(x2, r, w, b, 1e-5)
        xs = (x + r)
        ref2 = torch.nn.functional.layer_norm(xs.float(), (C,), w.float(), b.float(), 1e-5)
        out[f"add_C{C}"] = max((y2.float() - ref2).abs().max().item(), (x2.float() - xs.float()).abs().max().item())
    return out


@case
def tail():
    import torch
    from comfy_gen_server_b200 import _native as N
    g = torch.Generator().manual_seed(0)
    B, per = 3, 4 * 16 * 16
    x = torch.randn(B, 4, 16, 16, generator=g).cuda()
    eu = torch.randn(B, 4, 16, 16, generator=g).cuda().bfloat16()
    ec = torch.randn(B, 4, 16, 16, generator=g).cuda().bfloat16()
    sig = torch.full((B,), 3.7).cuda()
    out = {}
    for pt in (0, 1):
        d = N.cfg_denoise(x, eu, ec, sig, 7.0, pt)
        s = sig.view(B, 1, 1, 1)
        if pt == 0:
            dc = x - ec.float() * s
            du = x - eu.float() * s
        else:
            dc = x * 1.0 / (s ** 2 + 1.0) - ec.float() * s * 1.0 / (s ** 2 + 1.0) ** 0.5
            du = x * 1.0 / (s ** 2 + 1.0) - eu.float() * s * 1.0 / (s ** 2 + 1.0) ** 0.5
        ref = du + (dc - du) * 7.0
        out[f"cfg_pred{pt}"] = (d - ref).abs().max().item()
    d = torch.randn_like(x)
    dold = torch.randn_like(x)
    xe = N.sampler_step_euler(x, d, 3.7, -0.9)
    out["euler"] = (xe - (x + (x - d) / 3.7 * -0.9)).abs().max().item()
    xd = N.sampler_step_dpmpp2m(x, d, dold, 0.8, -0.3, 1.4, 0.4)
    out["dpmpp2m"] = (xd - (0.8 * x - (-0.3) * (1.4 * d - 0.4 * dold))).abs().max().item()
    xi = N.unet_input(x, sig, 2)
    out["unet_input"] = (xi.float() - torch.cat([x, x]) / (3.7 ** 2 + 1) ** 0.5).abs().max().item()
    return out


def _linear_case(M, N_, K, bias=True, residual=False):
    import torch
    from comfy_gen_server_b200 import _native as N
    x = _rand((M, K), 1)
    w = _rand((N_, K), 2, K ** -0.5)
    b = _rand((N_,), 3) if bias else None
    r = _rand((M, N_), 4) if residual else None
    y = N.linear(x, w, b, r)
    torch.cuda.synchronize()
    ref = x.float() @ w.float().t()
    if bias:
        ref = ref + b.float()
    if residual:
        ref = ref + r.float()
    return (y.float() - ref).abs().max().item()


@case
def linear_small():
    return {"M128_N256_K64": _linear_case(128, 256, 64, bias=False)}


@case
def linear_mid():
    return {"M256_N320_K320": _linear_case(256, 320, 320),
            "M1000_N640_K1280_res": _linear_case(1000, 640, 1280, residual=True)}


@case
def linear_big():
    return {"M16384_N1280_K1280": _linear_case(16384, 1280, 1280)}


def _geglu_case(M, N_, K):
    import torch
    from comfy_gen_server_b200 import _native as N
    x = _rand((M, K), 1)
    w = _rand((2 * N_, K), 2, K ** -0.5)
    b = _rand((2 * N_,), 3, 0.1)
    y = N.geglu(x, w, b)
    torch.cuda.synchronize()
    h = x.float() @ w.float().t() + b.float()
    ref = h[:, :N_] * torch.nn.functional.gelu(h[:, N_:])
    return (y.float() - ref).abs().max().item()


@case
def geglu_small():
    return {"M128_N128_K64": _geglu_case(128, 128, 64)}


@case
def geglu_mid():
    return {"M300_N1280_K320": _geglu_case(300, 1280, 320), "M4096_N2560_K640": _geglu_case(4096, 2560, 640)}


def _attn_case(B, H, Nq, Nk, d, scale=1.0):
    import torch
    from comfy_gen_server_b200 import _native as N
    q = _rand((B, Nq, H * d), 1, scale)
    k = _rand((B, Nk, H * d), 2, scale)
    v = _rand((B, Nk, H * d), 3)
    o = N.attention(q, k, v, H)
    torch.cuda.synchronize()
    qf, kf, vf = (t.float().view(B, -1, H, d).transpose(1, 2) for t in (q, k, v))
    ref = torch.softmax(qf @ kf.transpose(-1, -2) * d ** -0.5, -1) @ vf
    ref = ref.transpose(1, 2).reshape(B, Nq, H * d)
    err = (o.float() - ref).abs()
    return {"max": err.max().item(), "mean": err.mean().item(), "nan": int(torch.isnan(o.float()).sum().item())}


@case
def attn_1tile():
    return {"B1H1_128x128_d64": _attn_case(1, 1, 128, 128, 64)}


@case
def attn_2qtiles():
    return {"B1H1_256x128_d64": _attn_case(1, 1, 256, 128, 64)}


@case
def attn_multi_kv():
    return {"B1H2_256x512_d64": _attn_case(1, 2, 256, 512, 64), "B2H3_1024x1024_d64": _attn_case(2, 3, 1024, 1024, 64)}


@case
def attn_rescale():
    # large logits force the lazy-rescale path
    return {"B1H1_256x1024_d64_x4": _attn_case(1, 1, 256, 1024, 64, scale=4.0)}


@case
def attn_cross():
    return {"B2H10_1024x77_d64": _attn_case(2, 10, 1024, 77, 64), "B2H10_4096x154_d64": _attn_case(2, 10, 4096, 154, 64)}


@case
def attn_ragged():
    return {"B2H5_576x576_d64": _attn_case(2, 5, 576, 576, 64), "B1H20_144x144_d64": _attn_case(1, 20, 144, 144, 64),
            "B1H2_64x64_d64": _attn_case(1, 2, 64, 64, 64)}


@case
def attn_d40():
    return {"B2H8_4096x4096_d40": _attn_case(2, 8, 4096, 4096, 40), "B2H8_1024x77_d40": _attn_case(2, 8, 1024, 77, 40)}


@case
def attn_d80():
    return {"B2H8_1024x1024_d80": _attn_case(2, 8, 1024, 1024, 80)}


@case
def attn_d160():
    return {"B2H8_256x256_d160": _attn_case(2, 8, 256, 256, 160), "B2H8_64x77_d160": _attn_case(2, 8, 64, 77, 160)}


def _run_one(name):
    import torch
    torch.cuda.init()
    t0 = time.time()
    res = CASES[name]()
    torch.cuda.synchronize()
    print("RESULT " + json.dumps({"case": name, "ok": True, "res": res, "s": round(time.time() - t0, 2)}))


def main():
    args = sys.argv[1:]
    if args and args[0] == "--one":
        _run_one(args[1])
        return
    names = args or list(CASES)
    results = []
    for n in names:
        try:
            p = subprocess.run([sys.executable, __file__, "--one", n], capture_output=True, text=True, timeout=180)
            line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            if p.returncode == 0 and line:
                results.append(json.loads(line[-1][7:]))
            else:
                results.append({"case": n, "ok": False, "rc": p.returncode,
                                "tail": (p.stdout + p.stderr)[-1500:]})
        except subprocess.TimeoutExpired:
            results.append({"case": n, "ok": False, "rc": "timeout"})
        print(json.dumps(results[-1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w") as f:
        json.dump(results, f, indent=1)


if __name__ == "__main__":
    main()
