"""Per-kernel micro-benchmarks at the BASELINE.json shapes: CUDA-event timing on the launching
stream, >= 3 warm-ups, L2 flushed between timed iterations (writes a 256 MiB buffer).
Prints one JSON line per case and writes gpurun_out/kernel_bench.json.

usage: python tools/kernel_bench.py [attn|geglu|linear|ln|tail ...]
"""
from __future__ import annotations

import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from comfy_gen_server_b200 import _native as N  # noqa: E402

if os.environ.get("CGS_LIB"):          # a private variant build (tools/gemm_dbg.sh, tools/attn_variants.py)
    N._lib = N.load_library(os.path.join(ROOT, os.environ["CGS_LIB"]))

PEAKS = {"hbm_gbs": 6554.6, "bf16_tflops": 1650.7}
try:
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        PEAKS.update({k: v for k, v in json.load(f).items() if k in PEAKS})
except OSError:
    pass

_flush = None


def timeit(fn, iters=10, warmup=3, flush=True):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        if flush:
            _flush.zero_()          # flush L2: write a buffer larger than L2 (profiling recipe)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        times.append(s.elapsed_time(e))
    times.sort()
    return times[len(times) // 2]


def rnd(shape, dtype=torch.bfloat16, scale=1.0):
    return (torch.randn(shape, device="cuda") * scale).to(dtype)


def bench_attn():
    out = []
    # (label, B2, H, Nq, Nk, d)
    cases = [("sdxl1024_L1_self", 16, 10, 4096, 4096, 64), ("sdxl1024_L2_self", 16, 20, 1024, 1024, 64),
             ("sdxl2048_L1_self", 2, 10, 16384, 16384, 64), ("sdxl1024_L1_cross77", 16, 10, 4096, 77, 64),
             ("sdxl1024_L2_cross77", 16, 20, 1024, 77, 64), ("sd15_L0_self_d40", 32, 8, 4096, 4096, 40),
             ("sd15_L1_self_d80", 32, 8, 1024, 1024, 80), ("sd21_L0_self", 16, 5, 9216, 9216, 64),
             # one / two images per GPU (strong scaling on 8 / 4 GPUs)
             ("b1_sdxl_L2_self", 2, 20, 1024, 1024, 64), ("b1_sdxl_L1_self", 2, 10, 4096, 4096, 64),
             ("b2_sdxl_L2_self", 4, 20, 1024, 1024, 64), ("b2_sdxl_L1_self", 4, 10, 4096, 4096, 64)]
    only = os.environ.get("ATTN_CASES")
    if only:
        cases = [c for c in cases if any(c[0].startswith(o) for o in only.split(","))]
    for label, B, H, Nq, Nk, d in cases:
        q, k, v = rnd((B, Nq, H * d)), rnd((B, Nk, H * d)), rnd((B, Nk, H * d))
        ms = timeit(lambda: N.attention(q, k, v, H))
        flops = 4.0 * B * H * Nq * Nk * d
        sd = q.view(B, Nq, H, d).transpose(1, 2), k.view(B, Nk, H, d).transpose(1, 2), v.view(B, Nk, H, d).transpose(1, 2)
        ms_sdpa = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(*sd).transpose(1, 2).reshape(B, Nq, H * d))
        out.append({"kernel": "attention", "case": label, "ms": ms, "tflops": flops / ms / 1e9,
                    "frac_of_measured_peak": flops / ms / 1e9 / PEAKS["bf16_tflops"],
                    "torch_sdpa_ms": ms_sdpa, "torch_sdpa_tflops": flops / ms_sdpa / 1e9})
    return out


def bench_gemm_variants():
    """one CTA per tile vs CTA pairs (cgs_gemm_set_variant) vs cuBLASLt, on the SDXL GEMM shapes"""
    out = []
    lib = N.load_library()
    shapes = [("geglu", "sdxl_L2", 16 * 1024, 1280, 5120), ("geglu", "sdxl_L1", 16 * 4096, 640, 2560),
              ("linear", "sdxl_L2_ff_out", 16 * 1024, 5120, 1280), ("linear", "sdxl_L1_ff_out", 16 * 4096, 2560, 640),
              ("linear", "sdxl_L2_to_q", 16 * 1024, 1280, 1280), ("linear", "sdxl_L1_to_q", 16 * 4096, 640, 640),
              ("linear", "sdxl_L2_qkv", 16 * 1024, 1280, 3840), ("linear", "sdxl_L1_qkv", 16 * 4096, 640, 1920),
              ("linear", "sdxl_L2_ff_out_b1", 2 * 1024, 5120, 1280), ("linear", "sdxl_L2_to_q_b1", 2 * 1024, 1280, 1280),
              ("linear", "sdxl_L2_qkv_b1", 2 * 1024, 1280, 3840), ("geglu", "sdxl_L2_b1", 2 * 1024, 1280, 5120)]
    for kind, label, M, K, Nn in shapes:
        row = {"kernel": "gemm_variants", "kind": kind, "case": label, "M": M, "K": K, "N": Nn}
        if kind == "geglu":
            x, w, b = rnd((M, K)), rnd((2 * Nn, K), scale=K ** -0.5), rnd((2 * Nn,), scale=0.1)
            fn = lambda: N.geglu(x, w, b)
            flops = 2.0 * M * K * 2 * Nn
        else:
            x, w, b = rnd((M, K)), rnd((Nn, K), scale=K ** -0.5), rnd((Nn,), scale=0.1)
            fn = lambda: N.linear(x, w, b)
            flops = 2.0 * M * K * Nn
            row["cublaslt_tflops"] = round(flops / timeit(lambda: torch.nn.functional.linear(x, w, b)) / 1e9, 1)
        for v, name in ((1, "single_cta"), (2, "cta_pair"), (0, "auto")):
            lib.cgs_gemm_set_variant(v)
            row[name + "_tflops"] = round(flops / timeit(fn) / 1e9, 1)
        lib.cgs_gemm_set_variant(0)
        out.append(row)
    return out


def bench_conv():
    """3x3 convolutions of the SDXL UNet (channels-last, UNet batch 16 and 2) against cuDNN (F.conv2d)"""
    out = []
    lib = N.load_library()
    cl = torch.channels_last
    # (label, n, hw, Cin, Cout)
    cases = [("L0_320_320", 16, 128, 320, 320), ("L0_960_320", 16, 128, 960, 320), ("L0_640_320", 16, 128, 640, 320),
             ("L1_640_640", 16, 64, 640, 640), ("L1_320_640", 16, 64, 320, 640), ("L1_1920_640", 16, 64, 1920, 640),
             ("L1_1280_640", 16, 64, 1280, 640), ("L1_960_640", 16, 64, 960, 640),
             ("L2_1280_1280", 16, 32, 1280, 1280), ("L2_640_1280", 16, 32, 640, 1280), ("L2_2560_1280", 16, 32, 2560, 1280),
             ("L2_1920_1280", 16, 32, 1920, 1280), ("UP_1280_1280", 16, 64, 1280, 1280), ("UP_640_640", 16, 128, 640, 640),
             ("b1_L0_320_320", 2, 128, 320, 320), ("b1_L1_640_640", 2, 64, 640, 640), ("b1_L2_1280_1280", 2, 32, 1280, 1280),
             ("b1_L2_2560_1280", 2, 32, 2560, 1280)]
    only = os.environ.get("CONV_CASES")
    for label, n, hw, cin, cout in cases:
        if only and not any(label.startswith(o) for o in only.split(",")):
            continue
        x = rnd((n, cin, hw, hw)).contiguous(memory_format=cl)
        w = rnd((cout, cin, 3, 3), scale=(9 * cin) ** -0.5).contiguous(memory_format=cl)
        b = rnd((cout,), scale=0.1)
        r = rnd((n, cout, hw, hw)).contiguous(memory_format=cl)
        flops = 2.0 * n * hw * hw * 9 * cin * cout
        row = {"kernel": "conv3x3", "case": label, "n": n, "hw": hw, "cin": cin, "cout": cout}
        row["cudnn_nobias_tflops"] = round(flops / timeit(lambda: torch.nn.functional.conv2d(x, w, None, padding=1)) / 1e9, 1)
        row["cudnn_bias_tflops"] = round(flops / timeit(lambda: torch.nn.functional.conv2d(x, w, b, padding=1)) / 1e9, 1)
        variants = ((11, "tiles256"), (12, "tiles320")) if cout % 320 == 0 else ((10, "tiles256"),)
        nores = bool(os.environ.get("CONV_NORES"))          # (variant builds without the residual stage)
        for v, name in variants:
            lib.cgs_gemm_set_variant(v)
            for sv, sname in ((20, ""), (21, "_pieces")):
                lib.cgs_gemm_set_variant(sv)
                row[name + sname + "_tflops"] = round(flops / timeit(lambda: N.conv3x3(x, w, b)) / 1e9, 1)
                if not nores:
                    row[name + sname + "_residual_tflops"] = round(flops / timeit(lambda: N.conv3x3(x, w, b, r)) / 1e9, 1)
        lib.cgs_gemm_set_variant(10)
        lib.cgs_gemm_set_variant(20)
        if label.startswith("UP_"):      # Upsample: interpolate + conv; here on the half-size image, as four 2x2 convolutions
            xs = rnd((n, cin, hw // 2, hw // 2)).contiguous(memory_format=cl)
            w4 = N.upsample2x_conv_weights(w)
            ms = timeit(lambda: N.upsample2x_conv3x3(xs, w4, b))
            ms_ref = timeit(lambda: torch.nn.functional.conv2d(N.upsample_nearest2x(xs), w, None, padding=1))
            row["upconv_ms"], row["upsample_kernel_plus_cudnn_ms"] = round(ms, 4), round(ms_ref, 4)
            row["upconv_equiv_tflops"] = round(flops / ms / 1e9, 1)
        y = N.conv3x3(x, w, b)
        ref = torch.nn.functional.conv2d(x, w, b, padding=1)
        row["max_abs_vs_cudnn"] = float((y.float() - ref.float()).abs().max())
        out.append(row)
    return out


def bench_split():
    """How the last, partly filled round of tiles is scheduled (cgs_gemm_set_variant 30-32 linear, 20-22 convolution):
    64-column groups dealt out evenly / whole tiles / halves or quarters with at most one piece per worker."""
    out = []
    lib = N.load_library()
    cl = torch.channels_last
    for label, M, K, Nn in [("L2_to_out", 16384, 1280, 1280), ("L2_ff_out", 16384, 5120, 1280), ("L1_to_out", 65536, 640, 640),
                            ("L1_ff_out", 65536, 2560, 640), ("L2_qkv", 16384, 1280, 3840), ("b4_L2_ff_out", 8192, 5120, 1280),
                            ("b1_L2_ff_out", 2048, 5120, 1280)]:
        x, w, b, r = rnd((M, K)), rnd((Nn, K), scale=K ** -0.5), rnd((Nn,), scale=0.1), rnd((M, Nn))
        flops = 2.0 * M * K * Nn
        row = {"kernel": "split_policy", "case": label, "M": M, "K": K, "N": Nn}
        for v, name in ((30, "groups64"), (31, "whole"), (32, "halves_quarters")):
            lib.cgs_gemm_set_variant(v)
            row[name + "_tflops"] = round(flops / timeit(lambda: N.linear(x, w, b, r), iters=30) / 1e9, 1)
        lib.cgs_gemm_set_variant(30)
        out.append(row)
    for label, n, hw, cin, cout in [("conv_L2_1280_1280", 16, 32, 1280, 1280), ("conv_L2_2560_1280", 16, 32, 2560, 1280),
                                    ("conv_b1_L2_1280_1280", 2, 32, 1280, 1280), ("conv_b1_L2_2560_1280", 2, 32, 2560, 1280)]:
        x = rnd((n, cin, hw, hw)).contiguous(memory_format=cl)
        w = rnd((cout, cin, 3, 3), scale=(9 * cin) ** -0.5).contiguous(memory_format=cl)
        b = rnd((cout,), scale=0.1)
        flops = 2.0 * n * hw * hw * 9 * cin * cout
        row = {"kernel": "split_policy", "case": label}
        for v, name in ((21, "groups64"), (20, "whole"), (22, "halves_quarters")):
            lib.cgs_gemm_set_variant(v)
            row[name + "_tflops"] = round(flops / timeit(lambda: N.conv3x3(x, w, b), iters=30) / 1e9, 1)
        lib.cgs_gemm_set_variant(20)
        out.append(row)
    return out


def bench_geglu():
    out = []
    for label, M, K, Nn in [("sdxl1024_L2", 16 * 1024, 1280, 5120), ("sdxl1024_L1", 16 * 4096, 640, 2560),
                            ("sd15_L0", 32 * 4096, 320, 1280)]:
        x, w, b = rnd((M, K)), rnd((2 * Nn, K), scale=K ** -0.5), rnd((2 * Nn,), scale=0.1)
        ms = timeit(lambda: N.geglu(x, w, b))
        flops = 2.0 * M * K * 2 * Nn

        def ref():
            h = torch.nn.functional.linear(x, w, b)
            a, g = h.chunk(2, dim=-1)
            return a * torch.nn.functional.gelu(g)
        ms_ref = timeit(ref)
        out.append({"kernel": "geglu_gemm", "case": label, "ms": ms, "tflops": flops / ms / 1e9,
                    "frac_of_measured_peak": flops / ms / 1e9 / PEAKS["bf16_tflops"],
                    "torch_ms": ms_ref, "torch_tflops": flops / ms_ref / 1e9})
    return out


def bench_linear():
    out = []
    for label, M, K, Nn in [("sdxl_L2_to_q", 16 * 1024, 1280, 1280), ("sdxl_L2_ff_out", 16 * 1024, 5120, 1280),
                            ("sdxl_L1_to_q", 16 * 4096, 640, 640), ("sdxl_L1_ff_out", 16 * 4096, 2560, 640)]:
        x, w, b = rnd((M, K)), rnd((Nn, K), scale=K ** -0.5), rnd((Nn,), scale=0.1)
        ms = timeit(lambda: N.linear(x, w, b))
        ms_ref = timeit(lambda: torch.nn.functional.linear(x, w, b))
        flops = 2.0 * M * K * Nn
        out.append({"kernel": "linear", "case": label, "ms": ms, "tflops": flops / ms / 1e9,
                    "frac_of_measured_peak": flops / ms / 1e9 / PEAKS["bf16_tflops"],
                    "torch_ms": ms_ref, "torch_tflops": flops / ms_ref / 1e9})
    return out


def bench_ln():
    out = []
    for label, rows, C in [("sdxl_L2", 16 * 1024, 1280), ("sdxl_L1", 16 * 4096, 640), ("sd15_L0", 32 * 4096, 320),
                           ("synthetic_256MiB", 104858, 1280)]:
        x, w, b = rnd((rows, C)), rnd((C,)), rnd((C,))
        ms = timeit(lambda: N.layer_norm(x, w, b, 1e-5))
        ms_ref = timeit(lambda: torch.nn.functional.layer_norm(x, (C,), w, b, 1e-5))
        byt = 2.0 * rows * C * 2
        out.append({"kernel": "layernorm", "case": label, "ms": ms, "gbs": byt / ms / 1e6,
                    "frac_of_measured_hbm": byt / ms / 1e6 / PEAKS["hbm_gbs"], "torch_ms": ms_ref,
                    "torch_gbs": byt / ms_ref / 1e6})
        r = rnd((rows, C))
        ms2 = timeit(lambda: N.add_layer_norm_(x, r, w, b, 1e-5))
        byt2 = 5.0 * rows * C * 2
        out.append({"kernel": "add_layernorm", "case": label, "ms": ms2, "gbs": byt2 / ms2 / 1e6,
                    "frac_of_measured_hbm": byt2 / ms2 / 1e6 / PEAKS["hbm_gbs"]})
    return out


def bench_tail():
    out = []
    for label, B, hw in [("sdxl1024_B8", 8, 128), ("sdxl2048_B4", 4, 256), ("synthetic_B4096", 4096, 128)]:
        x = torch.randn(B, 4, hw, hw, device="cuda")
        eu, ec = rnd((B, 4, hw, hw)), rnd((B, 4, hw, hw))
        dold = torch.randn_like(x)
        sig = torch.full((B,), 3.0, device="cuda")
        xo = torch.empty_like(x)
        ms = timeit(lambda: N.cfg_sampler_tail(x, eu, ec, sig, 7.0, 0, dold, 3, 0.8, -0.3, 1.4, 0.4, x_out=xo))
        byt = x.numel() * (4 + 2 + 2 + 4 + 4 + 4)
        out.append({"kernel": "cfg_sampler_tail_dpmpp2m", "case": label, "ms": ms, "gbs": byt / ms / 1e6,
                    "frac_of_measured_hbm": byt / ms / 1e6 / PEAKS["hbm_gbs"]})
        ms = timeit(lambda: N.cfg_denoise(x, eu, ec, sig, 7.0, 0))
        byt = x.numel() * (4 + 2 + 2 + 4)
        out.append({"kernel": "cfg_denoise", "case": label, "ms": ms, "gbs": byt / ms / 1e6,
                    "frac_of_measured_hbm": byt / ms / 1e6 / PEAKS["hbm_gbs"]})
    return out


def main():
    which = sys.argv[1:] or ["attn", "geglu", "linear", "ln", "tail"]
    fns = {"split": bench_split, "conv": bench_conv, "gemmv": bench_gemm_variants, "attn": bench_attn, "geglu": bench_geglu, "linear": bench_linear, "ln": bench_ln, "tail": bench_tail}
    results = []
    for w in which:
        for r in fns[w]():
            print(json.dumps(r), flush=True)
            results.append(r)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "kernel_bench.json"), "w") as f:
        json.dump({"peaks": PEAKS, "results": results}, f, indent=1)


if __name__ == "__main__":
    main()
