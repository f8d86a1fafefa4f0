"""Kernel parity on the GPU: every C-ABI entry point (through the ctypes binding) against the oracle
on the same seeded inputs and against the reference's golden outputs. Tolerances:
  attention                    2e-2 max-abs in bf16 (BASELINE.json north_star)
  LayerNorm / GEGLU / linear   1 bf16 ulp of the output magnitude (inputs are bf16-rounded first)
  CFG / sampler tail           bit-exact (fp32, reference operation order)
"""
import math

import pytest
import torch

from conftest import load_golden, randn
from oracle import oracle as O

pytestmark = pytest.mark.gpu
ATTN_TOL = 2e-2


def N():
    from comfy_gen_server_b200 import _native
    return _native


def bf(t):
    return t.to("cuda", torch.bfloat16)


def ulp_tol(ref):
    return float(ref.abs().max()) * 2 ** -7 + 1e-3


# ------------------------------------------------------------------------------------------ attention
@pytest.mark.parametrize("name", ["self_d64", "cross77_d40", "cross154_d80", "self_d160_ragged"])
def test_attention_golden(name):
    g = load_golden("attention")
    q, k, v, h = g[f"{name}.q"], g[f"{name}.k"], g[f"{name}.v"], int(g[f"{name}.heads"])
    out = N().attention(bf(q), bf(k), bf(v), h).float().cpu()
    assert (out - g[f"{name}.out"]).abs().max() < ATTN_TOL          # vs the reference's fp32 output
    ref = O.attention(bf(q).float().cpu(), bf(k).float().cpu(), bf(v).float().cpu(), h)
    assert (out - ref).abs().max() < ATTN_TOL / 2                   # vs the oracle on the same bf16 inputs


@pytest.mark.parametrize("name", ["mask_causal_4d", "mask_float_2d", "mask_float_per_head", "mask_bool_keys"])
def test_masked_attention_golden(name):
    """optimized_attention_masked: additive float and boolean keep-masks in the reference's shapes."""
    g = load_golden("attention")
    q, k, v, h, mask = g[f"{name}.q"], g[f"{name}.k"], g[f"{name}.v"], int(g[f"{name}.heads"]), g[f"{name}.mask"]
    out = N().attention(bf(q), bf(k), bf(v), h, mask.cuda()).float().cpu()
    assert (out - g[f"{name}.out"]).abs().max() < ATTN_TOL
    ref = O.attention(bf(q).float().cpu(), bf(k).float().cpu(), bf(v).float().cpu(), h, mask)
    assert (out - ref).abs().max() < ATTN_TOL / 2


def test_masked_attention_large_and_fully_masked_rows():
    """Several key tiles with a mask (the two-stage variant), a broadcast key dimension, and rows whose keys are all
    masked by a boolean mask: a uniform average, as attention.py:113-117 gives (-finfo.max fill)."""
    n = N()
    B, H, Nq, Nk, d = 2, 4, 300, 700, 64
    q, k, v = bf(randn((B, Nq, H * d), 1)), bf(randn((B, Nk, H * d), 2)), bf(randn((B, Nk, H * d), 3))
    mask = randn((B, 1, Nq, Nk), 4).cuda() * 3.0
    ref = O.attention(q.float().cpu(), k.float().cpu(), v.float().cpu(), H, mask.cpu())
    assert (n.attention(q, k, v, H, mask).float().cpu() - ref).abs().max() < ATTN_TOL
    per_query = randn((Nq, 1), 5).cuda()                       # key dimension broadcast: no effect on the softmax
    base = n.attention(q, k, v, H)
    assert (n.attention(q, k, v, H, per_query).float() - base.float()).abs().max() < 2e-2
    keep = torch.ones((B, 1, Nq, Nk), dtype=torch.bool, device="cuda")
    keep[0, 0, 5] = False                                         # one fully masked row
    keep[:, :, :, 350:] = False
    keep[0, 0, 5] = False
    out = n.attention(q, k, v, H, keep).float().cpu()
    ref = O.attention(q.float().cpu(), k.float().cpu(), v.float().cpu(), H, keep.cpu())
    assert torch.isfinite(out).all() and (out - ref).abs().max() < ATTN_TOL


def test_masked_attention_minus_inf_float_mask_first_tile_fully_masked():
    """An additive float mask with -inf entries (what SDPA callers pass): rows whose first 128-key tile is entirely
    -inf while later tiles hold valid keys must come out finite and equal to the reference on the valid keys."""
    n = N()
    B, H, Nq, Nk, d = 1, 2, 200, 400, 64
    q, k, v = bf(randn((B, Nq, H * d), 1)), bf(randn((B, Nk, H * d), 2)), bf(randn((B, Nk, H * d), 3))
    mask = torch.zeros((B, 1, Nq, Nk))
    mask[:, :, 10:60, :128] = float("-inf")          # first key tile fully masked for 50 rows
    mask[:, :, 100:120, :300] = float("-inf")        # first two tiles and part of the third
    out = n.attention(q, k, v, H, mask.cuda()).float().cpu()
    ref = O.attention(q.float().cpu(), k.float().cpu(), v.float().cpu(), H, mask)
    assert torch.isfinite(out).all()
    assert (out - ref).abs().max() < ATTN_TOL


ATTN_SHAPES = [
    # B, H, Nq, Nk, d          -- the geometries of SURVEY.md section 8 (scaled-down batch)
    (2, 10, 4096, 4096, 64), (1, 20, 1024, 1024, 64), (2, 10, 4096, 77, 64), (2, 20, 1024, 154, 64),
    (2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80), (2, 8, 256, 256, 160), (2, 8, 64, 64, 160),
    (2, 8, 64, 77, 160), (1, 5, 9216, 9216, 64), (2, 10, 2304, 2304, 64), (2, 20, 576, 576, 64),
    (2, 20, 144, 144, 64), (2, 20, 144, 77, 64), (1, 1, 1, 1, 64), (3, 2, 129, 255, 64), (1, 2, 300, 231, 8),
    (1, 3, 257, 16, 16),
    # instantiations the first round never ran: <2,2,2,128> (64 < d <= 128 with more than 8 key tiles: SD1.5 above
    # 512^2), <3,1,1> (d = 160) with many key tiles and ragged ends
    (1, 8, 2304, 2304, 80), (1, 4, 1500, 3000, 128), (1, 8, 1300, 1300, 160), (1, 2, 700, 2100, 96),
    # the 96-key-tile variants of text cross-attention (Nk <= 192, d <= 128) and their borders
    (2, 8, 1024, 77, 80), (1, 4, 300, 96, 64), (1, 4, 300, 97, 64), (1, 4, 200, 192, 40), (1, 4, 200, 193, 64),
    (2, 5, 9216, 77, 64), (1, 3, 130, 1, 128),
]


@pytest.mark.parametrize("B,H,Nq,Nk,d", ATTN_SHAPES)
def test_attention_vs_fp32(B, H, Nq, Nk, d):
    q, k, v = bf(randn((B, Nq, H * d), 1)), bf(randn((B, Nk, H * d), 2)), bf(randn((B, Nk, H * d), 3))
    out = N().attention(q, k, v, H).float()
    qf, kf, vf = (t.float().view(B, -1, H, d).transpose(1, 2) for t in (q, k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * d ** -0.5, -1) @ vf).transpose(1, 2).reshape(B, Nq, H * d)
    assert not torch.isnan(out).any()
    assert (out - ref).abs().max() < ATTN_TOL


def test_attention_config5_16384_tokens_vs_fp32_rows():
    """BASELINE.json configs[4]: the 16384-token self-attention of SDXL 2048^2 (10 heads, d 64), compared with the
    fp32 computation on a subset of query rows (all keys): first / last tiles, tile borders, and a random sample."""
    B, H, Nn, d = 1, 10, 16384, 64
    q, k, v = bf(randn((B, Nn, H * d), 1)), bf(randn((B, Nn, H * d), 2)), bf(randn((B, Nn, H * d), 3))
    out = N().attention(q, k, v, H).float()
    rows = torch.cat([torch.arange(0, 130), torch.arange(383, 387), torch.arange(8190, 8194),
                      torch.arange(Nn - 130, Nn),
                      torch.randint(0, Nn, (256,), generator=torch.Generator().manual_seed(5))]).unique().cuda()
    qf = q.float().view(B, Nn, H, d)[:, rows].transpose(1, 2)                   # [B, H, R, d]
    kf, vf = (t.float().view(B, Nn, H, d).transpose(1, 2) for t in (k, v))
    ref = (torch.softmax(qf @ kf.transpose(-1, -2) * d ** -0.5, -1) @ vf).transpose(1, 2).reshape(B, len(rows), H * d)
    assert torch.isfinite(out).all()
    assert (out[:, rows] - ref).abs().max() < ATTN_TOL
    # size-independent property on the whole output: constant values -> constant output (softmax rows sum to one)
    half = torch.full_like(v, 0.5)
    assert (N().attention(q, k, half, H).float() - 0.5).abs().max() < 4e-3


def test_attention_large_logits_rescale_path():
    B, H, Nq, Nk, d = 1, 2, 384, 2048, 64
    q, k, v = bf(randn((B, Nq, H * d), 1, 4.0)), bf(randn((B, Nk, H * d), 2, 4.0)), bf(randn((B, Nk, H * d), 3))
    out = N().attention(q, k, v, H).float()
    ref = O.attention(q.float().cpu(), k.float().cpu(), v.float().cpu(), H).cuda()
    assert (out - ref).abs().max() < ATTN_TOL


def test_attention_properties_full_size():
    """Size-independent properties at a BASELINE.json full size (SDXL 1024^2 level 1: 4096 tokens):
    rows of softmax sum to one (v = const -> out = const); permuting the keys does not change the result."""
    B, H, Nn, d = 2, 10, 4096, 64
    q, k = bf(randn((B, Nn, H * d), 1)), bf(randn((B, Nn, H * d), 2))
    ones = torch.full((B, Nn, H * d), 0.5, device="cuda", dtype=torch.bfloat16)
    out = N().attention(q, k, ones, H).float()
    assert (out - 0.5).abs().max() < 4e-3
    v = bf(randn((B, Nn, H * d), 3))
    perm = torch.randperm(Nn, generator=torch.Generator().manual_seed(0)).cuda()
    a = N().attention(q, k, v, H).float()
    b = N().attention(q, k[:, perm].contiguous(), v[:, perm].contiguous(), H).float()
    assert (a - b).abs().max() < 8e-3


def test_attention_strided_views_of_fused_qkv():
    B, Nn, H, d = 2, 512, 4, 64
    qkv = bf(randn((B, Nn, 3 * H * d), 5))
    q, k, v = qkv[..., :H * d], qkv[..., H * d:2 * H * d], qkv[..., 2 * H * d:]
    a = N().attention(q, k, v, H)
    b = N().attention(q.contiguous(), k.contiguous(), v.contiguous(), H)
    assert torch.equal(a, b)
    assert torch.equal(qkv, bf(randn((B, Nn, 3 * H * d), 5)))      # inputs untouched


def test_attention_errors_and_empty():
    n = N()
    q = bf(randn((1, 16, 64), 1))
    with pytest.raises(RuntimeError):
        n.attention(q.cpu(), q.cpu(), q.cpu(), 1)                  # no CPU fallback
    with pytest.raises(RuntimeError):
        n.attention(q.float(), q.float(), q.float(), 1)            # bf16 only
    with pytest.raises(RuntimeError):
        n.attention(bf(randn((1, 16, 12), 1)), bf(randn((1, 16, 12), 1)), bf(randn((1, 16, 12), 1)), 1)  # d % 8
    with pytest.raises(RuntimeError):
        n.attention(q, q[:, :0], q[:, :0], 1)                      # empty key set
    assert n.attention(q[:, :0], q, q, 1).shape == (1, 0, 64)


# ------------------------------------------------------------------------------------------ LayerNorm
@pytest.mark.parametrize("rows,C", [(64, 128), (2 * 4096, 320), (2 * 1024, 640), (2 * 1024, 1280), (7, 2560), (1, 8)])
def test_layernorm_vs_oracle(rows, C):
    x, w, b = bf(randn((rows, C), 1, 2.0)), bf(randn((C,), 2) * 0.2 + 1), bf(randn((C,), 3) * 0.2)
    y = N().layer_norm(x, w, b, 1e-5).float().cpu()
    ref = O.layer_norm(x.float().cpu(), w.float().cpu(), b.float().cpu())
    assert (y - ref).abs().max() <= ulp_tol(ref)


def test_layernorm_golden_and_fused_residual():
    g = load_golden("blocks")
    from comfy_gen_server_b200 import unet_config as uc
    dim = int(g["dim"])
    sd = uc.seeded_state_dict({"weight": (dim,), "bias": (dim,)}, 100)
    y = N().layer_norm(bf(g["x"]), bf(sd["weight"]), bf(sd["bias"]), 1e-5).float().cpu()
    assert (y - g["layernorm.out"]).abs().max() < 3e-2
    x, r = bf(randn((300, 640), 1)), bf(randn((300, 640), 2))
    w, b = bf(sd["weight"].repeat(5)), bf(sd["bias"].repeat(5))
    xs = x.clone()
    y2 = N().add_layer_norm_(xs, r, w, b, 1e-5)
    assert torch.equal(xs, x + r)                                  # bf16 in-place residual, as the reference
    # (the plain kernel evaluates y = x * (rstd * gamma) + (beta - mean * rstd * gamma) on packed fp32 pairs, the fused
    # one (x - mean) * rstd * gamma + beta: the same function, rounded differently in the last bf16 place)
    y3 = N().layer_norm(x + r, w, b, 1e-5)
    assert (y2.float() - y3.float()).abs().max() <= ulp_tol(y3.float().cpu())


def test_layernorm_errors_and_empty():
    n = N()
    w = bf(torch.ones(12))
    with pytest.raises(RuntimeError):
        n.layer_norm(bf(torch.ones(4, 12)), w, w)                  # C % 8 != 0
    with pytest.raises(RuntimeError):
        n.layer_norm(torch.ones(4, 16), w, w)                      # CPU tensor
    assert n.layer_norm(bf(torch.ones(0, 16)), bf(torch.ones(16)), bf(torch.ones(16))).shape == (0, 16)


# ------------------------------------------------------------------------------------------ GroupNorm
@pytest.mark.parametrize("N_,C,H,W,silu", [(2, 320, 32, 32, True), (2, 640, 16, 16, True), (1, 1280, 8, 8, False),
                                           (3, 64, 12, 12, True), (2, 960, 16, 16, True), (1, 32, 2, 4, False)])
def test_groupnorm_silu_vs_oracle(N_, C, H, W, silu):
    x = bf(randn((N_, C, H, W), 1, 2.0) + 0.5)
    w, b = bf(randn((C,), 2) * 0.2 + 1), bf(randn((C,), 3) * 0.2)
    eps = 1e-5 if silu else 1e-6
    y = N().group_norm(x, w, b, 32, eps, silu).float().cpu()
    ref = torch.nn.functional.group_norm(x.float().cpu(), 32, w.float().cpu(), b.float().cpu(), eps)
    if silu:
        ref = torch.nn.functional.silu(ref)
    assert (y - ref).abs().max() <= ulp_tol(ref)


def test_groupnorm_with_channel_addend():
    """ResBlock form: GroupNorm(h + emb[:, :, None, None]) -> SiLU, the add done in bf16 like the reference."""
    x = bf(randn((2, 320, 16, 16), 1, 2.0))
    e = bf(randn((2, 320), 4))
    w, b = bf(randn((320,), 2) * 0.2 + 1), bf(randn((320,), 3) * 0.2)
    y = N().group_norm(x, w, b, 32, 1e-5, True, e).float().cpu()
    h = (x + e[..., None, None]).float().cpu()
    ref = torch.nn.functional.silu(torch.nn.functional.group_norm(h, 32, w.float().cpu(), b.float().cpu(), 1e-5))
    assert (y - ref).abs().max() <= ulp_tol(ref)


@pytest.mark.parametrize("N_,C,H,W,silu,add", [(2, 320, 32, 32, True, True), (2, 640, 16, 16, True, False),
                                               (1, 1280, 8, 8, False, False), (3, 960, 12, 12, True, True),
                                               (2, 2560, 16, 16, True, False), (1, 256, 3, 5, True, True),
                                               (2, 1920, 32, 32, True, True),
                                               # the VAE decoder's levels: 4 / 8 / 16 channels per group
                                               (1, 128, 64, 64, True, False), (2, 256, 24, 24, True, False),
                                               (1, 512, 16, 16, False, False),
                                               # 64 MB per sample: more than 128 partials, summed by the kernel in between
                                               (1, 128, 512, 512, True, False)])
def test_groupnorm_channels_last_vs_oracle(N_, C, H, W, silu, add):
    x = bf(randn((N_, C, H, W), 1, 2.0) + 0.5).contiguous(memory_format=torch.channels_last)
    e = bf(randn((N_, C), 4)) if add else None
    w, b = bf(randn((C,), 2) * 0.2 + 1), bf(randn((C,), 3) * 0.2)
    eps = 1e-5 if silu else 1e-6
    y = N().group_norm(x, w, b, 32, eps, silu, e)
    assert y.is_contiguous(memory_format=torch.channels_last) and y.shape == x.shape
    h = x if e is None else x + e[..., None, None]
    ref = torch.nn.functional.group_norm(h.float().cpu(), 32, w.float().cpu(), b.float().cpu(), eps)
    if silu:
        ref = torch.nn.functional.silu(ref)
    assert (y.float().cpu() - ref).abs().max() <= ulp_tol(ref)
    # same numbers as the NCHW kernel
    y2 = N().group_norm(x.contiguous(), w, b, 32, eps, silu, e) if (H * W) % 8 == 0 else None
    if y2 is not None:
        assert (y.float() - y2.float()).abs().max() <= ulp_tol(ref)


def test_groupnorm_channels_last_is_reproducible():
    x = bf(randn((4, 640, 64, 64), 1, 2.0)).contiguous(memory_format=torch.channels_last)
    w, b = bf(randn((640,), 2) * 0.2 + 1), bf(randn((640,), 3) * 0.2)
    a = N().group_norm(x, w, b, 32, 1e-5, True)
    for _ in range(3):
        assert torch.equal(a, N().group_norm(x, w, b, 32, 1e-5, True))


@pytest.mark.parametrize("N_,R,C", [(2, 4096, 640), (2, 1024, 1280), (3, 144, 1280), (1, 8, 8), (2, 576, 320), (1, 72, 200)])
def test_transpose_add_exact(N_, R, C):
    x = bf(randn((N_, R, C), 1))
    r = bf(randn((N_, C, R), 2))
    assert torch.equal(N().transpose_add(x), x.transpose(1, 2).contiguous())
    assert torch.equal(N().transpose_add(x, r), x.transpose(1, 2).contiguous() + r)


def test_residual_add_layouts():
    n = N()
    a, b = bf(randn((2, 320, 16, 16), 1)), bf(randn((2, 320, 16, 16), 2))
    assert torch.equal(n.residual_add(a, b), a + b)
    acl, bcl = a.contiguous(memory_format=torch.channels_last), b.contiguous(memory_format=torch.channels_last)
    out = n.residual_add(acl, bcl)
    assert out.is_contiguous(memory_format=torch.channels_last) and torch.equal(out, a + b)
    assert torch.equal(n.residual_add(acl, b), a + b)            # mixed layouts: falls back to torch's add
    tok = bf(randn((2, 256, 320), 3))
    img = tok.view(2, 16, 16, 320).permute(0, 3, 1, 2)            # the channels-last view used by SpatialTransformer
    assert torch.equal(n.residual_add(img, acl), img + acl)
    # per-channel bias (the folded conv bias): one rounding of the fp32 sum
    bias = bf(randn((320,), 4))
    want = (a.float() + b.float() + bias.float().view(1, -1, 1, 1)).to(torch.bfloat16)
    assert torch.equal(n.residual_add(a, b, bias), want)
    out = n.residual_add(acl, bcl, bias)
    assert out.is_contiguous(memory_format=torch.channels_last) and torch.equal(out, want)
    a2, b2 = bf(randn((2, 12, 3, 4), 5)), bf(randn((2, 12, 3, 4), 6))     # H*W and C not multiples of 8: torch path
    want2 = a2 + b2 + bf(randn((12,), 7)).view(1, -1, 1, 1)
    assert torch.equal(n.residual_add(a2, b2, bf(randn((12,), 7))), want2)


def test_cat_channels_and_upsample_nhwc():
    n = N()
    cl = lambda t: t.contiguous(memory_format=torch.channels_last)
    a, b = bf(randn((2, 320, 12, 20), 1)), bf(randn((2, 640, 12, 20), 2))
    out = n.cat_channels(cl(a), cl(b))
    assert out.is_contiguous(memory_format=torch.channels_last) and torch.equal(out, torch.cat([a, b], dim=1))
    assert torch.equal(n.cat_channels(a, b), torch.cat([a, b], dim=1))              # NCHW: torch path
    bias = bf(randn((320,), 9))
    want = torch.cat([a + bias.view(1, -1, 1, 1), b], dim=1)
    assert torch.equal(n.cat_channels(cl(a), cl(b), bias), want)
    assert torch.equal(n.cat_channels(a, b, bias), want)
    up = n.upsample_nearest2x(cl(a))
    assert up.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(up, torch.nn.functional.interpolate(a, scale_factor=2, mode="nearest"))
    assert torch.equal(n.upsample_nearest2x(a), torch.nn.functional.interpolate(a, scale_factor=2, mode="nearest"))
    big = bf(randn((1, 8, 3, 5), 3))                                                  # odd spatial sizes
    assert torch.equal(n.upsample_nearest2x(cl(big)), torch.nn.functional.interpolate(big, scale_factor=2, mode="nearest"))


def test_resblock_bias_folding_matches_plain_path(installed):
    """The reference's ResBlock with install() active (conv biases folded into GroupNorm's addend / the residual add,
    GroupNorm + SiLU fused) vs the same block through the reference's own _forward, op by op."""
    import comfy.ops
    import comfy.ldm.modules.diffusionmodules.openaimodel as U
    from comfy_gen_server_b200.layers import ORIGINAL
    torch.manual_seed(0)
    for cin, cout, cl in ((320, 320, False), (320, 640, True), (640, 320, True)):
        blk = U.ResBlock(cin, 1280, 0.0, out_channels=cout, dtype=torch.bfloat16, device="cuda",
                         operations=comfy.ops.disable_weight_init)
        for prm in blk.parameters():
            torch.nn.init.normal_(prm, std=0.05)
        x = bf(randn((2, cin, 16, 16), 11))
        emb = bf(randn((2, 1280), 12))
        if cl:
            blk = blk.to(memory_format=torch.channels_last)
            x = x.contiguous(memory_format=torch.channels_last)
        got = blk(x, emb)
        want = ORIGINAL["ResBlock._forward"](blk, x, emb)
        err = (got.float() - want.float()).abs().max().item()
        assert err <= 4 * ulp_tol(want), (cin, cout, cl, err)


def test_groupnorm_errors():
    n = N()
    w = bf(torch.ones(64))
    with pytest.raises(RuntimeError):
        n.group_norm(bf(torch.ones(1, 64, 3, 3)), w, w, 32, 1e-5)          # H*W % 8 != 0
    with pytest.raises(RuntimeError):
        n.group_norm(torch.ones(1, 64, 4, 4), w.cpu(), w.cpu(), 32, 1e-5)   # CPU tensor


# ------------------------------------------------------------------------------------------ GEMMs
@pytest.mark.parametrize("M,K,Nn", [(128, 64, 128), (300, 320, 1280), (2 * 1024, 640, 2560), (2 * 256, 1280, 5120),
                                    (1, 128, 128), (4096 + 17, 320, 1280)])
def test_geglu_vs_oracle(M, K, Nn):
    x, w, b = bf(randn((M, K), 1)), bf(randn((2 * Nn, K), 2, K ** -0.5)), bf(randn((2 * Nn,), 3, 0.1))
    y = N().geglu(x, w, b).float().cpu()
    ref = O.geglu(x.float().cpu(), w.float().cpu(), b.float().cpu())
    assert (y - ref).abs().max() <= ulp_tol(ref)


@pytest.mark.parametrize("M,K,Nn", [(256, 64, 128), (300, 320, 1280), (4096 + 17, 640, 2560), (1, 128, 128),
                                    (130, 1280, 256), (16384, 1280, 1280), (2048, 5120, 1280), (8192, 2560, 640),
                                    (19000, 320, 1184), (2048, 640, 32), (40000, 64, 96)])
def test_gemm_cta_pair_mainloop_equals_single_cta(M, K, Nn):
    """The cta_group::2 mainloop (M256 tiles on CTA pairs, half of B per CTA) against the one-CTA mainloop and the
    oracle, for GEGLU and for linear + bias + residual, incl. M that leaves the pair's second CTA partly or wholly
    empty, N that is not a multiple of the 256-column tile (640, 1184, 96, 32: a narrower MMA for what is left) and
    tile counts whose last round is dealt out in 64-column pieces. Same products, same order per output element
    whatever the tiling: bit-equal."""
    n = N()
    lib = n.load_library()
    x, w, b = bf(randn((M, K), 1)), bf(randn((2 * Nn, K), 2, K ** -0.5)), bf(randn((2 * Nn,), 3, 0.1))
    r = bf(randn((M, Nn), 4))
    glu = Nn % 128 == 0
    try:
        assert lib.cgs_gemm_set_variant(1) == 0
        g1, l1 = (n.geglu(x, w, b) if glu else None), n.linear(x, w[:Nn], b[:Nn], residual=r)
        assert lib.cgs_gemm_set_variant(2) == 0
        g2, l2 = (n.geglu(x, w, b) if glu else None), n.linear(x, w[:Nn], b[:Nn], residual=r)
    finally:
        lib.cgs_gemm_set_variant(0)
    if glu:
        ref = O.geglu(x.float().cpu(), w.float().cpu(), b.float().cpu())
        assert (g2.float().cpu() - ref).abs().max() <= ulp_tol(ref)
        assert torch.equal(g1, g2)
    assert torch.equal(l1, l2)
    ref_l = O.linear(x.float().cpu(), w[:Nn].float().cpu(), b[:Nn].float().cpu()) + r.float().cpu()
    assert (l2.float().cpu() - ref_l).abs().max() <= ulp_tol(ref_l)


def test_geglu_golden():
    g = load_golden("blocks")
    from comfy_gen_server_b200 import unet_config as uc
    dim = int(g["dim"])
    sd = uc.seeded_state_dict({"proj.weight": (dim * 8, dim), "proj.bias": (dim * 8,)}, 101)
    y = N().geglu(bf(g["x"]), bf(sd["proj.weight"]), bf(sd["proj.bias"])).float().cpu()
    # vs the reference's fp32 output on fp32 inputs: bf16 rounding of x, W and y (1 ulp each of |y|max ~ 8)
    assert (y - g["geglu.out"]).abs().max() <= 2 * ulp_tol(g["geglu.out"])


@pytest.mark.parametrize("M,K,Nn,res", [(128, 64, 256, False), (256, 320, 320, True), (1000, 1280, 640, True),
                                        (2 * 1024, 5120, 1280, False), (77, 2048, 1280, False)])
def test_linear_vs_oracle(M, K, Nn, res):
    x, w, b = bf(randn((M, K), 1)), bf(randn((Nn, K), 2, K ** -0.5)), bf(randn((Nn,), 3, 0.1))
    r = bf(randn((M, Nn), 4)) if res else None
    y = N().linear(x, w, b, r).float().cpu()
    ref = O.linear(x.float().cpu(), w.float().cpu(), b.float().cpu())
    if res:
        ref = ref + r.float().cpu()
    assert (y - ref).abs().max() <= ulp_tol(ref)


# ------------------------------------------------------------------------------------------ row softmax
@pytest.mark.parametrize("rows,cols,scale", [(7, 8, 1.0), (300, 1024, 512 ** -0.5), (64, 16384, 0.044), (3, 32768, 0.1),
                                             (5, 4104, 2.0)])
def test_softmax_rows_vs_torch(rows, cols, scale):
    """cgs_softmax_rows_fwd against torch.softmax in fp32 (the reference's slice_attention, model.py:175-180)."""
    x = bf(randn((rows, cols), 1, 6.0))
    ref = torch.softmax(x.float().cpu() * scale, dim=-1)
    y = N().softmax_rows_(x.clone(), scale).float().cpu()
    assert (y - ref).abs().max() <= float(ref.max()) * 2 ** -7 + 1e-6
    assert ((y.sum(-1) - 1).abs() < 2e-2).all()
    # a strided view (rows of a wider matrix)
    wide = bf(randn((rows, cols + 64), 2, 3.0))
    keep = wide[:, cols:].clone()
    N().softmax_rows_(wide[:, :cols], scale)
    assert torch.equal(wide[:, cols:], keep)
    assert (wide[:, :cols].float().sum(-1) - 1).abs().max() < 2e-2


# ------------------------------------------------------------------------------------------ 3x3 convolution
CONV_SHAPES = [
    # n, H, W, Cin, Cout, bias, residual
    (2, 16, 16, 64, 64, True, False),         # 8 x 16 pixel tiles
    (1, 32, 32, 320, 320, True, True),        # one 320-column tile (two MMAs per k-step), residual in the epilogue
    (2, 8, 8, 128, 640, False, True),         # tile taller than the image (TW 8, TH 16)
    (1, 24, 24, 64, 96, True, False),         # W = 24: TW 8, TH 16, the second tile row half outside; Cout = 96
    (3, 12, 20, 72, 32, True, True),          # W = 20: TW 4, TH 32; Cin not a multiple of 64 (partial channel block)
    (1, 128, 128, 64, 320, True, False),      # one image row per tile
    (1, 8, 48, 64, 64, False, False),         # three pixel tiles: the last pair's second CTA has none
    (1, 64, 64, 960, 320, True, True),        # SDXL output-block shape (960 -> 320), K = 8640
    (2, 32, 32, 640, 1280, True, False),      # five tile columns
    (1, 16, 16, 64, 960, True, True),         # Cout = 960: three 320-column tiles
]


@pytest.mark.parametrize("n,H,W,Cin,Cout,use_b,use_r", CONV_SHAPES)
def test_conv3x3_vs_oracle(n, H, W, Cin, Cout, use_b, use_r):
    """cgs_conv3x3_nhwc_fwd against F.conv2d in fp32 (the reference's operator, comfy/ops.py:67-80)."""
    cl = torch.channels_last
    x = bf(randn((n, Cin, H, W), 1)).contiguous(memory_format=cl)
    w = bf(randn((Cout, Cin, 3, 3), 2, (9 * Cin) ** -0.5)).contiguous(memory_format=cl)
    b = bf(randn((Cout,), 3, 0.1)) if use_b else None
    r = bf(randn((n, Cout, H, W), 4)).contiguous(memory_format=cl) if use_r else None
    y = N().conv3x3(x, w, b, r)
    assert y.shape == (n, Cout, H, W) and y.is_contiguous(memory_format=cl)
    ref = O.conv3x3(x.cpu(), w.cpu(), None if b is None else b.cpu(), None if r is None else r.cpu())
    assert (y.float().cpu() - ref).abs().max() <= ulp_tol(ref)


def test_conv3x3_channel_slice_and_tile_shapes():
    """The input may be a channel slice of a wider channels-last buffer, and the two tilings of Cout = 320 / 640
    (256-column tiles, 320-column tiles) give bit-identical results."""
    cl = torch.channels_last
    n = N()
    big = bf(randn((2, 192, 16, 32), 1)).contiguous(memory_format=cl)
    x = big[:, 64:128]
    for cout in (320, 640):
        w = bf(randn((cout, 64, 3, 3), 2, 24 ** -1.0)).contiguous(memory_format=cl)
        b = bf(randn((cout,), 3, 0.1))
        ref = O.conv3x3(x.cpu(), w.cpu(), b.cpu())
        outs = []
        for v in (11, 12, 10):
            assert n.load_library().cgs_gemm_set_variant(v) == 0
            outs.append(n.conv3x3(x, w, b))
        assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
        assert (outs[0].float().cpu() - ref).abs().max() <= ulp_tol(ref)


@pytest.mark.parametrize("n,H,W,Cin,Cout,use_b", [(1, 16, 16, 64, 64, True), (2, 8, 32, 128, 320, False),
                                                   (1, 32, 32, 640, 640, True), (2, 12, 20, 72, 96, True),
                                                   (1, 64, 64, 64, 1280, True)])
def test_upsample2x_conv3x3_vs_oracle(n, H, W, Cin, Cout, use_b):
    """cgs_upsample2x_conv3x3_nhwc_fwd (four 2x2 phase convolutions of the small image) against the reference's
    interpolate -> conv2d (openaimodel.py:83-96) in fp32."""
    cl = torch.channels_last
    x = bf(randn((n, Cin, H, W), 1)).contiguous(memory_format=cl)
    w = bf(randn((Cout, Cin, 3, 3), 2, (9 * Cin) ** -0.5)).contiguous(memory_format=cl)
    b = bf(randn((Cout,), 3, 0.1)) if use_b else None
    y = N().upsample2x_conv3x3(x, N().upsample2x_conv_weights(w), b)
    assert y.shape == (n, Cout, 2 * H, 2 * W) and y.is_contiguous(memory_format=cl)
    ref = O.upsample2x_conv3x3(x.cpu(), w.cpu(), None if b is None else b.cpu())
    assert (y.float().cpu() - ref).abs().max() <= ulp_tol(ref)


def test_conv3x3_errors():
    n = N()
    cl = torch.channels_last
    w = bf(randn((64, 64, 3, 3), 2)).contiguous(memory_format=cl)
    with pytest.raises(RuntimeError):
        n.conv3x3(bf(randn((1, 64, 8, 8), 1)), w)                                   # NCHW input
    with pytest.raises(RuntimeError):
        n.conv3x3(bf(randn((1, 64, 8, 8), 1)).contiguous(memory_format=cl), bf(randn((64, 64, 3, 3), 2)))   # NCHW weight
    with pytest.raises(RuntimeError):
        n.conv3x3(bf(randn((1, 60, 8, 8), 1)).contiguous(memory_format=cl),
                  bf(randn((64, 60, 3, 3), 2)).contiguous(memory_format=cl))        # Cin % 8


def test_gemm_errors():
    n = N()
    with pytest.raises(RuntimeError):
        n.geglu(bf(randn((8, 64), 1)), bf(randn((2 * 100, 64), 2)), None)     # N % 128
    with pytest.raises(RuntimeError):
        n.geglu(randn((8, 64), 1), randn((256, 64), 2), None)                 # CPU tensors


# ------------------------------------------------------------------------------------------ CFG / sampler tail
@pytest.mark.parametrize("pred", ["eps", "v"])
@pytest.mark.parametrize("out_dtype", [torch.bfloat16, torch.float32])
def test_cfg_denoise_bit_exact(pred, out_dtype):
    B = 3
    x = randn((B, 4, 24, 24), 1, 5.0)
    eu, ec = randn((B, 4, 24, 24), 2).to(out_dtype), randn((B, 4, 24, 24), 3).to(out_dtype)
    sig = torch.tensor([14.6146, 2.5, 0.0292])
    d = N().cfg_denoise(x.cuda(), eu.cuda(), ec.cuda(), sig.cuda(), 7.0, 0 if pred == "eps" else 1).cpu()
    dc = O.calculate_denoised(sig, ec.float(), x, pred)
    du = O.calculate_denoised(sig, eu.float(), x, pred)
    assert torch.equal(d, O.cfg_function(dc, du, 7.0))
    d1 = N().cfg_denoise(x.cuda(), None, ec.cuda(), sig.cuda(), 1.0, 0 if pred == "eps" else 1).cpu()
    assert torch.equal(d1, dc)


def test_unet_input_matches_reference_formula():
    g = load_golden("schedules")
    xi = N().unet_input(g["ci.x"].cuda(), g["ci.sigma"].cuda(), 2).float().cpu()
    ref = g["calculate_input"].to(torch.bfloat16).float()
    assert torch.equal(xi[:2], ref) and torch.equal(xi[2:], ref)


@pytest.mark.parametrize("key", ["euler_normal_20", "dpmpp_2m_karras_30", "euler_karras_7", "dpmpp_2m_normal_25"])
def test_sampler_known_answer_bit_exact_on_gpu(key, installed):
    """The reference's sampler loops with a Gaussian denoiser, driven through the product's sampler functions and
    the CUDA update kernels: final x bit-identical to the reference's (golden)."""
    from comfy_gen_server_b200 import sampling as ks
    g = load_golden("sampler_kat")
    fn = ks.sample_dpmpp_2m if key.startswith("dpmpp") else ks.sample_euler
    # the denoiser is evaluated with torch on the host copy so that only the update kernels are under test
    model = lambda x, sigma, **kw: (x.cpu() / (1.0 + sigma.cpu().view(-1, 1, 1, 1) ** 2)).cuda()
    trace = []
    xf = fn(model, g[key + ".x0"].cuda(), g[key + ".sigmas"].cuda(), extra_args={},
            callback=lambda d: trace.append(d["denoised"]))
    assert torch.equal(xf.cpu(), g[key + ".final"])
    assert torch.equal(trace[-1].cpu(), g[key + ".denoised_last"])
    const = fn(lambda x, s, **kw: torch.full_like(x, 0.25), g[key + ".x0"].cuda(), g[key + ".sigmas"].cuda())
    assert (const - 0.25).abs().max() == 0.0


@pytest.mark.parametrize("mode", [1, 2, 3])
def test_fused_tail_equals_separate_kernels(mode):
    n = N()
    B = 2
    x = randn((B, 4, 32, 32), 1, 3.0).cuda()
    eu, ec = bf(randn((B, 4, 32, 32), 2)), bf(randn((B, 4, 32, 32), 3))
    dold = randn((B, 4, 32, 32), 4).cuda()
    sig = torch.full((B,), 2.5).cuda()
    c = (2.5, -0.7, 0.0, 0.0) if mode == 1 else (0.8, -0.3, 1.4, 0.4)
    d, xn = n.cfg_sampler_tail(x, eu, ec, sig, 7.0, 0, dold if mode == 3 else None, mode, *c)
    d_ref = n.cfg_denoise(x, eu, ec, sig, 7.0, 0)
    if mode == 1:
        x_ref = n.sampler_step_euler(x, d_ref, c[0], c[1])
    else:
        x_ref = n.sampler_step_dpmpp2m(x, d_ref, dold if mode == 3 else None, *c)
    assert torch.equal(d, d_ref) and torch.equal(xn, x_ref)


def test_tail_large_and_ragged_sizes():
    """2048^2 latents (config 5) and a batch that does not fill a whole block."""
    n = N()
    for B, hw in [(4, 256), (1, 2)]:
        x = randn((B, 4, hw, hw), 1).cuda()
        d = randn((B, 4, hw, hw), 2).cuda()
        xe = n.sampler_step_euler(x, d, 3.0, -1.0).cpu()
        assert torch.equal(xe, x.cpu() + (x.cpu() - d.cpu()) / torch.tensor(3.0) * torch.tensor(-1.0))


def test_tail_errors():
    n = N()
    x = randn((1, 3, 5, 5), 1).cuda()       # 75 elements per sample: not a multiple of 4
    with pytest.raises(RuntimeError):
        n.cfg_denoise(x, None, x, torch.ones(1).cuda(), 1.0, 0)
    with pytest.raises(RuntimeError):
        n.sampler_step_euler(randn((1, 4, 4, 4), 1), randn((1, 4, 4, 4), 1), 1.0, 1.0)      # CPU tensors


# ------------------------------------------------------------------------------------------ guard bands
def test_kernels_write_only_their_output_window():
    """Outputs placed inside larger sentinel-filled buffers: ragged rows, partial column tiles and head dims that are
    not multiples of the 64-wide TMA box must not be written outside the logical output (TMA store clipping, row /
    column predicates). Called through the C-ABI with explicit strides."""
    n = N()
    lib = n.load_library()
    stream = torch.cuda.current_stream().cuda_stream
    SENT = 777.0

    # attention: B 2, H 3, Nq 200 (one full + one ragged query tile), Nk 77, d 40; output rows padded, columns padded
    B, H, Nq, Nk, d = 2, 3, 200, 77, 40
    q, k, v = bf(randn((B, Nq, H * d), 1)), bf(randn((B, Nk, H * d), 2)), bf(randn((B, Nk, H * d), 3))
    row, rows = H * d + 64, Nq + 56
    buf = torch.full((B, rows, row), SENT, dtype=torch.bfloat16, device="cuda")
    rc = lib.cgs_attention_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), buf.data_ptr(), B, H, Nq, Nk, d,
                               q.stride(0), q.stride(1), d, k.stride(0), k.stride(1), d, v.stride(0), v.stride(1), d,
                               buf.stride(0), buf.stride(1), d, float(d) ** -0.5, stream)
    assert rc == 0, lib.cgs_last_error()
    want = n.attention(q, k, v, H)
    assert torch.equal(buf[:, :Nq, :H * d], want)
    assert (buf[:, Nq:, :] == SENT).all() and (buf[:, :, H * d:] == SENT).all()
    # the same for a long key sequence (multi-tile variant, plain stores)
    Nq2, Nk2 = 300, 1500
    q2, k2, v2 = bf(randn((1, Nq2, 64), 4)), bf(randn((1, Nk2, 64), 5)), bf(randn((1, Nk2, 64), 6))
    buf2 = torch.full((1, Nq2 + 100, 64 + 64), SENT, dtype=torch.bfloat16, device="cuda")
    rc = lib.cgs_attention_fwd(q2.data_ptr(), k2.data_ptr(), v2.data_ptr(), buf2.data_ptr(), 1, 1, Nq2, Nk2, 64,
                               q2.stride(0), q2.stride(1), 64, k2.stride(0), k2.stride(1), 64, v2.stride(0),
                               v2.stride(1), 64, buf2.stride(0), buf2.stride(1), 64, 0.125, stream)
    assert rc == 0, lib.cgs_last_error()
    assert torch.equal(buf2[:, :Nq2, :64], n.attention(q2, k2, v2, 1))
    assert (buf2[:, Nq2:, :] == SENT).all() and (buf2[:, :, 64:] == SENT).all()

    # linear / GEGLU: M 200 (ragged row tile), N 96 (partial column tile), output with a wider leading dimension
    M, K, Nn = 200, 64, 96
    x, w, b = bf(randn((M, K), 7)), bf(randn((Nn, K), 8, 0.2)), bf(randn((Nn,), 9))
    ld = Nn + 40
    out = torch.full((M + 60, ld), SENT, dtype=torch.bfloat16, device="cuda")
    rc = lib.cgs_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), None, out.data_ptr(), M, Nn, K, K, K, ld, 0, stream)
    assert rc == 0, lib.cgs_last_error()
    assert torch.equal(out[:M, :Nn], n.linear(x, w, b))
    assert (out[M:] == SENT).all() and (out[:, Nn:] == SENT).all()
    N2 = 128
    w2, b2 = bf(randn((2 * N2, K), 10, 0.2)), bf(randn((2 * N2,), 11))
    ld2 = N2 + 24
    out2 = torch.full((M + 60, ld2), SENT, dtype=torch.bfloat16, device="cuda")
    rc = lib.cgs_geglu_gemm_fwd(x.data_ptr(), w2.data_ptr(), b2.data_ptr(), out2.data_ptr(), M, N2, K, K, K, ld2, stream)
    assert rc == 0, lib.cgs_last_error()
    assert torch.equal(out2[:M, :N2], n.geglu(x, w2, b2))
    assert (out2[M:] == SENT).all() and (out2[:, N2:] == SENT).all()
