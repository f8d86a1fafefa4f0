"""Generate tests/golden/*.npz by running the REFERENCE ITSELF (imported from /root/reference, CPU, fp32).

Run in the build container only (the reference does not exist on the GPU box):
    python oracle/make_golden.py [--skip-sd15]

The reference's own tests hold no golden vectors for this path, so these
fixtures -- outputs of the reference's functions on seeded inputs -- are what pins the oracle
(tests/test_oracle_golden.py) and, through it and directly, the CUDA path (tests -m gpu).
Nothing is copied from the reference: it is imported, called, and its outputs are saved.
"""
from __future__ import annotations

import os
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("CGS_REFERENCE", "/root/reference")
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

# shims documented in SURVEY.md section 8c
sys.modules["torchsde"] = types.ModuleType("torchsde")
import comfy.cli_args  # noqa: E402
comfy.cli_args.args.cpu = True
import comfy.model_management  # noqa: E402,F401
import comfy.ops  # noqa: E402
import comfy.utils  # noqa: E402
import comfy.model_base as mb  # noqa: E402
import comfy.model_detection as md  # noqa: E402
import comfy.model_patcher  # noqa: E402
import comfy.sample  # noqa: E402
import comfy.samplers  # noqa: E402
import comfy.model_sampling  # noqa: E402
import comfy.k_diffusion.sampling as kds  # noqa: E402
import comfy.ldm.modules.attention as ra  # noqa: E402

from comfy_gen_server_b200 import unet_config as uc  # noqa: E402

comfy.utils.PROGRESS_BAR_ENABLED = False
OPS = comfy.ops.disable_weight_init


def randn(shape, seed, scale=1.0):
    return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale


def save(name, **arrays):
    os.makedirs(GOLD, exist_ok=True)
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v))
                                 for k, v in arrays.items()})
    print(f"wrote {path} ({os.path.getsize(path) / 1024:.1f} KiB)")


def module_weights(module, seed):
    shapes = {k: tuple(v.shape) for k, v in module.state_dict().items()}
    sd = uc.seeded_state_dict(shapes, seed)
    module.load_state_dict(sd)
    return sd


# ----------------------------------------------------------------------------------------------
def gen_attention():
    out = {}
    cases = {"self_d64": (2, 64, 64, 2, 64), "cross77_d40": (2, 96, 77, 2, 40), "cross154_d80": (1, 50, 154, 3, 80),
             "self_d160_ragged": (1, 72, 72, 2, 160)}
    for name, (b, nq, nk, h, d) in cases.items():
        q, k, v = randn((b, nq, h * d), 1), randn((b, nk, h * d), 2), randn((b, nk, h * d), 3)
        o_basic = ra.attention_basic(q, k, v, h)
        o_torch = ra.attention_pytorch(q, k, v, h)
        assert (o_basic - o_torch).abs().max() < 1e-5
        out.update({f"{name}.q": q, f"{name}.k": k, f"{name}.v": v, f"{name}.out": o_torch,
                    f"{name}.heads": np.int64(h)})
    # masked attention (attention.py:111-123, 345): additive float (causal-like, per-head, 2-D) and boolean keep-masks
    g = torch.Generator().manual_seed(77)
    masked = {
        "mask_causal_4d": (2, 40, 40, 2, 64, torch.triu(torch.full((40, 40), float("-inf")), diagonal=1)[None, None]),
        "mask_float_2d": (2, 48, 77, 2, 40, torch.randn((48, 77), generator=g) * 2.0),
        "mask_float_per_head": (1, 33, 150, 3, 64, torch.randn((1, 3, 33, 150), generator=g)),
        "mask_bool_keys": (2, 24, 77, 2, 64, (torch.rand((2, 1, 1, 77), generator=g) > 0.3)),
    }
    for name, (b, nq, nk, h, d, mask) in masked.items():
        q, k, v = randn((b, nq, h * d), 4), randn((b, nk, h * d), 5), randn((b, nk, h * d), 6)
        o_torch = ra.attention_pytorch(q, k, v, h, mask=mask)
        if mask.dtype != torch.bool:        # attention_basic takes 2-D / 3-D / [bs, x, Nq, Nk] float masks
            o_basic = ra.attention_basic(q, k, v, h, mask=mask)
            assert (o_basic - o_torch).abs().max() < 1e-5, name
        out.update({f"{name}.q": q, f"{name}.k": k, f"{name}.v": v, f"{name}.out": o_torch, f"{name}.mask": mask,
                    f"{name}.heads": np.int64(h)})
    save("attention", **out)


def gen_blocks():
    dim, heads, dh, ctx_dim = 128, 2, 64, 96
    x = randn((2, 64, dim), 10)
    ctx = randn((2, 77, ctx_dim), 11)
    out = {"x": x, "context": ctx, "dim": np.int64(dim), "heads": np.int64(heads), "ctx_dim": np.int64(ctx_dim)}

    ln = OPS.LayerNorm(dim)
    module_weights(ln, 100)
    out["layernorm.out"] = ln(x)

    g = ra.GEGLU(dim, dim * 4, operations=OPS)
    module_weights(g, 101)
    out["geglu.out"] = g(x)

    ff = ra.FeedForward(dim, glu=True, operations=OPS)
    module_weights(ff, 102)
    out["ff.out"] = ff(x)

    ca = ra.CrossAttention(dim, context_dim=ctx_dim, heads=heads, dim_head=dh, operations=OPS)
    module_weights(ca, 103)
    out["cross_attention.out"] = ca(x, ctx)

    blk = ra.BasicTransformerBlock(dim, heads, dh, context_dim=ctx_dim, checkpoint=False, operations=OPS)
    module_weights(blk, 104)
    out["block.out"] = blk(x.clone(), context=ctx, transformer_options={})

    img = randn((2, dim, 8, 8), 12)
    out["image"] = img
    for use_linear in (False, True):
        st = ra.SpatialTransformer(dim, heads, dh, depth=2, context_dim=ctx_dim, use_linear=use_linear,
                                   use_checkpoint=False, operations=OPS)
        module_weights(st, 105)
        out[f"spatial_transformer.linear{int(use_linear)}.out"] = st(img.clone(), ctx, transformer_options={})
    save("blocks", **out)


def gen_schedules():
    ms = comfy.model_sampling.ModelSamplingDiscrete()
    out = {"sigmas_table": ms.sigmas}
    for sched, steps in (("karras", 30), ("normal", 20), ("simple", 20), ("sgm_uniform", 20), ("karras", 7),
                         ("normal", 25)):
        out[f"{sched}_{steps}"] = comfy.samplers.calculate_sigmas(ms, sched, steps)
    k30 = out["karras_30"]
    out["timestep_karras_30"] = ms.timestep(k30[:-1])
    out["sigma_of_t"] = ms.sigma(torch.tensor([0.0, 0.5, 10.25, 998.9, 999.0]))
    sig = torch.tensor([0.5, 3.0])
    x = randn((2, 4, 8, 8), 20)
    eps = randn((2, 4, 8, 8), 21)
    out["ci.sigma"], out["ci.x"], out["ci.model_out"] = sig, x, eps
    out["calculate_input"] = comfy.model_sampling.EPS.calculate_input(ms, sig, x)
    out["denoised_eps"] = comfy.model_sampling.EPS.calculate_denoised(ms, sig, eps, x)
    out["denoised_v"] = comfy.model_sampling.V_PREDICTION.calculate_denoised(ms, sig, eps, x)
    save("schedules", **out)


def gen_sampler_kat():
    ms = comfy.model_sampling.ModelSamplingDiscrete()
    out = {}
    for sampler, sched, steps in (("euler", "normal", 20), ("dpmpp_2m", "karras", 30), ("euler", "karras", 7),
                                  ("dpmpp_2m", "normal", 25)):
        sigmas = comfy.samplers.calculate_sigmas(ms, sched, steps)
        x0 = randn((2, 4, 8, 8), 0) * sigmas[0]
        trace = []
        model = lambda x, sigma, **kw: x / (1.0 + sigma.view(-1, 1, 1, 1) ** 2)
        fn = getattr(kds, "sample_" + sampler)
        xf = fn(model, x0.clone(), sigmas, extra_args={}, callback=lambda d: trace.append(d["denoised"].clone()),
                disable=True)
        key = f"{sampler}_{sched}_{steps}"
        out[key + ".x0"], out[key + ".final"], out[key + ".sigmas"] = x0, xf, sigmas
        out[key + ".denoised_last"] = trace[-1]
        const = fn(lambda x, s, **kw: torch.full_like(x, 0.25), x0.clone(), sigmas, extra_args={}, disable=True)
        assert (const - 0.25).abs().max() == 0.0
    save("sampler_kat", **out)


class _StubModel:
    """Records what calc_cond_batch feeds apply_model; returns a deterministic function of its inputs."""
    def __init__(self):
        self.calls = []

    def memory_required(self, shape):
        return 0

    def apply_model(self, x, t, c_crossattn=None, transformer_options=None, **kw):
        self.calls.append({"x": tuple(x.shape), "ctx": tuple(c_crossattn.shape),
                           "cond_or_uncond": list(transformer_options["cond_or_uncond"])})
        return x * 0.5 + c_crossattn.mean(dim=(1, 2)).view(-1, 1, 1, 1) + t.view(-1, 1, 1, 1) * 0.01


def gen_cfg():
    import comfy.conds
    x = randn((2, 4, 8, 8), 30)
    sigma = torch.tensor([2.5, 2.5])
    cpos, cneg = randn((1, 77, 32), 31), randn((1, 154, 32), 32)
    mk = lambda c: [{"model_conds": {"c_crossattn": comfy.conds.CONDCrossAttn(c)}}]
    stub = _StubModel()
    out_cfg = comfy.samplers.sampling_function(stub, x, sigma, mk(cneg), mk(cpos), 7.0, model_options={})
    assert stub.calls[0]["cond_or_uncond"] == [1, 0] and stub.calls[0]["ctx"] == (4, 154, 32)
    stub1 = _StubModel()
    out_1 = comfy.samplers.sampling_function(stub1, x, sigma, mk(cneg), mk(cpos), 1.0, model_options={})
    assert stub1.calls[0]["cond_or_uncond"] == [0]
    cp, up = randn((2, 4, 8, 8), 33), randn((2, 4, 8, 8), 34)
    save("cfg", x=x, sigma=sigma, cond=cpos, uncond=cneg, out_cfg7=out_cfg, out_cfg1=out_1,
         cond_pred=cp, uncond_pred=up, cfg_function=comfy.samplers.cfg_function(None, cp, up, 7.0, x, sigma),
         batch_order=np.array(stub.calls[0]["cond_or_uncond"]), batch_ctx_shape=np.array(stub.calls[0]["ctx"]))


# ----------------------------------------------------------------------------------------------
def build_reference_model(family: str, seed: int):
    fam = uc.FAMILIES[family]
    cfg = dict(fam["unet"])
    cfg["dtype"] = torch.float32
    mc = md.model_config_from_unet_config(dict(cfg))
    if mc is None:  # tiny configs are not in the reference's supported list: wrap the closest family
        import comfy.supported_models as sm
        base = {"tiny_sd15": sm.SD15, "tiny_sd21v": sm.SD20, "tiny_sdxl": sm.SDXL}[family]
        mc = base(dict(cfg))
    mc.set_inference_dtype(torch.float32, None)
    dev = torch.device("cpu")
    if fam["prediction"] == "v":
        model = mb.BaseModel(mc, model_type=mb.ModelType.V_PREDICTION, device=dev)
    elif fam["adm"]:
        model = mb.SDXL(mc, model_type=mb.ModelType.EPS, device=dev)
    else:
        model = mb.BaseModel(mc, model_type=mb.ModelType.EPS, device=dev)
    shapes = {k: tuple(v.shape) for k, v in model.diffusion_model.state_dict().items()}
    assert shapes == uc.param_shapes(fam["unet"]), "param_shapes diverged from the reference UNet"
    model.diffusion_model.load_state_dict(uc.seeded_state_dict(shapes, seed))
    return comfy.model_patcher.ModelPatcher(model, load_device=dev, offload_device=dev), model


def gen_e2e(family, seed, batch, hw, steps, sampler, scheduler, pos_len, neg_len, name, record_steps=True):
    fam = uc.FAMILIES[family]
    ctx_dim = fam["unet"]["context_dim"]
    patcher, model = build_reference_model(family, seed)
    latent = torch.zeros(batch, 4, hw, hw)
    noise = comfy.sample.prepare_noise(latent, seed)
    pos, neg = randn((1, pos_len, ctx_dim), seed + 1), randn((1, neg_len, ctx_dim), seed + 2)
    pp, pn = randn((1, 1280), seed + 3), randn((1, 1280), seed + 4)
    positive = [[pos, {"pooled_output": pp}]]
    negative = [[neg, {"pooled_output": pn}]]
    den = []
    cb = (lambda i, denoised, x, total: den.append(denoised.clone())) if record_steps else None
    t0 = time.time()
    with torch.inference_mode():
        out = comfy.sample.sample(patcher, noise, steps, 7.0, sampler, scheduler, positive, negative, latent,
                                  seed=seed, disable_pbar=True, callback=cb)
        # one raw UNet evaluation at the first sigma, to pin the network separately from the sampler
        sig0 = comfy.samplers.calculate_sigmas(model.model_sampling, scheduler, steps)[0]
        x0 = model.model_sampling.noise_scaling(sig0, noise.clone(), latent, True)
        kw = {}
        if fam["adm"]:
            kw["y"] = model.encode_adm(pooled_output=pp, width=hw * 8, height=hw * 8, device="cpu")
            kw["y"] = kw["y"].repeat(batch, 1)
        d0 = model.apply_model(x0, sig0 * torch.ones(batch), c_crossattn=pos.repeat(batch, 1, 1),
                               transformer_options={}, **kw)
    dt = time.time() - t0
    print(f"{name}: reference ran {steps} steps in {dt:.1f}s; out std {out.std():.4f} mean {out.mean():.4f} "
          f"cond-denoised(step0) std {d0.std():.4f}")
    arrays = dict(final=out, noise=noise, positive=pos, negative=neg, pooled_pos=pp, pooled_neg=pn,
                  seed=np.int64(seed), steps=np.int64(steps), cond_denoised_step0=d0,
                  ref_seconds=np.float64(dt), ref_threads=np.int64(torch.get_num_threads()))
    if record_steps:
        arrays["denoised_first"] = den[0]
        arrays["denoised_last"] = den[-1]
    save(name, **arrays)


def gen_workflows():
    """The sampling node's other call shapes (nodes.py:1420-1439, 1753-1785), each through comfy.sample.sample as
    common_ksampler drives it: img2img (denoise < 1), a KSamplerAdvanced window with and without leftover noise,
    in-painting (noise_mask) and per-image noise (batch_index)."""
    family, seed, hw, steps = "tiny_sd15", 45, 16, 10
    ctx_dim = uc.FAMILIES[family]["unet"]["context_dim"]
    patcher, model = build_reference_model(family, seed)
    pos, neg = randn((1, 77, ctx_dim), seed + 1), randn((1, 77, ctx_dim), seed + 2)
    positive, negative = [[pos, {}]], [[neg, {}]]
    image = randn((2, 4, hw, hw), seed + 5, 0.8)
    mask = torch.zeros(2, 1, hw, hw)
    mask[:, :, :, hw // 2:] = 1.0
    run = lambda noise, latent, **kw: comfy.sample.sample(patcher, noise, steps, 7.0, "euler", "normal", positive, negative,
                                                         latent, seed=seed, disable_pbar=True, **kw)
    out = dict(positive=pos, negative=neg, image=image, mask=mask, seed=np.int64(seed), steps=np.int64(steps))
    with torch.inference_mode():
        noise = comfy.sample.prepare_noise(image, seed)
        out["img2img_denoise06"] = run(noise, image, denoise=0.6)
        zeros = torch.zeros_like(image)
        out["advanced_3_7_leftover"] = run(zeros, image, disable_noise=True, start_step=3, last_step=7,
                                           force_full_denoise=False)
        out["advanced_0_6_full"] = run(noise, image, start_step=0, last_step=6, force_full_denoise=True)
        out["inpaint"] = run(noise, image, noise_mask=mask)
        inds = [3, 1]
        out["batch_index_noise"] = comfy.sample.prepare_noise(image, seed, inds)
        out["batch_index"] = run(out["batch_index_noise"], torch.zeros_like(image))
    save("workflows_tiny_sd15", **out)


def gen_sdxl_headline(seed=46, hw=128, steps=30, first_steps=2, name="e2e_sdxl_headline_2steps"):
    """The headline configuration at full size (BASELINE.json configs[2]: SDXL 1024^2, DPM++ 2M Karras, CFG 7; one
    image): the first `first_steps` steps of the 30-step schedule through the reference's fp32 CPU path
    (comfy.sample.sample with last_step), which pins the full-size network and the multistep update."""
    fam = uc.FAMILIES["sdxl"]
    ctx_dim = fam["unet"]["context_dim"]
    patcher, model = build_reference_model("sdxl", seed)
    latent = torch.zeros(1, 4, hw, hw)
    noise = comfy.sample.prepare_noise(latent, seed)
    pos, neg = randn((1, 77, ctx_dim), seed + 1), randn((1, 77, ctx_dim), seed + 2)
    pp, pn = randn((1, 1280), seed + 3), randn((1, 1280), seed + 4)
    t0 = time.time()
    with torch.inference_mode():
        out = comfy.sample.sample(patcher, noise, steps, 7.0, "dpmpp_2m", "karras", [[pos, {"pooled_output": pp}]],
                                  [[neg, {"pooled_output": pn}]], latent, seed=seed, disable_pbar=True,
                                  last_step=first_steps, force_full_denoise=False)
    dt = time.time() - t0
    print(f"e2e_sdxl_headline: reference ran {first_steps} of {steps} steps in {dt:.1f}s; out std {out.std():.4f}")
    save(name, x_after=out, noise=noise, positive=pos, negative=neg, pooled_pos=pp, pooled_neg=pn,
         seed=np.int64(seed), steps=np.int64(steps), first_steps=np.int64(first_steps), ref_seconds=np.float64(dt),
         ref_threads=np.int64(torch.get_num_threads()))


def gen_sd21v_full(seed=47, hw=96, steps=25, first_steps=3, name="e2e_sd21v_full_3steps"):
    """BASELINE.json configs[3] at full size (SD2.1-v 768^2, Euler / normal, CFG 7, v-prediction; one image): the first
    steps through the reference's fp32 CPU path. Token counts 9216 / 2304 / 576 / 144 are not multiples of 128."""
    fam = uc.FAMILIES["sd21v"]
    ctx_dim = fam["unet"]["context_dim"]
    patcher, model = build_reference_model("sd21v", seed)
    latent = torch.zeros(1, 4, hw, hw)
    noise = comfy.sample.prepare_noise(latent, seed)
    pos, neg = randn((1, 77, ctx_dim), seed + 1), randn((1, 77, ctx_dim), seed + 2)
    t0 = time.time()
    with torch.inference_mode():
        out = comfy.sample.sample(patcher, noise, steps, 7.0, "euler", "normal", [[pos, {}]], [[neg, {}]], latent,
                                  seed=seed, disable_pbar=True, last_step=first_steps, force_full_denoise=False)
    dt = time.time() - t0
    print(f"e2e_sd21v_full: reference ran {first_steps} of {steps} steps in {dt:.1f}s; out std {out.std():.4f}")
    save(name, x_after=out, noise=noise, positive=pos, negative=neg, seed=np.int64(seed),
         steps=np.int64(steps), first_steps=np.int64(first_steps), ref_seconds=np.float64(dt),
         ref_threads=np.int64(torch.get_num_threads()))


VAE_DDCONFIG = {'double_z': True, 'z_channels': 4, 'resolution': 256, 'in_channels': 3, 'out_ch': 3, 'ch': 128,
                'ch_mult': [1, 2, 4, 4], 'num_res_blocks': 2, 'attn_resolutions': [], 'dropout': 0.0}


def gen_vae_decode(seed=48, hw=32):
    """The step after the path: the SD / SDXL VAE decoder (comfy/ldm/modules/diffusionmodules/
    model.py:542-660, the configuration of comfy/sd.py:213) on a seeded latent, fp32 on the CPU."""
    import comfy.ldm.modules.diffusionmodules.model as vm
    dec = vm.Decoder(**VAE_DDCONFIG)
    shapes = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
    dec.load_state_dict(uc.seeded_state_dict(shapes, seed, residual_damp=1.0))
    z = randn((2, 4, hw, hw), seed + 1)
    with torch.inference_mode():
        img = dec(z)
    print(f"vae_decode: out {tuple(img.shape)} std {img.std():.4f}")
    save("vae_decode", z=z, image=img, seed=np.int64(seed))


def main():
    torch.manual_seed(0)
    if "--only-vae" in sys.argv:
        gen_vae_decode()
        return
    if "--only-sd21v-full" in sys.argv:
        gen_sd21v_full()
        return
    if "--only-sdxl-headline" in sys.argv:
        gen_sdxl_headline()
        return
    if "--only-workflows" in sys.argv:
        gen_workflows()
        return
    if "--only-full-schedules" in sys.argv:
        # north_star's bar is the cosine of the FINAL latents after the whole schedule: the headline configuration
        # (30 DPM++ 2M steps, ~20 min on 8 cores) and config 4 (25 Euler steps, v-prediction) at full size
        gen_sd21v_full(first_steps=25, name="e2e_sd21v_full_schedule")
        gen_sdxl_headline(first_steps=30, name="e2e_sdxl_headline_full_schedule")
        return
    gen_attention()
    if "--only-attention" in sys.argv:
        return
    gen_workflows()
    gen_blocks()
    gen_schedules()
    gen_sampler_kat()
    gen_cfg()
    gen_e2e("tiny_sd15", 42, 2, 16, 20, "euler", "normal", 77, 77, "e2e_tiny_sd15_euler")
    gen_e2e("tiny_sd21v", 43, 2, 12, 25, "euler", "normal", 77, 154, "e2e_tiny_sd21v_euler")
    gen_e2e("tiny_sdxl", 44, 2, 16, 30, "dpmpp_2m", "karras", 77, 77, "e2e_tiny_sdxl_dpmpp2m")
    if "--skip-sdxl" not in sys.argv:
        gen_sdxl_headline()
        gen_sd21v_full()
    if "--skip-sd15" not in sys.argv:
        # BASELINE.json configs[0]: SD1.5 512x512 batch 1, 20-step Euler, CFG 7.0, CPU reference path
        gen_e2e("sd15", 42, 1, 64, 20, "euler", "normal", 77, 77, "e2e_sd15_config1")


if __name__ == "__main__":
    main()
