"""The oracle (oracle/oracle.py, a CPU restatement) against outputs of the reference itself
(tests/golden/*.npz, produced by oracle/make_golden.py from /root/reference). CPU only."""
import numpy as np
import pytest
import torch

from conftest import load_golden, cosine
from oracle import oracle as O
from comfy_gen_server_b200 import unet_config as uc

TOL = 2e-5  # fp32 restatement vs fp32 reference: summation order only


def test_attention_matches_reference():
    g = load_golden("attention")
    for name in ("self_d64", "cross77_d40", "cross154_d80", "self_d160_ragged"):
        out = O.attention(g[f"{name}.q"], g[f"{name}.k"], g[f"{name}.v"], int(g[f"{name}.heads"]))
        assert (out - g[f"{name}.out"]).abs().max() < TOL, name


def test_masked_attention_matches_reference():
    """attention.py:111-123 / SDPA attn_mask: additive float masks (4-D causal, 2-D, per-head) and a boolean key mask."""
    g = load_golden("attention")
    for name in ("mask_causal_4d", "mask_float_2d", "mask_float_per_head", "mask_bool_keys"):
        out = O.attention(g[f"{name}.q"], g[f"{name}.k"], g[f"{name}.v"], int(g[f"{name}.heads"]), g[f"{name}.mask"])
        assert (out - g[f"{name}.out"]).abs().max() < TOL, name


def _weights(prefix_shapes, seed):
    return uc.seeded_state_dict(prefix_shapes, seed)


def test_layernorm_geglu_ff_attention_block_match_reference():
    g = load_golden("blocks")
    x, ctx = g["x"], g["context"]
    dim, heads, cd = int(g["dim"]), int(g["heads"]), int(g["ctx_dim"])
    sd = _weights({"weight": (dim,), "bias": (dim,)}, 100)
    assert (O.layer_norm(x, sd["weight"], sd["bias"]) - g["layernorm.out"]).abs().max() < TOL
    sd = _weights({"proj.weight": (dim * 8, dim), "proj.bias": (dim * 8,)}, 101)
    assert (O.geglu(x, sd["proj.weight"], sd["proj.bias"]) - g["geglu.out"]).abs().max() < TOL
    sd = _weights({"net.0.proj.weight": (dim * 8, dim), "net.0.proj.bias": (dim * 8,),
                   "net.2.weight": (dim, dim * 4), "net.2.bias": (dim,)}, 102)
    sd = {"ff." + k: v for k, v in sd.items()}
    assert (O.feed_forward(x, sd, "ff") - g["ff.out"]).abs().max() < TOL
    sd = _weights({"to_q.weight": (dim, dim), "to_k.weight": (dim, cd), "to_v.weight": (dim, cd),
                   "to_out.0.weight": (dim, dim), "to_out.0.bias": (dim,)}, 103)
    sd = {"a." + k: v for k, v in sd.items()}
    assert (O.cross_attention(x, ctx, sd, "a", heads) - g["cross_attention.out"]).abs().max() < TOL


def _block_shapes(dim, cd):
    s = {}
    for a, c in (("attn1", dim), ("attn2", cd)):
        s.update({f"{a}.to_q.weight": (dim, dim), f"{a}.to_k.weight": (dim, c), f"{a}.to_v.weight": (dim, c),
                  f"{a}.to_out.0.weight": (dim, dim), f"{a}.to_out.0.bias": (dim,)})
    s.update({"ff.net.0.proj.weight": (dim * 8, dim), "ff.net.0.proj.bias": (dim * 8,),
              "ff.net.2.weight": (dim, dim * 4), "ff.net.2.bias": (dim,)})
    for n in ("norm1", "norm2", "norm3"):
        s.update({f"{n}.weight": (dim,), f"{n}.bias": (dim,)})
    return s


def test_transformer_block_and_spatial_transformer_match_reference():
    g = load_golden("blocks")
    dim, heads, cd = int(g["dim"]), int(g["heads"]), int(g["ctx_dim"])
    sd = {"b." + k: v for k, v in _weights(_block_shapes(dim, cd), 104).items()}
    out = O.basic_transformer_block(g["x"], g["context"], sd, "b", heads)
    assert (out - g["block.out"]).abs().max() < 1e-4
    for use_linear in (False, True):
        shapes = {"norm.weight": (dim,), "norm.bias": (dim,)}
        pshape = (dim, dim) if use_linear else (dim, dim, 1, 1)
        shapes.update({"proj_in.weight": pshape, "proj_in.bias": (dim,), "proj_out.weight": pshape,
                       "proj_out.bias": (dim,)})
        for d in range(2):
            shapes.update({f"transformer_blocks.{d}.{k}": v for k, v in _block_shapes(dim, cd).items()})
        sd = {"st." + k: v for k, v in _weights(shapes, 105).items()}
        out = O.spatial_transformer(g["image"], g["context"], sd, "st", heads, 2, use_linear)
        assert (out - g[f"spatial_transformer.linear{int(use_linear)}.out"]).abs().max() < 1e-4


def test_schedules_bit_exact():
    g = load_golden("schedules")
    table = O.discrete_sigmas()
    logt = O.discrete_log_sigmas()
    assert torch.equal(table, g["sigmas_table"])
    assert abs(float(table[0]) - 0.0291672) < 1e-6 and abs(float(table[-1]) - 14.6146412) < 1e-5
    for key in ("karras_30", "normal_20", "simple_20", "sgm_uniform_20", "karras_7", "normal_25"):
        sched, steps = key.rsplit("_", 1)
        assert torch.equal(O.calculate_sigmas(table, sched, int(steps)), g[key]), key
    k30 = g["karras_30"]
    assert torch.equal(O.timestep_of_sigma(k30[:-1], logt), g["timestep_karras_30"])
    assert g["timestep_karras_30"][:5].tolist() == [999, 975, 949, 923, 894]   # SURVEY.md 8c KAT (3)
    assert torch.equal(O.sigma_of_timestep(torch.tensor([0.0, 0.5, 10.25, 998.9, 999.0]), logt), g["sigma_of_t"])


def test_model_sampling_formulas_bit_exact():
    g = load_golden("schedules")
    s, x, out = g["ci.sigma"], g["ci.x"], g["ci.model_out"]
    assert torch.equal(O.calculate_input(s, x), g["calculate_input"])
    assert torch.equal(O.calculate_denoised(s, out, x, "eps"), g["denoised_eps"])
    assert torch.equal(O.calculate_denoised(s, out, x, "v"), g["denoised_v"])


@pytest.mark.parametrize("key", ["euler_normal_20", "dpmpp_2m_karras_30", "euler_karras_7", "dpmpp_2m_normal_25"])
def test_sampler_known_answer_bit_exact(key):
    g = load_golden("sampler_kat")
    sampler = "dpmpp_2m" if key.startswith("dpmpp") else "euler"
    model = lambda x, sigma: x / (1.0 + sigma.view(-1, 1, 1, 1) ** 2)
    trace = []
    xf = O.SAMPLERS[sampler](model, g[key + ".x0"].clone(), g[key + ".sigmas"], lambda d: trace.append(d["denoised"]))
    assert torch.equal(xf, g[key + ".final"])
    assert torch.equal(trace[-1], g[key + ".denoised_last"])
    const = O.SAMPLERS[sampler](lambda x, s: torch.full_like(x, 0.25), g[key + ".x0"].clone(), g[key + ".sigmas"])
    assert (const - 0.25).abs().max() == 0.0        # SURVEY.md 8c KAT (1)


def test_sampler_kat_headline_numbers():
    """SURVEY.md 8c KAT (2): Gaussian denoiser, x0 = sigma0 * randn(seed 0)."""
    g = load_golden("sampler_kat")
    assert abs(float(g["euler_normal_20.final"].std()) - 0.944003) < 2e-2
    assert abs(float(g["dpmpp_2m_karras_30.final"].std()) - 1.018406) < 2e-2


def test_cfg_and_cond_batch_order():
    g = load_golden("cfg")
    assert g["batch_order"].tolist() == [1, 0] and g["batch_ctx_shape"].tolist() == [4, 154, 32]
    assert torch.equal(O.cfg_function(g["cond_pred"], g["uncond_pred"], 7.0), g["cfg_function"])
    calls = []

    def apply_model(x, t, ctx):
        calls.append(tuple(ctx.shape))
        return x * 0.5 + ctx.mean(dim=(1, 2)).view(-1, 1, 1, 1) + t.view(-1, 1, 1, 1) * 0.01
    out = O.sampling_function(apply_model, g["x"], g["sigma"], g["uncond"], g["cond"], 7.0)
    assert calls[0] == (4, 154, 32)
    assert (out - g["out_cfg7"]).abs().max() < 1e-6
    out1 = O.sampling_function(apply_model, g["x"], g["sigma"], g["uncond"], g["cond"], 1.0)
    assert (out1 - g["out_cfg1"]).abs().max() < 1e-6


@pytest.mark.parametrize("family,name,hw,steps,sampler,sched", [
    ("tiny_sd15", "e2e_tiny_sd15_euler", 16, 20, "euler", "normal"),
    ("tiny_sd21v", "e2e_tiny_sd21v_euler", 12, 25, "euler", "normal"),
    ("tiny_sdxl", "e2e_tiny_sdxl_dpmpp2m", 16, 30, "dpmpp_2m", "karras"),
])
def test_end_to_end_tiny_matches_reference(family, name, hw, steps, sampler, sched):
    g = load_golden(name)
    fam = uc.FAMILIES[family]
    sd = uc.seeded_state_dict(uc.param_shapes(fam["unet"]), int(g["seed"]))
    latent = torch.zeros(2, 4, hw, hw)
    noise = O.prepare_noise(latent, int(g["seed"]))
    assert torch.equal(noise, g["noise"])
    den = []
    out = O.sample(fam["unet"], sd, fam["prediction"], fam["scale_factor"], noise, steps, 7.0, sampler, sched,
                   g["positive"], g["negative"], latent, g["pooled_pos"], g["pooled_neg"],
                   callback=lambda d: den.append(d["denoised"]))
    assert cosine(den[0], g["denoised_first"]) > 0.999999
    assert cosine(out, g["final"]) > 0.99999
    assert (out - g["final"]).abs().max() < 2e-3 * float(g["final"].abs().max())


def test_workflow_variants_match_reference():
    """denoise < 1, KSamplerAdvanced windows, in-painting mask, batch_index noise: oracle.sample against the
    reference's comfy.sample.sample on the tiny SD1.5 UNet (fp32 CPU both)."""
    g = load_golden("workflows_tiny_sd15")
    fam = uc.FAMILIES["tiny_sd15"]
    seed, steps = int(g["seed"]), int(g["steps"])
    sd = uc.seeded_state_dict(uc.param_shapes(fam["unet"]), seed)
    image = g["image"]
    noise = O.prepare_noise(image, seed)
    run = lambda n, lat, **kw: O.sample(fam["unet"], sd, fam["prediction"], fam["scale_factor"], n, steps, 7.0,
                                        "euler", "normal", g["positive"], g["negative"], lat, **kw)
    cases = {
        "img2img_denoise06": run(noise, image, denoise=0.6),
        "advanced_3_7_leftover": run(torch.zeros_like(image), image, start_step=3, last_step=7),
        "advanced_0_6_full": run(noise, image, start_step=0, last_step=6, force_full_denoise=True),
        "inpaint": run(noise, image, noise_mask=g["mask"]),
        "batch_index": run(g["batch_index_noise"], torch.zeros_like(image)),
    }
    for name, out in cases.items():
        err = (out - g[name]).abs().max().item()
        assert err < 5e-4 * max(1.0, float(g[name].abs().max())), (name, err)
