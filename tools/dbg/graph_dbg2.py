import sys, os
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
from comfy_gen_server_b200 import integration, models, guider
integration.install()
import comfy.sample, comfy.utils
comfy.utils.PROGRESS_BAR_ENABLED = False
from conftest import load_golden
print("cudnn.benchmark", torch.backends.cudnn.benchmark, "deterministic", torch.backends.cudnn.deterministic,
      "tf32", torch.backends.cuda.matmul.allow_tf32, "bf16 reduced", torch.backends.cuda.matmul.allow_bf16_reduced_precision_reduction)
for fam, golden, sampler, sched in (("tiny_sd15", "e2e_tiny_sd15_euler", "euler", "normal"), ("tiny_sdxl", "e2e_tiny_sdxl_dpmpp2m", "dpmpp_2m", "karras")):
    g = load_golden(golden)
    seed = int(g["seed"])
    patcher, model = models.build_model(fam, "cuda", seed=seed)
    latent = torch.zeros(2, 4, 16, 16)
    noise = comfy.sample.prepare_noise(latent, seed)
    extra = lambda k: {"pooled_output": g[k]} if k in g else {}
    pos, neg = [[g["positive"], extra("pooled_pos")]], [[g["negative"], extra("pooled_neg")]]
    def run(opt):
        p2 = patcher.clone(); p2.model_options["cuda_graph"] = opt
        with torch.inference_mode():
            return comfy.sample.sample(p2, noise, 7, 7.0, sampler, sched, pos, neg, latent, seed=seed, disable_pbar=True)
    seq = [False, False, True, False, True, False]
    outs = [run(o) for o in seq]
    print(fam, "seq", seq, "equal to first:", [bool(torch.equal(outs[0], o)) for o in outs],
          "maxdiff", [round(float((outs[0] - o).abs().max()), 5) for o in outs])
    # direct UNet determinism: same input twice
    x = torch.randn(4, 4, 16, 16, device="cuda").bfloat16()
    t = torch.full((4,), 500.0, device="cuda")
    ctx = torch.randn(4, 77, g["positive"].shape[-1], device="cuda").bfloat16()
    kw = {"y": torch.randn(4, 2816, device="cuda").bfloat16()} if fam == "tiny_sdxl" else {}
    with torch.inference_mode():
        o1 = model.diffusion_model(x, t, context=ctx, transformer_options={}, **kw)
        torch.cuda.empty_cache()
        o2 = model.diffusion_model(x, t, context=ctx, transformer_options={}, **kw)
        junk = torch.empty(12345, device="cuda")
        o3 = model.diffusion_model(x, t, context=ctx, transformer_options={}, **kw)
    print(fam, "unet repeat equal", bool(torch.equal(o1, o2)), bool(torch.equal(o1, o3)))
