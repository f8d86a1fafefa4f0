import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tests"))
import torch
from comfy_gen_server_b200 import integration, models, guider
integration.install()
import comfy.sample, comfy.utils
comfy.utils.PROGRESS_BAR_ENABLED = False
from conftest import load_golden
fam, golden, sampler, sched = sys.argv[1:5] if len(sys.argv) > 4 else ("tiny_sd15", "e2e_tiny_sd15_euler", "euler", "normal")
g = load_golden(golden)
seed = int(g["seed"])
patcher, model = models.build_model(fam, "cuda", seed=seed)
latent = torch.zeros(2, 4, 16, 16)
noise = comfy.sample.prepare_noise(latent, seed)
extra = lambda k: {"pooled_output": g[k]} if k in g else {}
pos, neg = [[g["positive"], extra("pooled_pos")]], [[g["negative"], extra("pooled_neg")]]
def run(opt):
    p2 = patcher.clone(); p2.model_options["cuda_graph"] = opt
    xs = []
    with torch.inference_mode():
        out = comfy.sample.sample(p2, noise, 7, 7.0, sampler, sched, pos, neg, latent, seed=seed, disable_pbar=True,
                                  callback=lambda i, d, x, t: xs.append((d.clone(), x.clone())))
    return out, xs
a, xa = run(False)
b, xb = run(True)
for i, ((da, xa_), (db, xb_)) in enumerate(zip(xa, xb)):
    print(i, "denoised diff", float((da - db).abs().max()), "x diff", float((xa_ - xb_).abs().max()))
a2, _ = run(False)
print("eager repeat equal", torch.equal(a, a2), "graph equal", torch.equal(a, b))
model.__dict__.pop("_cgs_step_graphs", None)
b2, xb2 = run(True)
print("graph recapture equal to first graph", torch.equal(b, b2))
