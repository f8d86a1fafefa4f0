import sys, os
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
from comfy_gen_server_b200 import integration, models, _native
integration.install()
import comfy.sample, comfy.utils
comfy.utils.PROGRESS_BAR_ENABLED = False
from conftest import load_golden
g = load_golden("e2e_tiny_sd15_euler")
seed = int(g["seed"])
def fresh():
    return models.build_model("tiny_sd15", "cuda", seed=seed)
x = torch.randn(4, 4, 16, 16, generator=torch.Generator().manual_seed(1)).cuda().bfloat16()
t = torch.full((4,), 500.0, device="cuda")
ctx = torch.randn(4, 77, 96, generator=torch.Generator().manual_seed(2)).cuda().bfloat16()
def fw(model):
    with torch.inference_mode():
        return model.diffusion_model(x, t, context=ctx, transformer_options={})
for label, inst in (("installed", True), ("reference", False)):
    if not inst:
        integration.uninstall()
    p, m = fresh()
    outs = [fw(m) for _ in range(4)]
    print(label, "fresh model forwards equal to #1:", [bool(torch.equal(outs[0], o)) for o in outs],
          [float((outs[0].float() - o.float()).abs().max()) for o in outs])
    p2, m2 = fresh()
    o = fw(m2)
    print(label, "second fresh model forward #1 equal to first model's #1:", bool(torch.equal(o, outs[0])), "to #2:", bool(torch.equal(o, outs[1])))
integration.install()
# which op: hook every module's output on a fresh model, forward twice, find first differing module
p, m = fresh()
rec = []
def hook(name):
    def f(mod, inp, out):
        if isinstance(out, torch.Tensor):
            rec.append((name, out.detach().clone()))
    return f
hs = [mod.register_forward_hook(hook(n)) for n, mod in m.diffusion_model.named_modules() if n]
fw(m); first = list(rec); rec.clear()
fw(m); second = list(rec); rec.clear()
fw(m); third = list(rec)
for (n1, a), (n2, b), (n3, c) in zip(first, second, third):
    if not torch.equal(a, b) or not torch.equal(b, c):
        print("first differing module:", n1, "1v2", float((a.float() - b.float()).abs().max()), "2v3", float((b.float() - c.float()).abs().max()), tuple(a.shape))
        break
else:
    print("no module output differs between forwards 1, 2, 3 (with hooks)")
