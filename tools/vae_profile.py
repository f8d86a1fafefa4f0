"""Top GPU kernels of one 1024^2 VAE decode (the reference's Decoder with install() active): torch.profiler table."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from comfy_gen_server_b200 import integration, unet_config as uc
integration.install()
import comfy.ldm.modules.diffusionmodules.model as vm
dd = {'double_z': True, 'z_channels': 4, 'resolution': 256, 'in_channels': 3, 'out_ch': 3, 'ch': 128,
      'ch_mult': [1, 2, 4, 4], 'num_res_blocks': 2, 'attn_resolutions': [], 'dropout': 0.0}
dec = vm.Decoder(**dd)
shapes = {k: tuple(v.shape) for k, v in dec.state_dict().items()}
dec.load_state_dict(uc.seeded_state_dict(shapes, 1, residual_damp=1.0))
dec = dec.to("cuda", torch.bfloat16).eval()
z = torch.randn(int(os.environ.get("B", 1)), 4, 128, 128, device="cuda").bfloat16()
with torch.inference_mode():
    for _ in range(3):
        dec(z)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        dec(z)
    torch.cuda.synchronize()
    print(f"decode: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms")
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        dec(z)
        torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=90))
