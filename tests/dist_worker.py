"""torchrun worker for tests/test_gpu_multi.py: every rank runs sample_distributed on the tiny SDXL model and
rank 0 compares the gathered latents with the single-GPU sample() result."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from comfy_gen_server_b200 import sample as S, unet_config as uc  # noqa: E402
from comfy_gen_server_b200.model_base import build_model  # noqa: E402
from comfy_gen_server_b200.model_management import ModelPatcher  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    family, seed, steps = "tiny_sdxl", 44, 6
    fam = uc.FAMILIES[family]
    patcher = ModelPatcher(build_model(family, dtype=torch.bfloat16, device=dev, seed=seed), dev)
    g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
    batch = int(sys.argv[1])
    latent = torch.zeros(batch, 4, 16, 16)
    noise = S.prepare_noise(latent, seed)
    pos = [[g((1, 77, fam["unet"]["context_dim"]), 1), {"pooled_output": g((1, 1280), 3)}]]
    neg = [[g((1, 154, fam["unet"]["context_dim"]), 2), {"pooled_output": g((1, 1280), 4)}]]
    res = {}
    with torch.inference_mode():
        ref = S.sample(patcher, noise, steps, 7.0, "dpmpp_2m", "karras", pos, neg, latent, seed=seed)
        for mode, split in (("batch", False), ("cfg_split", True)):
            out = S.sample_distributed(patcher, noise, steps, 7.0, "dpmpp_2m", "karras", pos, neg, latent,
                                       cfg_split=split, seed=seed)
            a, b = out.double().flatten(), ref.double().flatten()
            res[mode] = {"max_abs": float((out - ref).abs().max()), "cos": float((a @ b) / (a.norm() * b.norm())),
                         "shape": list(out.shape), "equal": bool(torch.equal(out, ref))}
    if rank == 0:
        print("RESULT " + json.dumps(res), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
