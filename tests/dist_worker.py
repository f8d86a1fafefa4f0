"""torchrun worker for tests/test_gpu_multi.py: every rank runs placement.sample_distributed on the tiny SDXL model
(the reference's classes + install()) and rank 0 compares the gathered latents with the single-GPU
comfy.sample.sample result. Modes: batch sharding, CFG-pair split over NCCL, CFG-pair split with the exchange fused
into the tail kernel through peer memory."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from comfy_gen_server_b200 import integration, models, placement, unet_config as uc
    integration.install()
    import comfy.sample
    import comfy.utils
    comfy.utils.PROGRESS_BAR_ENABLED = False
    family, seed, steps = "tiny_sdxl", 44, 6
    fam = uc.FAMILIES[family]
    patcher, _ = models.build_model(family, dev, seed=seed)
    g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
    batch = int(sys.argv[1])
    latent = torch.zeros(batch, 4, 16, 16)
    noise = comfy.sample.prepare_noise(latent, seed)
    pos = [[g((1, 77, fam["unet"]["context_dim"]), 1), {"pooled_output": g((1, 1280), 3)}]]
    neg = [[g((1, 154, fam["unet"]["context_dim"]), 2), {"pooled_output": g((1, 1280), 4)}]]
    res = {}
    modes = [("batch", False, None, False), ("cfg_split", True, None, False), ("cfg_split_graph", True, None, True)]
    if hasattr(placement, "PeerExchange"):
        modes += [("cfg_split_peer", True, placement.PeerExchange, False),
                  ("cfg_split_peer_graph", True, placement.PeerExchange, True)]
    with torch.inference_mode():
        ref = comfy.sample.sample(patcher, noise, steps, 7.0, "dpmpp_2m", "karras", pos, neg, latent, seed=seed,
                                  disable_pbar=True)
        for mode, split, factory, graph in modes:
            p2 = patcher.clone()
            p2.model_options["cuda_graph"] = True if graph else False
            out = placement.sample_distributed(p2, noise, steps, 7.0, "dpmpp_2m", "karras", pos, neg, latent,
                                               cfg_split=split, seed=seed, exchange_factory=factory)
            a, b = out.double().flatten(), ref.double().flatten()
            res[mode] = {"max_abs": float((out - ref).abs().max()), "cos": float((a @ b) / (a.norm() * b.norm())),
                         "shape": list(out.shape), "equal": bool(torch.equal(out, ref)),
                         "sum": float(out.double().sum())}
    if rank == 0:
        print("RESULT " + json.dumps(res), flush=True)
    # captured step graphs hold NCCL kernels of the pair communicators: drop them before the communicators go, and do
    # not let a slow communicator teardown keep the test waiting
    patcher.model.__dict__.pop("_cgs_step_graphs", None)
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
