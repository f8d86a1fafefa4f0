"""Run BASELINE.json configurations through the public API on 1..8 GPUs and report images/s, the placement used, and
parity against the 1-GPU run of the same global batch.

    python tools/run_configs.py cfg3 [--steps N]                                     (1 GPU)
    torchrun --nproc-per-node 2 tools/run_configs.py cfg4 --exchange nccl,peer       (CFG pair split over 2 GPUs)
    torchrun --nproc-per-node 8 tools/run_configs.py cfg3 cfg4 cfg5

cfg2  SD1.5 512^2 batch 16, 20-step Euler            batch-sharded
cfg3  SDXL 1024^2 batch 8, 30-step DPM++ 2M Karras   batch-sharded
cfg4  SD2.1-v 768^2 batch 8, 25-step Euler           CFG pair split (2 GPUs: 8+8; 8 GPUs: 4 shards x 2 halves)
cfg5  SDXL 2048^2 batch 4, 30-step DPM++ 2M Karras   CFG pair split on 8 GPUs (one UNet sample per GPU)
Writes gpurun_out/configs_multi_gpu.json (rank 0)."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (family, global batch, latent hw, steps, sampler, scheduler, cfg_split)
    "cfg2": ("sd15", 16, 64, 20, "euler", "normal", False),
    "cfg3": ("sdxl", 8, 128, 30, "dpmpp_2m", "karras", False),
    "cfg4": ("sd21v", 8, 96, 25, "euler", "normal", True),
    "cfg5": ("sdxl", 4, 256, 30, "dpmpp_2m", "karras", True),
}


def main():
    args = sys.argv[1:]

    def opt(name, default):
        if name in args:
            i = args.index(name)
            v = args[i + 1]
            del args[i:i + 2]
            return v
        return default
    steps_override = opt("--steps", None)
    exchanges = opt("--exchange", "peer").split(",")
    graphs = [g == "1" for g in opt("--graph", "1").split(",")]
    check_parity = opt("--parity", "1") == "1"
    names = args or ["cfg3"]
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from comfy_gen_server_b200 import integration, models, placement, unet_config as uc
    integration.install()
    import comfy.sample
    import comfy.utils
    comfy.utils.PROGRESS_BAR_ENABLED = False
    out = []
    cache = {}
    for name in names:
        fam, batch, hw, steps, sampler, sched, split = CONFIGS[name]
        steps_run = int(steps_override) if steps_override else steps
        if fam not in cache:
            cache.clear()
            torch.cuda.empty_cache()
            cache[fam] = models.build_model(fam, dev, seed=1)[0]
        patcher = cache[fam]
        ctx = uc.FAMILIES[fam]["unet"]["context_dim"]
        g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
        latent = torch.zeros(batch, 4, hw, hw)
        noise = comfy.sample.prepare_noise(latent, 1)
        pos = [[g((1, 77, ctx), 1), {"pooled_output": g((1, 1280), 3)}]]
        neg = [[g((1, 77, ctx), 2), {"pooled_output": g((1, 1280), 4)}]]
        use_split = split and world > 1 and world % 2 == 0
        single = None
        for ex in (exchanges if use_split else ["-"]):
            for graph in graphs:
                p2 = patcher.clone()
                p2.model_options["cuda_graph"] = True if graph else False
                factory = placement.PeerExchange if ex == "peer" else None

                def call(n):
                    if world == 1:
                        return comfy.sample.sample(p2, noise, n, 7.0, sampler, sched, pos, neg, latent, seed=1,
                                                   disable_pbar=True)
                    return placement.sample_distributed(p2, noise, n, 7.0, sampler, sched, pos, neg, latent, seed=1,
                                                        cfg_split=use_split, exchange_factory=factory)
                with torch.inference_mode():
                    call(min(4, steps_run))                                  # warm-up (captures the step graph)
                    if world > 1:
                        dist.barrier()
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    res = call(steps_run)
                    torch.cuda.synchronize()
                    dt = time.perf_counter() - t0
                    if world > 1:
                        t = torch.tensor([dt], device=dev, dtype=torch.float64)
                        dist.all_reduce(t, op=dist.ReduceOp.MAX)
                        dt = float(t[0])
                    r = {"config": name, "n_gpus": world, "placement": "cfg_split" if use_split else "batch",
                         "exchange": ex, "step_graph": graph, "global_batch": batch, "steps_run": steps_run,
                         "ms_per_step": dt / steps_run * 1e3,
                         "images_per_s_at_full_schedule": batch / (dt / steps_run * steps),
                         "finite": bool(torch.isfinite(res).all()), "sum": float(res.double().sum()),
                         "peak_mem_gb": torch.cuda.max_memory_allocated() / 2 ** 30}
                    if check_parity and world > 1:
                        if single is None:
                            p1 = patcher.clone()
                            p1.model_options["cuda_graph"] = False
                            single = comfy.sample.sample(p1, noise, steps_run, 7.0, sampler, sched, pos, neg, latent,
                                                         seed=1, disable_pbar=True)
                        a, b = res.double().flatten(), single.double().flatten()
                        r["parity_vs_1gpu"] = {"cosine": float((a @ b) / (a.norm() * b.norm())),
                                               "max_abs": float((res - single).abs().max()),
                                               "ref_abs_max": float(single.abs().max())}
                if rank == 0:
                    print(json.dumps(r), flush=True)
                out.append(r)
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        path = os.path.join(ROOT, "gpurun_out", f"configs_multi_gpu_n{world}.json")
        json.dump(out, open(path, "w"), indent=1)
    if world > 1:
        for p in cache.values():
            p.model.__dict__.pop("_cgs_step_graphs", None)
        torch.cuda.synchronize()
        dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
