"""Step time of the benchmark workload (SDXL 1024^2, DPM++ 2M, CFG 7) at the per-GPU batch sizes of the strong-scaling
run (8 / 4 / 2 / 1 images per GPU), on ONE GPU: the timed loop of bench.py has no collective, so a rank of the N-GPU run
does exactly this. Reports launch-by-launch and step-graph-replay times and the host's enqueue time.
usage: python tools/small_batch.py [batches ...] [--family sdxl] [--hw 128] [--steps 10]"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    args = sys.argv[1:]
    def opt(name, default):
        if name in args:
            i = args.index(name)
            v = args[i + 1]
            del args[i:i + 2]
            return v
        return default
    family, hw, steps = opt("--family", "sdxl"), int(opt("--hw", "128")), int(opt("--steps", "10"))
    sampler, sched = opt("--sampler", "dpmpp_2m"), opt("--sched", "karras")
    profile = "--profile" in args
    if profile:
        args.remove("--profile")
    batches = [int(a) for a in args] or [8, 4, 2, 1]
    from comfy_gen_server_b200 import _native, integration, models, sampling as ks, unet_config as uc
    from comfy_gen_server_b200.guider import FusedDenoise
    integration.install()
    if os.environ.get("CGS_MIN_ROWS"):                              # A/B: from how many rows on the linears run on this library
        from comfy_gen_server_b200 import block
        block.NATIVE_LINEAR_MIN_ROWS = int(os.environ["CGS_MIN_ROWS"])
    for f in os.environ.get("CGS_FEATURES_OFF", "").split(","):      # A/B: switch routing features off (_native.FEATURES)
        if f:
            _native.FEATURES[f] = False
    import comfy.sampler_helpers, comfy.samplers, comfy.utils
    comfy.utils.PROGRESS_BAR_ENABLED = False
    dev = torch.device("cuda")
    fam = uc.FAMILIES[family]
    patcher, model = models.build_model(family, dev, seed=1)
    g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
    ctx = fam["unet"]["context_dim"]
    pos = [[g((1, 77, ctx), 1), {"pooled_output": g((1, 1280), 3)}]]
    neg = [[g((1, 77, ctx), 2), {"pooled_output": g((1, 1280), 4)}]]
    sched_steps = 30
    sig = comfy.samplers.calculate_sigmas(model.model_sampling, sched, sched_steps).cpu()
    out = []
    for b in batches:
        noise = g((b, 4, hw, hw), 5)
        row = {"per_gpu_batch": b}
        for mode in (False, True):
            p2 = patcher.clone()
            p2.model_options["cuda_graph"] = mode
            guider = comfy.samplers.CFGGuider(p2)
            guider.set_conds(pos, neg)
            guider.set_cfg(7.0)
            guider.conds = {k: [c.copy() for c in v] for k, v in guider.original_conds.items()}
            guider.inner_model, guider.conds, guider.loaded_models = comfy.sampler_helpers.prepare_sampling(
                p2, noise.shape, guider.conds)
            guider.conds = comfy.samplers.process_conds(guider.inner_model, noise.to(dev), guider.conds, dev,
                                                        torch.zeros_like(noise).to(dev), None, 1)
            model_k = comfy.samplers.KSamplerX0Inpaint(guider, sig.to(dev))
            x = (noise * float(torch.sqrt(1.0 + sig[0] ** 2.0))).to(dev).contiguous()
            fused = FusedDenoise.create(model_k, x, {"model_options": p2.model_options, "seed": 1}, sig)
            state = {"x": x, "old": None}

            def step(i):
                i = i % sched_steps
                if sampler == "euler":
                    res = fused.step(state["x"], i, sig[i], 1, float(sig[i]), float(sig[i + 1] - sig[i]), 0.0, 0.0, None)
                else:
                    m, cx, cd, cc, co = ks.dpmpp_2m_coefficients(sig, i, state["old"] is not None and i > 0)
                    res = fused.step(state["x"], i, sig[i], m, cx, cd, cc, co, state["old"] if m == 3 else None)
                state["old"], xn = res
                state["x"] = x if i == sched_steps - 1 else xn
            with torch.inference_mode():
                for w in range(4):
                    step(w)
                torch.cuda.synchronize()
                if profile:                       # ncu --profile-from-start off: one step (graph replay nodes)
                    if mode:
                        torch.cuda.profiler.start()
                        step(4)
                        torch.cuda.synchronize()
                        torch.cuda.profiler.stop()
                    continue
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0 = time.perf_counter()
                e0.record()
                for k in range(steps):
                    step(4 + k)
                e1.record()
                host_ms = (time.perf_counter() - t0) * 1e3 / steps
                torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            row["graph_ms" if mode else "eager_ms"] = ms
            row["graph_host_ms" if mode else "eager_host_ms"] = host_ms
            if mode:
                row["replayed"] = bool(fused.graph)
        if profile:
            return
        row["images_per_s_graph"] = b / (sched_steps * row["graph_ms"] / 1e3)
        print(json.dumps(row), flush=True)
        out.append(row)
    if out and out[0]["per_gpu_batch"] == 8:
        for r in out:
            r["strong_scaling_efficiency_vs_b8"] = (out[0]["graph_ms"] / 8 * r["per_gpu_batch"]) / r["graph_ms"]
        print(json.dumps([{k: r[k] for k in ("per_gpu_batch", "strong_scaling_efficiency_vs_b8")} for r in out]))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"small_batch_{family}.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
