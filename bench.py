"""bench.py -- the driver's measurement contract for the per-step UNet denoise hot path.

    python bench.py --gpus N --steps K --warmup W                       (this repo's CUDA path, as a drop-in)
    python bench.py --impl reference --gpus N --steps K --warmup W      (the reference's own CPU path, host cores)

Workload (config.workload): BASELINE.json configs[2], the configuration the metric is quoted on -- SDXL base,
1024x1024 (latents 4x128x128), GLOBAL batch 8, DPM++ 2M / Karras, CFG 7.0, bf16 UNet, random-init weights, synthetic
prompt embeddings -- on 1/2/4/8 GPUs: 8/4/2/1 images per GPU ("scaling": "strong"), weights replicated, batch sharded,
no data-path collective (one all-gather of the final latents in the e2e leg).

What runs is the REFERENCE's own model and sampler objects (the staged tree under baseline/_ref: comfy.model_base,
comfy.samplers, comfy.sample) with integration.install() active -- the path a user of the reference gets.

A "step" is one denoise step of the batch: UNet on the [uncond x b, cond x b] batch, eps -> denoised, CFG combine,
DPM++ 2M update. `value` = images/s = global batch / (30 steps x step time), device-timed with CUDA events, inputs
resident in HBM, max over ranks. `e2e` is the same metric through the public API (comfy.sample.sample, or
placement.sample_distributed at N > 1) from pinned HOST buffers: noise, latent and prompt embeddings are copied
host->device and the final latents device->host inside the timed region. `gpu_library_baseline` (N = 1) times the
UNMODIFIED reference (install() not active: SDPA / cuBLASLt / cuDNN / ATen) on the same GPU, model and inputs.
`parity_vs_1gpu` (N > 1) compares the gathered N-GPU latents with a 1-GPU run of the same global batch.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "SDXL 1024x1024 30-step DPM++ 2M Karras CFG 7.0 images/sec"
UNIT = "images/s"
SCHEDULE_STEPS = 30
FAMILY = "sdxl"
GLOBAL_BATCH = 8
LATENT_HW = 128
CFG_SCALE = 7.0
SEED = 1234


def load_traffic(kernel_substr):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/), if any."""
    for name in ("r02_ncu_top_kernels.json", "r01_final_ncu_top_kernels.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                for k in json.load(f)["kernels"]:
                    if kernel_substr in k["name"]:
                        return k
        except (OSError, ValueError, KeyError):
            pass
    return None


def load_peaks():
    peaks = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        peaks.update({k: m[k] for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in m})
        peaks["source"] = "measured"
    except (OSError, ValueError):
        pass
    return peaks


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            parts = [p.strip() for p in l.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------------------------
# synthetic inputs
# ----------------------------------------------------------------------------------------------
def synthetic_inputs(batch, ctx_dim):
    import torch
    g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
    latent = torch.zeros(batch, 4, LATENT_HW, LATENT_HW)
    noise = torch.randn(latent.shape, generator=torch.manual_seed(SEED))
    pos = [[g((1, 77, ctx_dim), 1), {"pooled_output": g((1, 1280), 3)}]]
    neg = [[g((1, 77, ctx_dim), 2), {"pooled_output": g((1, 1280), 4)}]]
    return latent, noise, pos, neg


def workload_config(n_gpus):
    per = GLOBAL_BATCH // n_gpus
    return {"workload": "BASELINE.json configs[2]: SDXL base 1024x1024, global batch 8, 30-step DPM++ 2M Karras, "
                        "CFG 7.0 (UNet batch = uncond+cond), bf16, random-init weights, synthetic prompt embeddings; "
                        "the reference's own model / sampler objects with integration.install() active",
            "per_gpu_batch": per, "global_batch": GLOBAL_BATCH, "latent": [4, LATENT_HW, LATENT_HW],
            "schedule_steps": SCHEDULE_STEPS, "sampler": "dpmpp_2m", "scheduler": "karras", "cfg": CFG_SCALE,
            "parallelism": f"batch-sharded x{n_gpus} ({per} image(s) per GPU), weights replicated",
            "l2": "activations per step (>0.1 GB per image) exceed the 126 MB L2; no explicit flush between steps"}


# ----------------------------------------------------------------------------------------------
# --impl reference: the reference's own CPU implementation of the path, on the host cores
# ----------------------------------------------------------------------------------------------
class _Budget(Exception):
    pass


def reference_cpu_seconds_per_step(max_steps, budget_s):
    """Time denoise steps of ONE 1024^2 image (UNet batch 2 = uncond + cond) through the reference's own code in its
    --cpu mode (fp32, all host threads): comfy.sample.sample on the 30-step schedule, stopped after `max_steps` steps
    or when the time budget is used up. Returns (seconds per step, steps timed, threads, kind)."""
    import torch
    from comfy_gen_server_b200 import unet_config as uc
    from comfy_gen_server_b200.reference import ensure_importable
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    ensure_importable(cpu=True)
    import comfy.sample
    import comfy.utils
    from comfy_gen_server_b200 import models
    comfy.utils.PROGRESS_BAR_ENABLED = False
    fam = uc.FAMILIES[FAMILY]
    patcher, _ = models.build_model(FAMILY, "cpu", seed=SEED, dtype=torch.float32)
    latent, noise, pos, neg = synthetic_inputs(1, fam["unet"]["context_dim"])
    stamps = [None]

    def cb(i, denoised, x, total):
        stamps.append(time.time())
        if len(stamps) - 1 >= max_steps or stamps[-1] - stamps[0] > budget_s:
            raise _Budget()
    try:
        with torch.inference_mode():
            stamps[0] = time.time()
            comfy.sample.sample(patcher, noise, SCHEDULE_STEPS, CFG_SCALE, "dpmpp_2m", "karras", pos, neg, latent,
                                seed=SEED, disable_pbar=True, callback=cb)
    except _Budget:
        pass
    done = len(stamps) - 1
    return (stamps[-1] - stamps[0]) / done, done, threads, "reference"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    budget = float(os.environ.get("CGS_REF_MAX_SECONDS", "240"))
    sec, timed, threads, kind = reference_cpu_seconds_per_step(args.steps + min(args.warmup, 1), budget)
    value = 1.0 / (SCHEDULE_STEPS * sec)
    sample = (f"the reference's own code (baseline/_ref, --cpu mode, fp32, {threads} threads): {timed} denoise step(s) "
              f"of ONE 1024^2 image (UNet batch 2: uncond + cond) of the 30-step DPM++ 2M schedule through "
              f"comfy.sample.sample, within a {budget:.0f}s budget; images/s = 1 / (30 steps x s per step)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": timed, "steps_requested": args.steps, "warmup": 0,
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------
# this repo's path
# ----------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from comfy_gen_server_b200 import _native, unet_config as uc

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: this path has no CPU fallback")
    _native.load_library()          # fail loudly if the CUDA extension is missing
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise RuntimeError(f"--gpus {args.gpus} needs torchrun (one process per GPU)")
    if GLOBAL_BATCH % world != 0:
        raise RuntimeError(f"global batch {GLOBAL_BATCH} does not divide over {world} GPUs")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    from comfy_gen_server_b200 import integration, models, placement
    from comfy_gen_server_b200.guider import FusedDenoise
    from comfy_gen_server_b200 import sampling as ks
    from comfy_gen_server_b200.reference import ensure_importable
    ensure_importable()
    import comfy.sample
    import comfy.sampler_helpers
    import comfy.samplers
    import comfy.utils
    comfy.utils.PROGRESS_BAR_ENABLED = False

    fam = uc.FAMILIES[FAMILY]
    patcher, model = models.build_model(FAMILY, dev, seed=SEED)
    per = GLOBAL_BATCH // world
    latent, noise, pos, neg = synthetic_inputs(GLOBAL_BATCH, fam["unet"]["context_dim"])
    my = slice(rank * per, (rank + 1) * per)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- the unmodified reference on this GPU (before install() touches anything) ---------------
    lib_base = None
    if world == 1 and not args.no_library_baseline:
        n = 4
        with torch.inference_mode():
            run = lambda k: comfy.sample.sample(patcher, noise, SCHEDULE_STEPS, CFG_SCALE, "dpmpp_2m", "karras", pos,
                                                neg, latent, seed=SEED, disable_pbar=True, last_step=k,
                                                force_full_denoise=False)
            run(2)                                                        # warm-up (cuDNN / cuBLASLt plan selection)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            run(n)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        lib_base = {"value": GLOBAL_BATCH / (SCHEDULE_STEPS * ms / 1e3), "unit": UNIT, "ms_per_step": ms,
                    "what": f"the unmodified reference (baseline/_ref, install() not active) on this GPU: bf16 UNet, "
                            f"attention_pytorch -> SDPA, cuBLASLt, cuDNN, ATen; comfy.sample.sample, {n} steps of the "
                            f"same schedule, batch and inputs, CUDA events around the call (incl. its per-step syncs)"}

    integration.install()
    if args.gemm_set:
        for code in args.gemm_set.split(","):
            assert _native.load_library().cgs_gemm_set_variant(int(code)) == 0, code
    if args.native is not None:         # A/B: which of the block's projections run on the library GEMM
        for f in ("qkv", "to_out", "ff_out", "conv", "skip", "upconv"):
            _native.FEATURES["native_" + f] = f in args.native.split(",")

    # ---- device-resident leg: K denoise steps through the fused step of the drop-in --------------
    guider = comfy.samplers.CFGGuider(patcher)
    guider.set_conds(pos, neg)
    guider.set_cfg(CFG_SCALE)
    guider.conds = {k: [c.copy() for c in v] for k, v in guider.original_conds.items()}
    guider.inner_model, guider.conds, guider.loaded_models = comfy.sampler_helpers.prepare_sampling(
        patcher, noise[my].shape, guider.conds)
    guider.conds = comfy.samplers.process_conds(guider.inner_model, noise[my].to(dev), guider.conds, dev,
                                                latent[my].to(dev), None, SEED)
    sig = comfy.samplers.calculate_sigmas(model.model_sampling, "karras", SCHEDULE_STEPS).cpu()
    model_k = comfy.samplers.KSamplerX0Inpaint(guider, sig.to(dev))
    x = (noise[my] * float(torch.sqrt(1.0 + sig[0] ** 2.0))).to(dev).contiguous()
    # The timed region runs the step the way sample() runs it: replayed as one CUDA graph once the probe finds the
    # host's enqueue time within reach of the GPU's (it does at every batch size here: ~1070 launches per step). A
    # replay has no per-kernel events, so the per-kernel rooflines come from separate launch-by-launch passes below.
    time_in_region = False
    if time_in_region:
        _native.time_family("geglu_gemm")       # (also keeps the step out of a CUDA graph: a replay has no per-kernel events)
    if args.profile:
        patcher.model_options["cuda_graph"] = False       # launch by launch under the profiler
    fused = FusedDenoise.create(model_k, x, {"model_options": patcher.model_options, "seed": SEED}, sig)
    if fused is None:
        raise RuntimeError("bench: the fused fast path was refused for the benchmark workload")
    state = {"x": x, "old": None}

    def denoise_step(i, f=fused):
        i = i % SCHEDULE_STEPS
        mode, c_x, c_d, c_cur, c_old = ks.dpmpp_2m_coefficients(sig, i, state["old"] is not None and i > 0)
        state["old"], xn = f.step(state["x"], i, sig[i], mode, c_x, c_d, c_cur, c_old,
                                  state["old"] if mode == 3 else None)
        # keep the synthetic trajectory bounded across schedule wrap-arounds: restart from x at i == last
        state["x"] = x if i == SCHEDULE_STEPS - 1 else xn

    with torch.inference_mode():
        for w in range(3):                       # set-up: one-time costs, the host/GPU probe, step-graph capture
            denoise_step(w)
        for w in range(args.warmup):
            denoise_step(3 + w)
        barrier()
        _native.reset_launch_counters()
        if time_in_region:
            _native.time_family("geglu_gemm")
        clocks = ClockSampler(local_rank)
        if rank == 0:
            clocks.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if args.profile:
            _native.time_family(None)
            torch.cuda.profiler.start()          # ncu --profile-from-start off captures only the timed steps
        ev0.record()
        for k in range(args.steps):
            denoise_step(3 + args.warmup + k)
        ev1.record()
        barrier()
        if args.profile:
            torch.cuda.profiler.stop()
            if rank == 0:
                print(json.dumps({"profile_run": True, "ms_per_step_under_profiler": ev0.elapsed_time(ev1) / args.steps}))
            return
        clock_info = clocks.stop() if rank == 0 else None
        elapsed_ms = ev0.elapsed_time(ev1)
        launches = dict(_native.LAUNCHES)
        replayed = bool(fused.graph)

        # per-kernel passes (launch by launch, one family per pass keeps the event overhead negligible)
        def family_pass(name, steps):
            p2 = patcher.clone()
            p2.model_options["cuda_graph"] = False
            f2 = FusedDenoise.create(model_k, x, {"model_options": p2.model_options, "seed": SEED}, sig)
            _native.reset_launch_counters()
            _native.time_family(name)
            state.update(x=x, old=None)
            for k in range(steps):
                denoise_step(k, f2)
            torch.cuda.synchronize()
            out = _native.timed_family_totals()
            _native.time_family(None)
            return out
        if time_in_region:
            geglu_ms, geglu_flops, geglu_n = _native.timed_family_totals()
            geglu_steps = args.steps
            _native.time_family(None)
        else:
            geglu_steps = min(args.steps, 6)
            geglu_ms, geglu_flops, geglu_n = family_pass("geglu_gemm", geglu_steps)
        attn_steps = min(args.steps, 6)
        attn_ms, attn_flops, attn_n = family_pass("attention", attn_steps)
        xattn_ms, xattn_bytes, xattn_n = family_pass("attention_cross", min(args.steps, 3))
        ln_ms, ln_bytes, ln_n = family_pass("add_layernorm", min(args.steps, 3))
        pln_ms, pln_bytes, pln_n = family_pass("layernorm", min(args.steps, 3))
        lin_ms, lin_flops, lin_n = family_pass("linear", min(args.steps, 3))
        conv_ms, conv_flops, conv_n = family_pass("conv3x3", min(args.steps, 3))

        # ---- end-to-end leg: public API, pinned host buffers -> final latents on the host -------
        noise_h, latent_h = noise.pin_memory(), latent.pin_memory()
        pos_h = [[pos[0][0].pin_memory(), {"pooled_output": pos[0][1]["pooled_output"].pin_memory()}]]
        neg_h = [[neg[0][0].pin_memory(), {"pooled_output": neg[0][1]["pooled_output"].pin_memory()}]]
        e2e_steps = args.steps

        def e2e_call():
            if world > 1:
                return placement.sample_distributed(patcher, noise_h, e2e_steps, CFG_SCALE, "dpmpp_2m", "karras",
                                                    pos_h, neg_h, latent_h, seed=SEED)
            return comfy.sample.sample(patcher, noise_h, e2e_steps, CFG_SCALE, "dpmpp_2m", "karras", pos_h, neg_h,
                                       latent_h, seed=SEED, disable_pbar=True)
        e2e_call()                               # warm-up call (a server's steady state: step graph already captured)
        barrier()
        t0 = time.perf_counter()
        out = e2e_call()
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        assert out.device.type == "cpu" and out.shape[0] == GLOBAL_BATCH and bool(torch.isfinite(out).all())

        # ---- N-GPU latents == 1-GPU latents -------------------------------------------------------
        parity = None
        if world > 1:
            single = comfy.sample.sample(patcher, noise_h, e2e_steps, CFG_SCALE, "dpmpp_2m", "karras", pos_h, neg_h,
                                         latent_h, seed=SEED, disable_pbar=True)
            a, b = out.double().flatten(), single.double().flatten()
            parity = {"max_abs": float((out - single).abs().max()), "cosine": float((a @ b) / (a.norm() * b.norm())),
                      "ref_abs_max": float(single.abs().max()), "bit_equal": bool(torch.equal(out, single)),
                      "what": f"gathered latents of sample_distributed on {world} GPUs vs comfy.sample.sample of the same "
                              f"global batch on one GPU (rank 0), {e2e_steps} steps"}

    # max over ranks
    if world > 1:
        t = torch.tensor([elapsed_ms, e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms, e2e_s = float(t[0]), float(t[1])

    if rank == 0:
        ms_per_step = elapsed_ms / args.steps
        value = GLOBAL_BATCH / (SCHEDULE_STEPS * ms_per_step / 1e3)
        e2e_value = GLOBAL_BATCH / (e2e_s * SCHEDULE_STEPS / e2e_steps)
        peaks = load_peaks()
        h2d = (noise_h[my].numel() + latent_h[my].numel()) * 4 + 2 * (77 * fam["unet"]["context_dim"] + 1280) * 4
        d2h = per * 4 * LATENT_HW * LATENT_HW * 4
        tf = geglu_flops / (geglu_ms * 1e-3) / 1e12 if geglu_ms > 0 else 0.0
        atf = attn_flops / (attn_ms * 1e-3) / 1e12 if attn_ms > 0 else 0.0
        lnbw = ln_bytes / (ln_ms * 1e-3) / 1e9 if ln_ms > 0 else 0.0
        plnbw = pln_bytes / (pln_ms * 1e-3) / 1e9 if pln_ms > 0 else 0.0
        peak_tf = peaks["bf16_tflops_sustained"]
        ncu_gemm = load_traffic("gemm_kernel<1")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(world),
            "clocks": clock_info,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // e2e_steps,
                    "d2h_bytes_per_step": d2h // e2e_steps,
                    "note": f"one public-API call of {e2e_steps} steps from pinned host buffers (after one warm-up "
                            f"call); bytes are per call / steps, per rank"},
            "gpu_launches": int(sum(launches.values())), "gpu_launches_by_kernel": launches,
            "step_graph_replay": replayed,
            "roofline": {"kernel": "gemm_kernel<EPI_GEGLU> (cgs_geglu_gemm_fwd)", "bound": "tensor",
                         "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf if peak_tf else None,
                         "traffic": ncu_gemm["dram_bytes_per_launch"] if ncu_gemm else None,
                         "traffic_note": (f"ncu --set full, {ncu_gemm['shape']}: algorithmic "
                                          f"{ncu_gemm['algorithmic_bytes']} B") if ncu_gemm else None,
                         "launches_timed": geglu_n, "ms_in_kernel_per_step": geglu_ms / geglu_steps,
                         "timed": "inside the timed region" if time_in_region else "separate launch-by-launch pass",
                         "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({peaks['source']})"},
            "roofline_attention": {"kernel": "attn_fwd_kernel (cgs_attention_fwd, self-attention: 4096 / 1024 tokens, d = 64)", "bound": "tensor",
                                   "achieved": atf, "peak": peak_tf, "unit": "TFLOP/s",
                                   "frac": atf / peak_tf if peak_tf else None, "launches_timed": attn_n,
                                   "ms_in_kernel_per_step": attn_ms / attn_steps},
            "roofline_attention_cross": {"kernel": "attn_fwd_kernel<.., 96> (cgs_attention_fwd, text cross-attention: 77 keys)",
                                         "bound": "hbm",
                                         "achieved": xattn_bytes / (xattn_ms * 1e-3) / 1e9 if xattn_ms > 0 else 0.0,
                                         "peak": peaks["hbm_gbs"], "unit": "GB/s", "launches_timed": xattn_n,
                                         "frac": (xattn_bytes / (xattn_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if xattn_ms > 0 else 0.0,
                                         "ms_in_kernel_per_step": xattn_ms / min(args.steps, 3),
                                         "note": "algorithmic bytes = Q + O + K + V once"},
            "roofline_add_layernorm": {"kernel": "layernorm_kernel<ADD> (cgs_add_layernorm_fwd)", "bound": "hbm",
                                       "achieved": lnbw, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                       "frac": lnbw / peaks["hbm_gbs"], "launches_timed": ln_n,
                                       "ms_in_kernel_per_step": ln_ms / min(args.steps, 3)},
            "roofline_linear": {"kernel": "gemm_kernel<EPI_LINEAR> (cgs_linear_fwd: fused QKV, to_out / ff.net.2 + residual)",
                                "bound": "tensor", "achieved": lin_flops / (lin_ms * 1e-3) / 1e12 if lin_ms > 0 else 0.0,
                                "peak": peak_tf, "unit": "TFLOP/s", "launches_timed": lin_n,
                                "frac": (lin_flops / (lin_ms * 1e-3) / 1e12 / peak_tf) if lin_ms > 0 else 0.0,
                                "ms_in_kernel_per_step": lin_ms / min(args.steps, 3)},
            "roofline_conv": {"kernel": "gemm_kernel<EPI_LINEAR, PAIR, CONV> (cgs_conv3x3_nhwc_fwd, "
                                        "cgs_upsample2x_conv3x3_nhwc_fwd: ResBlock / Upsample convolutions)",
                              "bound": "tensor", "achieved": conv_flops / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0,
                              "peak": peak_tf, "unit": "TFLOP/s", "launches_timed": conv_n,
                              "frac": (conv_flops / (conv_ms * 1e-3) / 1e12 / peak_tf) if conv_ms > 0 else 0.0,
                              "ms_in_kernel_per_step": conv_ms / min(args.steps, 3),
                              "note": "algorithmic FLOPs of the kernel as run: 2 x pixels x 9 Cin x Cout, and "
                                      "2 x small-image pixels x 16 Cin x Cout for interpolate + conv"},
            "roofline_layernorm": {"kernel": "layernorm_stream_kernel (cgs_layernorm_fwd)", "bound": "hbm",
                                   "achieved": plnbw, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                   "frac": plnbw / peaks["hbm_gbs"], "launches_timed": pln_n,
                                   "ms_in_kernel_per_step": pln_ms / min(args.steps, 3)},
        }
        if lib_base is not None:
            line["gpu_library_baseline"] = lib_base
        if parity is not None:
            line["parity_vs_1gpu"] = parity
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_subprocess()
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline_subprocess():
    """The reference's CPU path in a process of its own (its --cpu mode is a process-wide switch of the reference):
    one bounded sample, see run_reference."""
    env = dict(os.environ, CGS_REF_MAX_SECONDS="60", CUDA_VISIBLE_DEVICES="")
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        env.pop(k, None)
    try:
        p = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "1",
                            "--warmup", "0"], capture_output=True, text=True, timeout=600, env=env)
        for l in reversed(p.stdout.splitlines()):
            if l.startswith("{"):
                return json.loads(l)["cpu_baseline"]
        return {"value": None, "unit": UNIT, "error": (p.stderr or p.stdout)[-400:]}
    except (subprocess.TimeoutExpired, OSError) as e:
        return {"value": None, "unit": UNIT, "error": str(e)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true")
    ap.add_argument("--native", default=None, help="A/B: comma list of qkv,to_out,ff_out,conv,skip on the library GEMM")
    ap.add_argument("--gemm-set", default=None, help="A/B: comma list of cgs_gemm_set_variant codes (include/cgs_b200.h)")
    ap.add_argument("--profile", action="store_true",
                    help="only the timed denoise steps, bracketed by cudaProfilerStart/Stop (for ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
