"""bench.py -- the driver's measurement contract for the per-step UNet denoise hot path.

    python bench.py --gpus N --steps K --warmup W              (this repo's CUDA path)
    python bench.py --impl reference --gpus N --steps K --warmup W   (CPU oracle port of the reference)

Workload (config.workload): BASELINE.json configs[2], the configuration the metric is quoted on and the
largest that is meant for one GPU -- SDXL base, 1024x1024 (latents 4x128x128), batch 8 per GPU,
DPM++ 2M / Karras, CFG 7.0, bf16 UNet, random-init weights, synthetic prompt embeddings.

A "step" is one denoise step of that batch: UNet on the [uncond x 8, cond x 8] batch, eps -> denoised,
CFG combine, DPM++ 2M update. `value` = images/s = global batch / (30 steps x mean step time), device-timed
with CUDA events, inputs resident in HBM. `e2e` is the same metric through the public API
(comfy_gen_server_b200.sample.sample / sample_distributed) from pinned HOST buffers: noise, latent and
prompt embeddings are copied host->device and the final latents device->host inside the timed region.
N > 1: one process per GPU (torchrun), weights replicated, batch sharded, scaling "weak" (8 images per GPU);
the only collective is the all-gather of final latents in the e2e leg.
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
PER_GPU_BATCH = 8
LATENT_HW = 128
CFG_SCALE = 7.0
SEED = 1234


def load_traffic(kernel_substr):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (profiles/), if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "r01_final_ncu_top_kernels.json")) as f:
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


# ----------------------------------------------------------------------------------------------
# CPU baseline (the oracle port of the reference, timed on the host cores)
# ----------------------------------------------------------------------------------------------
def cpu_unet_eval_seconds(n_evals, warm=0, max_seconds=None):
    """Time UNet evaluations of ONE 1024^2 latent (cond half only) with the oracle, fp32, all host threads.
    Returns (seconds per evaluation, evaluations timed, threads)."""
    import torch
    from comfy_gen_server_b200 import unet_config as uc
    from oracle import oracle as O
    fam = uc.FAMILIES[FAMILY]
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    shapes = uc.param_shapes(fam["unet"])
    sd = uc.seeded_state_dict(shapes, SEED)
    g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
    x = g((1, 4, LATENT_HW, LATENT_HW), 5)
    ctx = g((1, 77, fam["unet"]["context_dim"]), 1)
    y = O.sdxl_adm(g((1, 1280), 3), LATENT_HW * 8, LATENT_HW * 8)
    table, logt = O.discrete_sigmas(), O.discrete_log_sigmas()
    sig = torch.tensor([float(table[-1])])

    def one():
        with torch.inference_mode():
            xc = O.calculate_input(sig, x)
            out = O.unet_forward(fam["unet"], sd, xc, O.timestep_of_sigma(sig, logt).float(), ctx, y)
            return O.calculate_denoised(sig, out, x, fam["prediction"])
    for _ in range(warm):
        one()
    t0 = time.time()
    done = 0
    for _ in range(n_evals):
        one()
        done += 1
        if max_seconds is not None and time.time() - t0 > max_seconds:
            break
    return (time.time() - t0) / done, done, threads


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port; the reference is pure
    Python and cannot travel to the GPU box, and it has no compiled code to build into oracle/_ref)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    budget = float(os.environ.get("CGS_REF_MAX_SECONDS", "240"))
    sec, timed, threads = cpu_unet_eval_seconds(args.steps, warm=min(args.warmup, 1), max_seconds=budget)
    # one denoise step of one image = 2 UNet evaluations (cond + uncond); an image = 30 steps
    value = 1.0 / (SCHEDULE_STEPS * 2 * sec)
    sample = (f"each step = ONE fp32 UNet evaluation of one 1024^2 latent (cond half), oracle port, "
              f"{timed} of {args.steps} requested steps timed within a {budget:.0f}s budget; "
              f"images/s = 1 / (30 steps x 2 CFG halves x s per evaluation)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": timed, "steps_requested": args.steps, "warmup": min(args.warmup, 1),
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(n_gpus):
    return {"workload": "BASELINE.json configs[2]: SDXL base 1024x1024, batch 8 per GPU, 30-step DPM++ 2M Karras, "
                        "CFG 7.0 (UNet batch 16 = uncond+cond), bf16, random-init weights, synthetic prompt embeddings",
            "per_gpu_batch": PER_GPU_BATCH, "global_batch": PER_GPU_BATCH * n_gpus, "latent": [4, LATENT_HW, LATENT_HW],
            "schedule_steps": SCHEDULE_STEPS, "sampler": "dpmpp_2m", "scheduler": "karras", "cfg": CFG_SCALE,
            "parallelism": f"batch-sharded x{n_gpus}, weights replicated",
            "l2": "activations per step (>1 GB) exceed the 126 MB L2; no explicit flush between steps"}


# ----------------------------------------------------------------------------------------------
# this repo's path
# ----------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from comfy_gen_server_b200 import _native, unet_config as uc
    from comfy_gen_server_b200 import sample as S, samplers
    from comfy_gen_server_b200.k_diffusion import sampling as ks
    from comfy_gen_server_b200.model_base import build_model
    from comfy_gen_server_b200.model_management import ModelPatcher

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: this path has no CPU fallback")
    _native.load_library()          # fail loudly if the CUDA extension is missing
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise RuntimeError(f"--gpus {args.gpus} needs torchrun (one process per GPU)")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    fam = uc.FAMILIES[FAMILY]
    model = build_model(FAMILY, dtype=torch.bfloat16, device=dev, seed=SEED)
    patcher = ModelPatcher(model, load_device=dev)
    global_batch = PER_GPU_BATCH * world
    latent, noise, pos, neg = synthetic_inputs(global_batch, fam["unet"]["context_dim"])
    my = slice(rank * PER_GPU_BATCH, (rank + 1) * PER_GPU_BATCH)

    # ---- device-resident leg: K denoise steps through the guider's fused step ------------------
    guider = samplers.CFGGuider(patcher)
    guider.set_conds(pos, neg)
    guider.set_cfg(CFG_SCALE)
    guider.conds = {k: [c.copy() for c in v] for k, v in guider.original_conds.items()}
    guider.inner_model = model
    guider.conds = samplers.process_conds(model, noise[my].to(dev), guider.conds, dev)
    sig = samplers.calculate_sigmas(model.model_sampling, "karras", SCHEDULE_STEPS)
    guider.prepare_schedule(sig)
    x = (noise[my] * float(torch.sqrt(1.0 + sig[0] ** 2.0))).to(dev).contiguous()
    state = {"x": x, "old": None}

    def denoise_step(i):
        i = i % SCHEDULE_STEPS
        mode, c_x, c_d, c_cur, c_old = ks.dpmpp_2m_coefficients(sig, i, state["old"] is not None and i > 0)
        res = guider.fused_step(state["x"], i, sig[i], mode, c_x, c_d, c_cur, c_old,
                                state["old"] if mode == 3 else None, model_options=patcher.model_options)
        if res is None:
            raise RuntimeError("bench: the fused fast path was refused for the benchmark workload")
        state["old"], xn = res
        # keep the synthetic trajectory bounded across schedule wrap-arounds: restart from x at i == last
        state["x"] = x if i == SCHEDULE_STEPS - 1 else xn

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.inference_mode():
        # the GEGLU launches are bracketed with CUDA events from the first step on (which also keeps the guider from
        # replaying the step as a CUDA graph: a replay has no per-kernel events; at this batch size its "auto" probe
        # declines anyway, the GPU being the limiter)
        _native.time_family("geglu_gemm")
        for w in range(args.warmup):
            denoise_step(w)
        barrier()
        _native.reset_launch_counters()
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
            denoise_step(args.warmup + k)
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
        geglu_ms, geglu_flops, geglu_n = _native.timed_family_totals()
        _native.time_family(None)

        # second pass, timing the attention family (one family per pass keeps the event overhead negligible)
        _native.reset_launch_counters()
        _native.time_family("attention")
        for k in range(min(args.steps, 6)):
            denoise_step(k)
        torch.cuda.synchronize()
        attn_ms, attn_flops, attn_n = _native.timed_family_totals()
        _native.time_family(None)

        # ---- end-to-end leg: public API, pinned host buffers -> final latents on the host -------
        noise_h, latent_h = noise.pin_memory(), latent.pin_memory()
        pos_h = [[pos[0][0].pin_memory(), {"pooled_output": pos[0][1]["pooled_output"].pin_memory()}]]
        neg_h = [[neg[0][0].pin_memory(), {"pooled_output": neg[0][1]["pooled_output"].pin_memory()}]]
        e2e_steps = args.steps

        def e2e_call():
            if world > 1:
                return S.sample_distributed(patcher, noise_h, e2e_steps, CFG_SCALE, "dpmpp_2m", "karras", pos_h, neg_h,
                                            latent_h, seed=SEED)
            return S.sample(patcher, noise_h, e2e_steps, CFG_SCALE, "dpmpp_2m", "karras", pos_h, neg_h, latent_h,
                            seed=SEED, disable_pbar=True)
        barrier()
        t0 = time.perf_counter()
        out = e2e_call()
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        assert out.device.type == "cpu" and out.shape[0] == global_batch and bool(torch.isfinite(out).all())

    # max over ranks
    if world > 1:
        t = torch.tensor([elapsed_ms, e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_ms, e2e_s = float(t[0]), float(t[1])

    if rank == 0:
        ms_per_step = elapsed_ms / args.steps
        value = global_batch / (SCHEDULE_STEPS * ms_per_step / 1e3)
        e2e_value = global_batch / (e2e_s * SCHEDULE_STEPS / e2e_steps)
        peaks = load_peaks()
        per_rank_b = PER_GPU_BATCH
        h2d = (noise_h[my].numel() + latent_h[my].numel()) * 4 + 2 * (77 * fam["unet"]["context_dim"] + 1280) * 4
        d2h = per_rank_b * 4 * LATENT_HW * LATENT_HW * 4
        tf = geglu_flops / (geglu_ms * 1e-3) / 1e12 if geglu_ms > 0 else 0.0
        atf = attn_flops / (attn_ms * 1e-3) / 1e12 if attn_ms > 0 else 0.0
        peak_tf = peaks["bf16_tflops_sustained"]
        ncu_gemm = load_traffic("gemm_kernel")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": workload_config(world),
            "clocks": clock_info,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // e2e_steps,
                    "d2h_bytes_per_step": d2h // e2e_steps,
                    "note": f"one sample() call of {e2e_steps} steps from pinned host buffers; bytes are per call / steps"},
            "gpu_launches": int(sum(launches.values())), "gpu_launches_by_kernel": launches,
            "roofline": {"kernel": "gemm_kernel<EPI_GEGLU> (cgs_geglu_gemm_fwd)", "bound": "tensor",
                         "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf if peak_tf else None,
                         "traffic": ncu_gemm["dram_bytes_per_launch"] if ncu_gemm else None,
                         "traffic_note": (f"ncu --set full, {ncu_gemm['shape']}: algorithmic "
                                          f"{ncu_gemm['algorithmic_bytes']} B") if ncu_gemm else None,
                         "launches_timed": geglu_n, "ms_in_kernel_per_step": geglu_ms / args.steps,
                         "peak_source": f"MEASURED_PEAKS.json bf16_tflops_sustained ({peaks['source']})"},
            "roofline_attention": {"kernel": "attn_fwd_kernel (cgs_attention_fwd, self + cross)", "bound": "tensor",
                                   "achieved": atf, "peak": peak_tf, "unit": "TFLOP/s",
                                   "frac": atf / peak_tf if peak_tf else None, "launches_timed": attn_n,
                                   "ms_in_kernel_per_step": attn_ms / max(1, min(args.steps, 6))},
        }
        if world == 1 and not args.no_cpu_baseline:
            sec, timed, threads = cpu_unet_eval_seconds(2, warm=0, max_seconds=45)
            line["cpu_baseline"] = {
                "value": 1.0 / (SCHEDULE_STEPS * 2 * sec), "unit": UNIT, "cores": threads, "kind": "port",
                "sample": f"{timed} fp32 UNet evaluation(s) of one 1024^2 latent with the oracle port "
                          f"({sec:.1f} s each); images/s = 1 / (30 steps x 2 CFG halves x s per evaluation)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile", action="store_true",
                    help="only the timed denoise steps, bracketed by cudaProfilerStart/Stop (for ncu)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
