"""Classifier-free guidance, cond batching and the sampler wrappers, with the reference's public
API (comfy/samplers.py): calc_cond_batch, cfg_function, sampling_function, CFGGuider, KSAMPLER,
KSampler, calculate_sigmas, sample -- same names, argument meaning and hook behaviour
(`sampler_cfg_function`, `sampler_post_cfg_function`, `model_function_wrapper`,
`denoise_mask_function`, `disable_cfg1_optimization`, cond areas / strength / timestep ranges).

Fast path (CFGGuider.fused_step): when every item of the predicate in SURVEY.md section 8b holds --
plain full-frame conds, no hooks -- one UNet call on the [uncond x B, cond x B] batch (built once per
schedule, text K/V cached inside the attention modules) is followed by ONE kernel that does eps/v ->
denoised, the CFG combine and the sampler update (cgs_cfg_sampler_tail_fwd). Anything else runs the
reference's composition below, which still reaches the B200 kernels through the modules.

ControlNet / GLIGEN side networks are outside this hot path and raise.
"""
from __future__ import annotations

import collections
import logging
import time
import math

import torch

from . import _native
from . import model_management
from .conds import CONDCrossAttn, CONDRegular
from .k_diffusion import sampling as k_diffusion_sampling
from .model_sampling import ModelSamplingDiscrete


# ----------------------------------------------------------------------------------------------
# cond plumbing (comfy/samplers.py:10-129)
# ----------------------------------------------------------------------------------------------
cond_obj = collections.namedtuple('cond_obj', ['input_x', 'mult', 'conditioning', 'area', 'control', 'patches'])


def get_area_and_mult(conds, x_in, timestep_in):
    area = (x_in.shape[2], x_in.shape[3], 0, 0)
    strength = 1.0
    if 'timestep_start' in conds:
        timestep_start = conds['timestep_start']
        if timestep_in[0] > timestep_start:
            return None
    if 'timestep_end' in conds:
        timestep_end = conds['timestep_end']
        if timestep_in[0] < timestep_end:
            return None
    if 'area' in conds:
        area = conds['area']
    if 'strength' in conds:
        strength = conds['strength']

    input_x = x_in[:, :, area[2]:area[0] + area[2], area[3]:area[1] + area[3]]
    if 'mask' in conds:
        mask_strength = conds.get("mask_strength", 1.0)
        mask = conds['mask']
        assert mask.shape[1] == x_in.shape[2]
        assert mask.shape[2] == x_in.shape[3]
        mask = mask[:, area[2]:area[0] + area[2], area[3]:area[1] + area[3]] * mask_strength
        mask = mask.unsqueeze(1).repeat(input_x.shape[0] // mask.shape[0], input_x.shape[1], 1, 1)
    else:
        mask = torch.ones_like(input_x)
    mult = mask * strength

    if 'mask' not in conds:
        rr = 8
        if area[2] != 0:
            for t in range(rr):
                mult[:, :, t:1 + t, :] *= ((1.0 / rr) * (t + 1))
        if (area[0] + area[2]) < x_in.shape[2]:
            for t in range(rr):
                mult[:, :, area[0] - 1 - t:area[0] - t, :] *= ((1.0 / rr) * (t + 1))
        if area[3] != 0:
            for t in range(rr):
                mult[:, :, :, t:1 + t] *= ((1.0 / rr) * (t + 1))
        if (area[1] + area[3]) < x_in.shape[3]:
            for t in range(rr):
                mult[:, :, :, area[1] - 1 - t:area[1] - t] *= ((1.0 / rr) * (t + 1))

    conditioning = {}
    model_conds = conds["model_conds"]
    for c in model_conds:
        conditioning[c] = model_conds[c].process_cond(batch_size=x_in.shape[0], device=x_in.device, area=area)

    if conds.get('control', None) is not None or 'gligen' in conds:
        raise NotImplementedError("ControlNet / GLIGEN conds are outside the B200 hot path")
    return cond_obj(input_x, mult, conditioning, area, None, None)


def cond_equal_size(c1, c2):
    if c1 is c2:
        return True
    if c1.keys() != c2.keys():
        return False
    for k in c1:
        if not c1[k].can_concat(c2[k]):
            return False
    return True


def can_concat_cond(c1, c2):
    if c1.input_x.shape != c2.input_x.shape:
        return False
    return cond_equal_size(c1.conditioning, c2.conditioning)


def cond_cat(c_list):
    temp = {}
    for x in c_list:
        for k in x:
            temp.setdefault(k, []).append(x[k])
    return {k: conds[0].concat(conds[1:]) for k, conds in temp.items()}


def calc_cond_batch(model, conds, x_in, timestep, model_options):
    """comfy/samplers.py:131-228. Returns one tensor per cond list; batches concat-compatible entries
    (order [uncond..., cond...] because the candidate list is reversed, as in the reference)."""
    out_conds = []
    out_counts = []
    to_run = []
    for i in range(len(conds)):
        out_conds.append(torch.zeros_like(x_in))
        out_counts.append(torch.ones_like(x_in) * 1e-37)
        cond = conds[i]
        if cond is not None:
            for x in cond:
                p = get_area_and_mult(x, x_in, timestep)
                if p is None:
                    continue
                to_run += [(p, i)]

    while len(to_run) > 0:
        first = to_run[0]
        first_shape = first[0][0].shape
        to_batch_temp = []
        for x in range(len(to_run)):
            if can_concat_cond(to_run[x][0], first[0]):
                to_batch_temp += [x]
        to_batch_temp.reverse()
        to_batch = to_batch_temp[:1]

        free_memory = model_management.get_free_memory(x_in.device)
        for i in range(1, len(to_batch_temp) + 1):
            batch_amount = to_batch_temp[:len(to_batch_temp) // i]
            input_shape = [len(batch_amount) * first_shape[0]] + list(first_shape)[1:]
            if model.memory_required(input_shape) < free_memory:
                to_batch = batch_amount
                break

        input_x, mult, c, cond_or_uncond, area = [], [], [], [], []
        for x in to_batch:
            o = to_run.pop(x)
            p = o[0]
            input_x.append(p.input_x)
            mult.append(p.mult)
            c.append(p.conditioning)
            area.append(p.area)
            cond_or_uncond.append(o[1])

        batch_chunks = len(cond_or_uncond)
        input_x = torch.cat(input_x)
        c = cond_cat(c)
        timestep_ = torch.cat([timestep] * batch_chunks)

        transformer_options = {}
        if 'transformer_options' in model_options:
            transformer_options = model_options['transformer_options'].copy()
        transformer_options["cond_or_uncond"] = cond_or_uncond[:]
        transformer_options["sigmas"] = timestep
        c['transformer_options'] = transformer_options

        if 'model_function_wrapper' in model_options:
            output = model_options['model_function_wrapper'](
                model.apply_model, {"input": input_x, "timestep": timestep_, "c": c,
                                    "cond_or_uncond": cond_or_uncond}).chunk(batch_chunks)
        else:
            output = model.apply_model(input_x, timestep_, **c).chunk(batch_chunks)

        for o in range(batch_chunks):
            cond_index = cond_or_uncond[o]
            a = area[o]
            out_conds[cond_index][:, :, a[2]:a[0] + a[2], a[3]:a[1] + a[3]] += output[o] * mult[o]
            out_counts[cond_index][:, :, a[2]:a[0] + a[2], a[3]:a[1] + a[3]] += mult[o]

    for i in range(len(out_conds)):
        out_conds[i] /= out_counts[i]
    return out_conds


def cfg_function(model, cond_pred, uncond_pred, cond_scale, x, timestep, model_options={}, cond=None, uncond=None):
    """comfy/samplers.py:234-247."""
    if "sampler_cfg_function" in model_options:
        args = {"cond": x - cond_pred, "uncond": x - uncond_pred, "cond_scale": cond_scale, "timestep": timestep,
                "input": x, "sigma": timestep, "cond_denoised": cond_pred, "uncond_denoised": uncond_pred,
                "model": model, "model_options": model_options}
        cfg_result = x - model_options["sampler_cfg_function"](args)
    else:
        cfg_result = uncond_pred + (cond_pred - uncond_pred) * cond_scale

    for fn in model_options.get("sampler_post_cfg_function", []):
        args = {"denoised": cfg_result, "cond": cond, "uncond": uncond, "model": model,
                "uncond_denoised": uncond_pred, "cond_denoised": cond_pred, "sigma": timestep,
                "model_options": model_options, "input": x}
        cfg_result = fn(args)
    return cfg_result


def sampling_function(model, x, timestep, uncond, cond, cond_scale, model_options={}, seed=None):
    """comfy/samplers.py:251-259. Returns denoised."""
    if math.isclose(cond_scale, 1.0) and model_options.get("disable_cfg1_optimization", False) == False:
        uncond_ = None
    else:
        uncond_ = uncond
    conds = [cond, uncond_]
    out = calc_cond_batch(model, conds, x, timestep, model_options)
    return cfg_function(model, out[0], out[1], cond_scale, x, timestep, model_options=model_options, cond=cond,
                        uncond=uncond_)


class KSamplerX0Inpaint:
    """comfy/samplers.py:262-275."""
    def __init__(self, model, sigmas):
        self.inner_model = model
        self.sigmas = sigmas

    def __call__(self, x, sigma, denoise_mask, model_options={}, seed=None):
        if denoise_mask is not None:
            if "denoise_mask_function" in model_options:
                denoise_mask = model_options["denoise_mask_function"](
                    sigma, denoise_mask, extra_options={"model": self.inner_model, "sigmas": self.sigmas})
            latent_mask = 1. - denoise_mask
            x = x * denoise_mask + self.inner_model.inner_model.model_sampling.noise_scaling(
                sigma.reshape([sigma.shape[0]] + [1] * (len(self.noise.shape) - 1)), self.noise,
                self.latent_image) * latent_mask
        out = self.inner_model(x, sigma, model_options=model_options, seed=seed)
        if denoise_mask is not None:
            out = out * denoise_mask + self.latent_image * latent_mask
        return out

    # fast-path plumbing: only a pass-through when no inpainting mask is active
    def prepare_schedule(self, sig):
        fn = getattr(self.inner_model, "prepare_schedule", None)
        if fn is not None:
            fn(sig)

    def fused_step(self, x, i, sigma_hat, sampler, c0, c1, c2, c3, old_denoised, denoise_mask=None,
                   model_options={}, seed=None):
        fn = getattr(self.inner_model, "fused_step", None)
        if denoise_mask is not None or fn is None:
            return None
        return fn(x, i, sigma_hat, sampler, c0, c1, c2, c3, old_denoised, model_options=model_options, seed=seed)


# ----------------------------------------------------------------------------------------------
# schedules (comfy/samplers.py:277-313, 663-678)
# ----------------------------------------------------------------------------------------------
def simple_scheduler(model_sampling, steps):
    s = model_sampling
    sigs = []
    ss = len(s.sigmas) / steps
    for x in range(steps):
        sigs += [float(s.sigmas[-(1 + int(x * ss))])]
    sigs += [0.0]
    return torch.FloatTensor(sigs)


def ddim_scheduler(model_sampling, steps):
    s = model_sampling
    sigs = []
    ss = max(len(s.sigmas) // steps, 1)
    x = 1
    while x < len(s.sigmas):
        sigs += [float(s.sigmas[x])]
        x += ss
    sigs = sigs[::-1]
    sigs += [0.0]
    return torch.FloatTensor(sigs)


def normal_scheduler(model_sampling, steps, sgm=False, floor=False):
    s = model_sampling
    start = s.timestep(s.sigma_max)
    end = s.timestep(s.sigma_min)
    if sgm:
        timesteps = torch.linspace(start, end, steps + 1)[:-1]
    else:
        timesteps = torch.linspace(start, end, steps)
    sigs = []
    for x in range(len(timesteps)):
        sigs.append(s.sigma(timesteps[x]))
    sigs += [0.0]
    return torch.FloatTensor(sigs)


SCHEDULER_NAMES = ["normal", "karras", "exponential", "sgm_uniform", "simple", "ddim_uniform"]
KSAMPLER_NAMES = ["euler", "dpmpp_2m"]      # the samplers of this hot path; the other 17 are out of scope
SAMPLER_NAMES = KSAMPLER_NAMES


def calculate_sigmas(model_sampling, scheduler_name, steps):
    if scheduler_name == "karras":
        return k_diffusion_sampling.get_sigmas_karras(n=steps, sigma_min=float(model_sampling.sigma_min),
                                                      sigma_max=float(model_sampling.sigma_max))
    if scheduler_name == "exponential":
        return k_diffusion_sampling.get_sigmas_exponential(n=steps, sigma_min=float(model_sampling.sigma_min),
                                                           sigma_max=float(model_sampling.sigma_max))
    if scheduler_name == "normal":
        return normal_scheduler(model_sampling, steps)
    if scheduler_name == "simple":
        return simple_scheduler(model_sampling, steps)
    if scheduler_name == "ddim_uniform":
        return ddim_scheduler(model_sampling, steps)
    if scheduler_name == "sgm_uniform":
        return normal_scheduler(model_sampling, steps, sgm=True)
    raise ValueError("invalid scheduler {}".format(scheduler_name))


# ----------------------------------------------------------------------------------------------
# cond preparation (comfy/sampler_helpers.py:15-33, comfy/samplers.py:476-590)
# ----------------------------------------------------------------------------------------------
def convert_cond(cond):
    """CONDITIONING [[tensor, dict], ...] -> list of cond dicts (sampler_helpers.py:15-27)."""
    out = []
    for c in cond:
        temp = c[1].copy()
        model_conds = temp.get("model_conds", {})
        if c[0] is not None:
            model_conds["c_crossattn"] = CONDCrossAttn(c[0])
            temp["cross_attn"] = c[0]
        temp["model_conds"] = model_conds
        out.append(temp)
    return out


def calculate_start_end_timesteps(model, conds):
    s = model.model_sampling
    for t in range(len(conds)):
        x = conds[t]
        timestep_start = timestep_end = None
        if 'start_percent' in x:
            timestep_start = s.percent_to_sigma(x['start_percent'])
        if 'end_percent' in x:
            timestep_end = s.percent_to_sigma(x['end_percent'])
        if (timestep_start is not None) or (timestep_end is not None):
            n = x.copy()
            if timestep_start is not None:
                n['timestep_start'] = timestep_start
            if timestep_end is not None:
                n['timestep_end'] = timestep_end
            conds[t] = n


def encode_model_conds(model_function, conds, noise, device, prompt_type, **kwargs):
    for t in range(len(conds)):
        x = conds[t]
        params = x.copy()
        params["device"] = device
        params["noise"] = noise
        params["width"] = params.get("width", noise.shape[3] * 8)
        params["height"] = params.get("height", noise.shape[2] * 8)
        params["prompt_type"] = params.get("prompt_type", prompt_type)
        for k in kwargs:
            if k not in params:
                params[k] = kwargs[k]
        out = model_function(**params)
        x = x.copy()
        model_conds = x['model_conds'].copy()
        for k in out:
            model_conds[k] = out[k]
        x['model_conds'] = model_conds
        conds[t] = x
    return conds


def process_conds(model, noise, conds, device, latent_image=None, denoise_mask=None, seed=None):
    for k in conds:
        conds[k] = conds[k][:]
        for c in conds[k]:
            if 'area' in c and isinstance(c['area'][0], str):
                raise NotImplementedError("percentage areas are outside the B200 hot path")
    for k in conds:
        calculate_start_end_timesteps(model, conds[k])
    if hasattr(model, 'extra_conds'):
        for k in conds:
            conds[k] = encode_model_conds(model.extra_conds, conds[k], noise, device, k, latent_image=latent_image,
                                          denoise_mask=denoise_mask, seed=seed)
    return conds


# ----------------------------------------------------------------------------------------------
# guider
# ----------------------------------------------------------------------------------------------
_PLAIN_COND_KEYS_FORBIDDEN = ('area', 'mask', 'control', 'gligen', 'timestep_start', 'timestep_end')
_HOOK_KEYS = ("sampler_cfg_function", "model_function_wrapper", "denoise_mask_function")


class CFGGuider:
    """comfy/samplers.py:592-650; subclasses override predict_noise / set_conds like the reference's guiders."""
    def __init__(self, model_patcher):
        self.model_patcher = model_patcher
        self.model_options = model_patcher.model_options
        self.original_conds = {}
        self.cfg = 1.0
        self._fused = None

    def set_conds(self, positive, negative):
        self.inner_set_conds({"positive": positive, "negative": negative})

    def set_cfg(self, cfg):
        self.cfg = cfg

    def inner_set_conds(self, conds):
        for k in conds:
            self.original_conds[k] = convert_cond(conds[k])

    def __call__(self, *args, **kwargs):
        return self.predict_noise(*args, **kwargs)

    def predict_noise(self, x, timestep, model_options={}, seed=None):
        return sampling_function(self.inner_model, x, timestep, self.conds.get("negative", None),
                                 self.conds.get("positive", None), self.cfg, model_options=model_options, seed=seed)

    # -- fused fast path -----------------------------------------------------------------------
    def _plain_cond(self, key):
        lst = self.conds.get(key, None)
        if lst is None or len(lst) != 1:
            return None
        c = lst[0]
        if any(k in c for k in _PLAIN_COND_KEYS_FORBIDDEN) or c.get('strength', 1.0) != 1.0:
            return None
        mc = c.get("model_conds", {})
        if "c_crossattn" not in mc or any(k not in ("c_crossattn", "y") for k in mc):
            return None
        return mc

    def _fused_predicate(self, x, model_options):
        if type(self).predict_noise is not CFGGuider.predict_noise:
            return False
        if any(k in model_options for k in _HOOK_KEYS) or model_options.get("sampler_post_cfg_function"):
            return False
        m = self.inner_model
        if not isinstance(m.model_sampling, ModelSamplingDiscrete) or not hasattr(m, "unet_raw"):
            return False
        if m.manual_cast_dtype is not None or m.get_dtype() != torch.bfloat16:
            return False
        return x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 4

    def prepare_schedule(self, sig):
        """Per-step sigma and timestep vectors for the whole schedule, uploaded once."""
        self._schedule = sig

    def _build_fused(self, x, model_options):
        pos = self._plain_cond("positive")
        if pos is None:
            return None
        use_uncond = not (math.isclose(self.cfg, 1.0) and not model_options.get("disable_cfg1_optimization", False))
        neg = self._plain_cond("negative") if use_uncond else None
        if use_uncond and neg is None:
            return None
        if use_uncond and set(pos.keys()) != set(neg.keys()):
            return None
        b = x.shape[0]
        dev, dt = x.device, torch.bfloat16
        proc = lambda mc: {k: v.process_cond(batch_size=b, device=dev, area=None) for k, v in mc.items()}
        parts = ([proc(neg)] if use_uncond else []) + [proc(pos)]      # [uncond, cond]: cond_or_uncond == [1, 0]
        if use_uncond and not cond_equal_size(parts[0], parts[1]):
            return None
        parts, order = self._select_parts(parts, use_uncond)
        cat = cond_cat(parts)
        sig = getattr(self, "_schedule", None)
        ms = self.inner_model.model_sampling
        state = {
            "copies": len(parts), "context": cat["c_crossattn"].to(dt).contiguous(),
            "y": cat["y"].to(dt).contiguous() if "y" in cat else None,
            "cond_or_uncond": order, "batch": b, "tables": None,
            # CUDA-graph replay of the whole step ("auto": only when the host, not the GPU, limits the step)
            "calls": 0, "graph": None, "graph_opt": model_options.get("cuda_graph", "auto"), "probe": None,
        }
        if sig is not None:
            valid = sig[sig > 0]
            ts = torch.zeros(len(sig))
            ts[:len(valid)] = ms.timestep(valid).float()
            state["tables"] = (sig.view(-1, 1).repeat(1, b).to(dev),
                               ts.view(-1, 1).repeat(1, b * len(parts)).to(dev), sig.clone(), ts)
        return state

    def _select_parts(self, parts, use_uncond):
        """Which cond parts this rank evaluates, and their cond_or_uncond tags."""
        return parts, ([1, 0] if use_uncond else [0])

    def fused_step(self, x, i, sigma_hat, sampler, c0, c1, c2, c3, old_denoised, model_options={}, seed=None):
        """UNet on the [uncond, cond] batch + one kernel for denoise/CFG/update. Returns (denoised, x_next), or
        None when the fast-path predicate does not hold (the caller then runs the generic composition)."""
        if not self._fused_predicate(x, model_options):
            return None
        st = self._fused
        if st is False:                 # this sample() call's conds are not plain: the generic composition runs
            return None
        if st is None or st["batch"] != x.shape[0] or st["context"].device != x.device:
            st = self._fused = self._build_fused(x, model_options)
            if st is None:
                self._fused = False
                return None
        m = self.inner_model
        ms = m.model_sampling
        tables = st["tables"]
        sigma_f = float(sigma_hat)
        st["calls"] += 1
        if st["graph"] is None and st["calls"] == 1 and st["graph_opt"] not in (False, None):
            st["graph"] = self._reuse_step_graph(st, x, model_options)      # captured by an earlier sample() call?
            if st["graph"] is None and st["graph_opt"] == "auto" and type(self) is CFGGuider:
                # an earlier call already found the GPU to be the limiter for these shapes: do not probe again
                verdicts = self.inner_model.__dict__.get("_step_graph_verdicts", {})
                if verdicts.get(self._graph_key(st, x, model_options)) is False:
                    st["graph"] = False
        if st["graph"] is None and st["calls"] >= 3:
            st["graph"] = self._maybe_capture_step(st, x, model_options)
        if st["graph"]:
            return self._graph_step(st, x, i, sigma_f, sampler, c0, c1, c2, c3, old_denoised)
        # (the first step also pays one-time costs -- cross-attention K/V, cuDNN plan selection -- so the second is probed)
        probe = st["calls"] == 2 and st["graph_opt"] == "auto" and st["graph"] is None
        if probe:
            # with work still queued on the GPU the launch queue fills up and the host's enqueue time reads as the
            # GPU's: drain it once (once per model and shape, the verdict is remembered)
            torch.cuda.synchronize()
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
            t_host = time.perf_counter()
        if tables is not None and i < len(tables[2]) and float(tables[2][i]) == sigma_f:
            sigma_vec, t_vec = tables[0][i], tables[1][i]
        else:
            sigma_vec = torch.full((x.shape[0],), sigma_f, device=x.device)
            t = float(ms.timestep(torch.tensor([sigma_f]))[0])
            t_vec = torch.full((x.shape[0] * st["copies"],), t, device=x.device)
        xc = _native.unet_input(x, sigma_vec, st["copies"], float(ms.sigma_data))
        topt = dict(model_options.get("transformer_options", {}))
        topt["cond_or_uncond"] = list(st["cond_or_uncond"])
        topt["sigmas"] = sigma_vec
        out = m.unet_raw(xc, t_vec, st["context"], st["y"], topt)
        if not out.is_contiguous():          # channels_last UNets: the 4-channel output is tiny, make it NCHW
            out = out.contiguous()
        out_u, out_c = self._split_outputs(out, x.shape[0], st)
        res = _native.cfg_sampler_tail(x, out_u, out_c, sigma_vec, float(self.cfg), ms.pred_type, old_denoised,
                                       sampler, c0, c1, c2, c3, float(ms.sigma_data))
        if probe:
            st["probe"] = ((time.perf_counter() - t_host) * 1e3, ev)
            ev[1].record()
        return res

    # -- CUDA-graph replay of the step -------------------------------------------------------------
    # One denoise step is ~1100 kernel launches (SDXL); the host needs ~24 ms to enqueue them. With eight 1024^2
    # images per GPU the GPU works 120 ms on them and the host never matters; with one image the GPU is done in
    # 15 ms and waits for the host. The step is shape-static and, once the sampler coefficients are read from device
    # memory (cgs_cfg_sampler_tail_dev_fwd), identical for every step of the schedule: capture it once, replay it.
    def _maybe_capture_step(self, st, x, model_options):
        opt = st["graph_opt"]
        if opt is False or opt is None or type(self) is not CFGGuider or _native._timed_family is not None:
            return False
        if opt == "auto":
            if st["probe"] is None:
                return False
            host_ms, ev = st["probe"]
            ev[1].synchronize()
            host_bound = host_ms >= 0.8 * ev[0].elapsed_time(ev[1])
            self.inner_model.__dict__.setdefault("_step_graph_verdicts", {})[self._graph_key(st, x, model_options)] = host_bound
            if not host_bound:                                  # the GPU is the limiter: nothing to gain
                return False
        try:
            return self._capture_step(st, x, model_options)
        except Exception as e:      # capture is an optimisation of the same native path; run it launch by launch
            logging.warning("CUDA graph capture of the denoise step failed (%s); continuing without it", e)
            return False

    def _capture_step(self, st, x, model_options):
        m = self.inner_model
        ms = m.model_sampling
        b, copies = x.shape[0], st["copies"]
        n_par = b + b * copies + 8                    # [sigma x b | timestep x b*copies | mode, c0..c3, pad]
        g = {
            "P": torch.zeros(n_par, device=x.device, dtype=torch.float32),
            "X": torch.empty_like(x), "XN": torch.empty_like(x), "DEN": torch.empty_like(x),
            "OLD": torch.zeros_like(x), "last_den": None,
            # the conditioning lives in buffers of the graph's own, so that a later sample() call with other prompts
            # can reuse the graph: copy the new context in, recompute the cross-attention K/V in place, replay
            "CTX": st["context"].clone(), "Y": None if st["y"] is None else st["y"].clone(),
        }
        g["kv"] = self._prime_kv_caches(g["CTX"], None)
        sigma_vec, t_vec, step = g["P"][:b], g["P"][b:b + b * copies], g["P"][b + b * copies:]
        topt = dict(model_options.get("transformer_options", {}))
        topt["cond_or_uncond"] = list(st["cond_or_uncond"])
        topt["sigmas"] = sigma_vec
        cfg, sigma_data = float(self.cfg), float(ms.sigma_data)

        def body():
            xc = _native.unet_input(g["X"], sigma_vec, copies, sigma_data)
            out = m.unet_raw(xc, t_vec, g["CTX"], g["Y"], topt)
            if not out.is_contiguous():
                out = out.contiguous()
            out_u, out_c = self._split_outputs(out, b, st)
            _native.cfg_sampler_tail_dev(g["X"], out_u, out_c, sigma_vec, cfg, ms.pred_type, g["OLD"], step,
                                         g["DEN"], g["XN"], sigma_data)
            g["OLD"].copy_(g["DEN"])                     # the next step's old_denoised

        before = dict(_native.LAUNCHES)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            body()
        # capture launched nothing; remember what one replay launches
        g["counts"] = {k: v - before.get(k, 0) for k, v in _native.LAUNCHES.items() if v != before.get(k, 0)}
        for k, v in g["counts"].items():
            _native.LAUNCHES[k] -= v
        g["graph"] = graph
        cache = m.__dict__.setdefault("_step_graphs", collections.OrderedDict())
        cache[self._graph_key(st, x, model_options)] = g
        while len(cache) > 2:                              # a graph pins its activation pool: keep the two latest
            cache.popitem(last=False)
        return g

    # -- graphs that outlive a sample() call ----------------------------------------------------------
    def _graph_key(self, st, x, model_options):
        """Everything that is baked into a captured step: shapes, scalars that are kernel arguments, and the
        identity of the weights (a LoRA patch or a reload replaces or rewrites parameter tensors)."""
        m = self.inner_model
        params = list(m.diffusion_model.parameters())
        weights = hash(tuple((p.data_ptr(), p._version) for p in params))      # every tensor: ~1.7 k pointers per call
        topt = model_options.get("transformer_options", {})
        return (tuple(x.shape), str(x.device), st["copies"], tuple(st["cond_or_uncond"]), float(self.cfg),
                m.model_sampling.pred_type, float(m.model_sampling.sigma_data), tuple(st["context"].shape),
                None if st["y"] is None else tuple(st["y"].shape),
                tuple(sorted((str(k), id(v)) for k, v in topt.items())),
                len(params), weights,
                # the code the capture recorded: feature switches and the op bindings (tools and tests rebind them)
                tuple(sorted(_native.FEATURES.items())),
                tuple(id(getattr(_native, f)) for f in ("attention", "layer_norm", "add_layer_norm_", "group_norm",
                                                        "geglu", "linear", "residual_add", "cat_channels")))

    def _cross_attention_modules(self):
        from .attention import CrossAttention
        return [mod for mod in self.inner_model.diffusion_model.modules() if isinstance(mod, CrossAttention)]

    def _prime_kv_caches(self, ctx, kv):
        """Cross-attention K/V of `ctx` for every layer that caches them. First time (kv None): fresh tensors, handed
        to the layers' caches, so that the capture bakes their addresses. Later: recomputed into the same tensors."""
        out = []
        if kv is None:
            for mod in self._cross_attention_modules():
                if mod._kv_cache is None:
                    continue                                   # a layer that did not take the cached path
                k, v = mod.to_k(ctx), mod.to_v(ctx)
                mod.set_kv_cache(ctx, k, v)
                out.append((mod, k, v))
            return out
        for mod, k, v in kv:
            k.copy_(mod.to_k(ctx))
            v.copy_(mod.to_v(ctx))
            mod.set_kv_cache(ctx, k, v)
        return kv

    def _reuse_step_graph(self, st, x, model_options):
        if type(self) is not CFGGuider or _native._timed_family is not None:
            return None
        cache = self.inner_model.__dict__.get("_step_graphs")
        if not cache:
            return None
        key = self._graph_key(st, x, model_options)
        g = cache.get(key)
        if g is None:
            return None
        cache.move_to_end(key)
        g["CTX"].copy_(st["context"])
        if g["Y"] is not None:
            g["Y"].copy_(st["y"])
        self._prime_kv_caches(g["CTX"], g["kv"])
        g["last_den"] = None
        return g

    def _graph_step(self, st, x, i, sigma_f, sampler, c0, c1, c2, c3, old_denoised):
        g = st["graph"]
        b, copies = x.shape[0], st["copies"]
        if st.get("stage") is None:                       # pinned rows of this sample() call (not the graph's: the host
            st["stage"] = torch.zeros(64, g["P"].numel(), dtype=torch.float32).pin_memory()   # runs steps ahead)
        if i >= st["stage"].shape[0]:
            grown = torch.zeros(2 * (i + 1), st["stage"].shape[1], dtype=torch.float32).pin_memory()
            grown[:st["stage"].shape[0]] = st["stage"]
            st["stage"] = grown
        tables = st["tables"]
        if tables is not None and i < len(tables[2]) and float(tables[2][i]) == sigma_f:
            t = float(tables[3][i])
        else:
            t = float(self.inner_model.model_sampling.timestep(torch.tensor([sigma_f]))[0])
        row = st["stage"][i]                             # one pinned row per step: never rewritten while in flight
        row[:b] = sigma_f
        row[b:b + b * copies] = t
        row[b + b * copies:b + b * copies + 5] = torch.tensor([float(sampler), c0, c1, c2, c3])
        g["P"].copy_(row, non_blocking=True)
        g["X"].copy_(x)
        if sampler == 3 and old_denoised is not None and old_denoised is not g["last_den"]:
            g["OLD"].copy_(old_denoised)                 # first replayed step: denoised of the last eager step
        g["graph"].replay()
        _native.add_replayed_launches(g["counts"])
        g["last_den"] = g["DEN"].clone()
        return g["last_den"], g["XN"].clone()

    def _split_outputs(self, out, b, st):
        """(uncond, cond) halves of the batched model output; uncond is None when CFG is off."""
        if st["copies"] == 2:
            return out[:b], out[b:]
        return None, out

    # -- reference flow ------------------------------------------------------------------------
    def inner_sample(self, noise, latent_image, device, sampler, sigmas, denoise_mask, callback, disable_pbar, seed):
        if latent_image is not None and torch.count_nonzero(latent_image) > 0:
            latent_image = self.inner_model.process_latent_in(latent_image)
        self.conds = process_conds(self.inner_model, noise, self.conds, device, latent_image, denoise_mask, seed)
        extra_args = {"model_options": self.model_options, "seed": seed}
        samples = sampler.sample(self, sigmas, extra_args, callback, noise, latent_image, denoise_mask, disable_pbar)
        return self.inner_model.process_latent_out(samples.to(torch.float32))

    def sample(self, noise, latent_image, sampler, sigmas, denoise_mask=None, callback=None, disable_pbar=False,
               seed=None):
        if sigmas.shape[-1] == 0:
            return latent_image
        self.conds = {}
        for k in self.original_conds:
            self.conds[k] = list(map(lambda a: a.copy(), self.original_conds[k]))
        self._fused = None
        self._schedule = None
        self.inner_model, self.loaded_models = model_management.prepare_sampling(self.model_patcher, noise.shape)
        device = self.model_patcher.load_device
        if denoise_mask is not None:
            denoise_mask = model_management.prepare_mask(denoise_mask, noise.shape, device)
        noise = noise.to(device)
        latent_image = latent_image.to(device)
        sigmas = sigmas.to(device)
        output = self.inner_sample(noise, latent_image, device, sampler, sigmas, denoise_mask, callback,
                                   disable_pbar, seed)
        del self.inner_model
        del self.conds
        del self.loaded_models
        self._fused = None
        return output


class SplitCFGGuider(CFGGuider):
    """CFG pair split across two ranks (model_management.Placement mode "cfg_split"): this rank evaluates only
    its half (cond or uncond) of the UNet batch; a 2-rank all-gather of the raw outputs per step gives both
    ranks (uncond, cond), and both run the identical fused tail, so x stays bit-identical on the pair.
    Only the fused fast path is supported (plain conds, no hooks); anything else must run unsplit."""
    def __init__(self, model_patcher, placement, group):
        super().__init__(model_patcher)
        self.placement = placement
        self.group = group

    def _build_fused(self, x, model_options):
        """Build the batched [uncond, cond] conditioning exactly as the unsplit guider does (including the LCM
        repeat-padding of mismatched text lengths, comfy/conds.py:50-63), then keep only this rank's half, so a
        split run sees the same per-sample inputs as the batched run."""
        st = super()._build_fused(x, model_options)
        if st is None:
            return None
        if st["copies"] != 2:
            raise RuntimeError("cfg_split placement needs cfg != 1 (there is no uncond half to split off)")
        b = st["batch"]
        half = slice(0, b) if self.placement.role == "uncond" else slice(b, 2 * b)
        st["context"] = st["context"][half].contiguous()
        if st["y"] is not None:
            st["y"] = st["y"][half].contiguous()
        if st["tables"] is not None:
            st["tables"] = (st["tables"][0], st["tables"][1][:, :b].contiguous(), st["tables"][2])
        st["copies"] = 1
        st["cond_or_uncond"] = [1] if self.placement.role == "uncond" else [0]
        return st

    def _split_outputs(self, out, b, st):
        return model_management.exchange_cfg_halves(out, self.placement, self.group)

    def predict_noise(self, x, timestep, model_options={}, seed=None):
        raise RuntimeError("SplitCFGGuider supports only the fused fast path (plain conds, no sampler hooks)")

    def _fused_predicate(self, x, model_options):
        if any(k in model_options for k in _HOOK_KEYS) or model_options.get("sampler_post_cfg_function"):
            return False
        m = self.inner_model
        return isinstance(m.model_sampling, ModelSamplingDiscrete) and x.is_cuda and x.dtype == torch.float32 \
            and x.is_contiguous() and m.get_dtype() == torch.bfloat16


# ----------------------------------------------------------------------------------------------
# sampler objects (comfy/samplers.py:498-558, 680-689)
# ----------------------------------------------------------------------------------------------
class Sampler:
    def sample(self):
        pass

    def max_denoise(self, model_wrap, sigmas):
        max_sigma = float(model_wrap.inner_model.model_sampling.sigma_max)
        sigma = float(sigmas[0])
        return math.isclose(max_sigma, sigma, rel_tol=1e-05) or sigma > max_sigma


class KSAMPLER(Sampler):
    def __init__(self, sampler_function, extra_options={}, inpaint_options={}):
        self.sampler_function = sampler_function
        self.extra_options = extra_options
        self.inpaint_options = inpaint_options

    def sample(self, model_wrap, sigmas, extra_args, callback, noise, latent_image=None, denoise_mask=None,
               disable_pbar=False):
        extra_args["denoise_mask"] = denoise_mask
        model_k = KSamplerX0Inpaint(model_wrap, sigmas)
        model_k.latent_image = latent_image
        if self.inpaint_options.get("random", False):
            generator = torch.manual_seed(extra_args.get("seed", 41) + 1)
            model_k.noise = torch.randn(noise.shape, generator=generator, device="cpu").to(noise.dtype).to(noise.device)
        else:
            model_k.noise = noise
        noise = model_wrap.inner_model.model_sampling.noise_scaling(sigmas[0], noise, latent_image,
                                                                    self.max_denoise(model_wrap, sigmas))
        k_callback = None
        total_steps = len(sigmas) - 1
        if callback is not None:
            k_callback = lambda x: callback(x["i"], x["denoised"], x["x"], total_steps)
        samples = self.sampler_function(model_k, noise, sigmas, extra_args=extra_args, callback=k_callback,
                                        disable=disable_pbar, **self.extra_options)
        return model_wrap.inner_model.model_sampling.inverse_noise_scaling(sigmas[-1], samples)


def ksampler(sampler_name, extra_options={}, inpaint_options={}):
    fn = getattr(k_diffusion_sampling, "sample_{}".format(sampler_name), None)
    if fn is None:
        raise NotImplementedError(f"sampler '{sampler_name}' is outside the B200 hot path "
                                  f"(available: {KSAMPLER_NAMES})")
    return KSAMPLER(fn, extra_options, inpaint_options)


def sampler_object(name):
    if name == "ddim":
        return ksampler("euler", inpaint_options={"random": True})
    return ksampler(name)


def sample(model, noise, positive, negative, cfg, device, sampler, sigmas, model_options={}, latent_image=None,
           denoise_mask=None, callback=None, disable_pbar=False, seed=None):
    cfg_guider = CFGGuider(model)
    cfg_guider.set_conds(positive, negative)
    cfg_guider.set_cfg(cfg)
    return cfg_guider.sample(noise, latent_image, sampler, sigmas, denoise_mask, callback, disable_pbar, seed)


class KSampler:
    """comfy/samplers.py:691-755."""
    SCHEDULERS = SCHEDULER_NAMES
    SAMPLERS = SAMPLER_NAMES

    def __init__(self, model, steps, device, sampler=None, scheduler=None, denoise=None, model_options={}):
        self.model = model
        self.device = device
        if scheduler not in self.SCHEDULERS:
            scheduler = self.SCHEDULERS[0]
        if sampler not in self.SAMPLERS:
            sampler = self.SAMPLERS[0]
        self.scheduler = scheduler
        self.sampler = sampler
        self.set_steps(steps, denoise)
        self.denoise = denoise
        self.model_options = model_options

    def calculate_sigmas(self, steps):
        return calculate_sigmas(self.model.get_model_object("model_sampling"), self.scheduler, steps)

    def set_steps(self, steps, denoise=None):
        self.steps = steps
        if denoise is None or denoise > 0.9999:
            self.sigmas = self.calculate_sigmas(steps).to(self.device)
        else:
            if denoise <= 0.0:
                self.sigmas = torch.FloatTensor([])
            else:
                new_steps = int(steps / denoise)
                sigmas = self.calculate_sigmas(new_steps).to(self.device)
                self.sigmas = sigmas[-(steps + 1):]

    def sample(self, noise, positive, negative, cfg, latent_image=None, start_step=None, last_step=None,
               force_full_denoise=False, denoise_mask=None, sigmas=None, callback=None, disable_pbar=False, seed=None):
        if sigmas is None:
            sigmas = self.sigmas
        if last_step is not None and last_step < (len(sigmas) - 1):
            sigmas = sigmas[:last_step + 1]
            if force_full_denoise:
                sigmas[-1] = 0
        if start_step is not None:
            if start_step < (len(sigmas) - 1):
                sigmas = sigmas[start_step:]
            else:
                if latent_image is not None:
                    return latent_image
                return torch.zeros_like(noise)
        sampler = sampler_object(self.sampler)
        return sample(self.model, noise, positive, negative, cfg, self.device, sampler, sigmas, self.model_options,
                      latent_image=latent_image, denoise_mask=denoise_mask, callback=callback,
                      disable_pbar=disable_pbar, seed=seed)
