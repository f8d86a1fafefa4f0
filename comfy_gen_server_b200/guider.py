"""The fused denoise step of the B200 path: one UNet call on the [uncond x B, cond x B] batch followed by ONE kernel
that does eps/v -> denoised, the CFG combine and the sampler update -- and, when the host is the limiter, the whole
step replayed as a CUDA graph.

It stands in for the reference's per-step composition

    KSamplerX0Inpaint.__call__ -> CFGGuider.predict_noise -> sampling_function -> calc_cond_batch
        -> BaseModel.apply_model -> cfg_function                  (comfy/samplers.py:131-275, model_base.py:74-98)
    + the sampler's own update                                     (comfy/k_diffusion/sampling.py:126-144, 570-593)

but only when nothing can observe the difference ("fast-path predicate"): a stock CFGGuider
(predict_noise not overridden), no sampler_cfg / post_cfg / model_function_wrapper / denoise_mask hooks, one plain
full-frame cond per side (no area, mask, strength, timestep range, ControlNet, GLIGEN), the stock discrete
eps / v-prediction model_sampling, a bf16 UNet. FusedDenoise.create() returns None otherwise and the sampler
functions (sampling.py) call the model exactly as the reference's samplers do.

With plain full-frame conds the reference's accumulate / divide (samplers.py:137-138, 222-226) is the identity
(mult == 1, counts == 1 + 1e-37 == 1 in fp32), so skipping it is exact; the tail kernel evaluates the reference's
fp32 expressions in the reference's order and is bit-identical to them (tests/test_gpu_kernels.py).
"""
from __future__ import annotations

import collections
import logging
import math
import time

import torch

from . import _native
from . import block as B
from . import layers

_PLAIN_COND_KEYS_FORBIDDEN = ('area', 'mask', 'control', 'gligen', 'timestep_start', 'timestep_end')
_HOOK_KEYS = ("sampler_cfg_function", "model_function_wrapper", "denoise_mask_function")


def prediction_type(ms):
    """0 (eps) / 1 (v) when `ms` is the reference's stock discrete-schedule model_sampling, else None."""
    import comfy.model_sampling as M
    t = type(ms)
    if not isinstance(ms, M.ModelSamplingDiscrete) or t.timestep is not M.ModelSamplingDiscrete.timestep \
            or t.calculate_input is not M.EPS.calculate_input:
        return None
    if t.calculate_denoised is M.EPS.calculate_denoised:
        return 0
    if t.calculate_denoised is M.V_PREDICTION.calculate_denoised:
        return 1
    return None


def _plain_model_conds(cond_list):
    """model_conds of a cond list that is one plain full-frame entry, else None."""
    if cond_list is None or len(cond_list) != 1:
        return None
    c = cond_list[0]
    if any(k in c for k in _PLAIN_COND_KEYS_FORBIDDEN) or c.get('strength', 1.0) != 1.0:
        return None
    mc = c.get("model_conds", {})
    if "c_crossattn" not in mc or any(k not in ("c_crossattn", "y") for k in mc):
        return None
    return mc


class FusedDenoise:
    """Per-sample()-call state of the fused step (conditioning batch, sigma / timestep tables, step graph)."""

    # ------------------------------------------------------------------------------------------
    @classmethod
    def create(cls, model_k, x, extra_args, sig):
        """model_k: what the reference hands a sampler function (KSamplerX0Inpaint around the guider);
        sig: host copy of the schedule. Returns None when the fast-path predicate does not hold."""
        import comfy.samplers as S
        guider = getattr(model_k, "inner_model", None)
        if not isinstance(guider, S.CFGGuider) or not isinstance(model_k, S.KSamplerX0Inpaint):
            return None
        split = getattr(guider, "cgs_split", None)                  # placement.SplitCFGGuider
        if split is None and type(guider).predict_noise is not S.CFGGuider.predict_noise:
            return None
        if extra_args.get("denoise_mask") is not None:
            return None
        model_options = extra_args.get("model_options", {})
        if any(k in model_options for k in _HOOK_KEYS) or model_options.get("sampler_post_cfg_function"):
            return None
        m = getattr(guider, "inner_model", None)
        if m is None or not hasattr(m, "diffusion_model") or m.manual_cast_dtype is not None \
                or m.get_dtype() != layers.ACT_DTYPE or getattr(m, "concat_keys", ()):
            return None
        pred = prediction_type(m.model_sampling)
        if pred is None:
            return None
        if not (layers.is_latent(x) and x.dim() == 4 and (x.numel() // max(x.shape[0], 1)) % 4 == 0):
            return None
        self = cls()
        self.guider, self.model, self.ms, self.pred_type = guider, m, m.model_sampling, pred
        self.model_options, self.split = model_options, split
        self.cfg = float(guider.cfg)
        if not self._build(x, sig):
            return None
        return self

    def _build(self, x, sig) -> bool:
        import comfy.samplers as S
        conds = self.guider.conds
        pos = _plain_model_conds(conds.get("positive"))
        if pos is None:
            return False
        use_uncond = not (math.isclose(self.cfg, 1.0) and not self.model_options.get("disable_cfg1_optimization", False))
        neg = _plain_model_conds(conds.get("negative")) if use_uncond else None
        if use_uncond and (neg is None or set(pos.keys()) != set(neg.keys())):
            return False
        b, dev, dt = x.shape[0], x.device, layers.ACT_DTYPE
        proc = lambda mc: {k: v.process_cond(batch_size=b, device=dev, area=None) for k, v in mc.items()}
        parts = ([proc(neg)] if use_uncond else []) + [proc(pos)]        # [uncond, cond]: cond_or_uncond == [1, 0]
        if use_uncond and not S.cond_equal_size(parts[0], parts[1]):
            return False
        cat = S.cond_cat(parts)          # incl. the LCM repeat-padding of unequal text lengths (comfy/conds.py:50-63)
        self.copies = len(parts)
        self.cond_or_uncond = [1, 0] if use_uncond else [0]
        self.context = cat["c_crossattn"].to(dt).contiguous()
        self.y = cat["y"].to(dt).contiguous() if "y" in cat else None
        if self.split is not None:
            if self.copies != 2:
                raise RuntimeError("cfg_split placement needs cfg != 1 (there is no uncond half to split off)")
            # built exactly as the unsplit guider builds it, then this rank keeps its half: a split run sees the
            # same per-sample inputs as the batched run
            half = slice(0, b) if self.split.placement.role == "uncond" else slice(b, 2 * b)
            self.context = self.context[half].contiguous()
            self.y = None if self.y is None else self.y[half].contiguous()
            self.copies = 1
            self.cond_or_uncond = [1] if self.split.placement.role == "uncond" else [0]
        self.batch = b
        valid = sig[sig > 0]
        ts = torch.zeros(len(sig))
        ts[:len(valid)] = self.ms.timestep(valid).float()
        self.sig_host, self.ts_host = sig.clone(), ts
        self.sigma_table = sig.view(-1, 1).repeat(1, b).to(dev)
        self.ts_table = ts.view(-1, 1).repeat(1, b * self.copies).to(dev)
        self.calls = 0
        self.graph = None
        self.graph_opt = self.model_options.get("cuda_graph", "auto")
        self.probe = None
        self.stage = None
        return True

    # ------------------------------------------------------------------------------------------
    def _unet(self, xc, t_vec, context, y, sigma_vec):
        topt = dict(self.model_options.get("transformer_options", {}))
        topt["cond_or_uncond"] = list(self.cond_or_uncond)
        topt["sigmas"] = sigma_vec
        kw = {} if y is None else {"y": y}
        out = self.model.diffusion_model(xc, t_vec, context=context, control=None, transformer_options=topt, **kw)
        if not out.is_contiguous():          # channels-last UNets: the 4-channel output is tiny, make it NCHW
            out = out.contiguous()
        return out

    def _halves(self, out, b):
        """(uncond, cond) raw model outputs for the tail kernel; uncond is None when CFG is off."""
        if self.split is not None:
            return self.split.exchange(out)
        if self.copies == 2:
            return out[:b], out[b:]
        return None, out

    def _sigma_and_t(self, i, sigma_f, x):
        if i < len(self.sig_host) and float(self.sig_host[i]) == sigma_f:
            return self.sigma_table[i], self.ts_table[i], float(self.ts_host[i])
        t = float(self.ms.timestep(torch.tensor([sigma_f]))[0])
        return (torch.full((x.shape[0],), sigma_f, device=x.device),
                torch.full((x.shape[0] * self.copies,), t, device=x.device), t)

    def step(self, x, i, sigma_hat, sampler, c0, c1, c2, c3, old_denoised):
        """One denoise step. sampler: 1 Euler (c0 = sigma_hat, c1 = dt), 2 / 3 DPM++ 2M first-order / multistep
        (c0..c3 = c_x, c_d, c_cur, c_old). Returns (denoised, x_next)."""
        sigma_f = float(sigma_hat)
        self.calls += 1
        if self.graph is None and self.calls == 1 and self.graph_opt not in (False, None):
            self.graph = self._reuse_step_graph(x)                  # captured by an earlier sample() call?
            if self.graph is None and self.graph_opt == "auto":
                # an earlier call already found the GPU to be the limiter for these shapes: do not probe again
                verdicts = self.model.__dict__.get("_cgs_graph_verdicts", {})
                if verdicts.get(self._graph_key(x)) is False:
                    self.graph = False
        if self.graph is None and (self.calls >= 3 or (self.graph_opt is True and self.calls >= 2)):
            self.graph = self._maybe_capture_step(x)
        if self.graph:
            return self._graph_step(x, i, sigma_f, sampler, c0, c1, c2, c3, old_denoised)
        # (the first step also pays one-time costs -- cross-attention K/V, cuDNN plan selection -- so the second is probed)
        probe = self.calls == 2 and self.graph_opt == "auto" and self.graph is None
        if probe:
            # with work still queued on the GPU the launch queue fills up and the host's enqueue time reads as the
            # GPU's: drain it once (once per model and shape, the verdict is remembered)
            torch.cuda.synchronize()
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
            t_host = time.perf_counter()
        sigma_vec, t_vec, _ = self._sigma_and_t(i, sigma_f, x)
        xc = _native.unet_input(x, sigma_vec, self.copies, float(self.ms.sigma_data))
        out = self._unet(xc, t_vec, self.context, self.y, sigma_vec)
        if self.split is not None and self.split.fuses_tail:
            res = self.split.tail_eager(x, out, sigma_vec, self.cfg, self.pred_type, old_denoised, sampler,
                                        c0, c1, c2, c3, float(self.ms.sigma_data))
        else:
            out_u, out_c = self._halves(out, x.shape[0])
            res = _native.cfg_sampler_tail(x, out_u, out_c, sigma_vec, self.cfg, self.pred_type, old_denoised,
                                           sampler, c0, c1, c2, c3, float(self.ms.sigma_data))
        if probe:
            self.probe = ((time.perf_counter() - t_host) * 1e3, ev)
            ev[1].record()
        return res

    # -- CUDA-graph replay of the step -------------------------------------------------------------
    # One denoise step is ~1100 kernel launches (SDXL); the host needs ~24 ms to enqueue them. With eight 1024^2
    # images per GPU the GPU works 115 ms on them and the host never matters; with one image the GPU is done in
    # 15 ms and waits for the host. The step is shape-static and, once the sampler coefficients are read from device
    # memory (cgs_cfg_sampler_tail_dev_fwd), identical for every step of the schedule: capture it once, replay it.
    def _graph_allowed(self) -> bool:
        if self.graph_opt in (False, None) or _native._timed_family is not None:
            return False
        return self.split is None or self.split.graph_capturable

    def _maybe_capture_step(self, x):
        if not self._graph_allowed():
            return False
        if self.graph_opt == "auto":
            if self.probe is None:
                return False
            host_ms, ev = self.probe
            ev[1].synchronize()
            # (a replay also closes the gaps between the ~1070 launches on the GPU side: worth it from about half)
            host_bound = host_ms >= 0.5 * ev[0].elapsed_time(ev[1])
            if self.split is not None:
                host_bound = self.split.agree(host_bound)          # both ranks of a pair must take the same path
            self.model.__dict__.setdefault("_cgs_graph_verdicts", {})[self._graph_key(x)] = host_bound
            if not host_bound:                                  # the GPU is the limiter: nothing to gain
                return False
        try:
            return self._capture_step(x)
        except Exception as e:      # capture is an optimisation of the same native path; run it launch by launch
            if self.split is not None:
                raise               # (a pair whose ranks disagree about replaying would dead-lock in the exchange)
            logging.warning("CUDA graph capture of the denoise step failed (%s); continuing without it", e)
            return False

    def _capture_step(self, x):
        b, copies = x.shape[0], self.copies
        n_par = b + b * copies + 8                    # [sigma x b | timestep x b*copies | mode, c0..c3, pad]
        g = {
            "P": torch.zeros(n_par, device=x.device, dtype=torch.float32),
            "X": torch.empty_like(x), "XN": torch.empty_like(x), "DEN": torch.empty_like(x),
            "OLD": torch.zeros_like(x), "last_den": None,
            # the conditioning lives in buffers of the graph's own, so that a later sample() call with other prompts
            # can reuse the graph: copy the new context in, recompute the cross-attention K/V in place, replay
            "CTX": self.context.clone(), "Y": None if self.y is None else self.y.clone(),
        }
        g["kv"] = self._prime_kv_caches(g["CTX"], None)
        sigma_vec, t_vec, step = g["P"][:b], g["P"][b:b + b * copies], g["P"][b + b * copies:]
        sigma_data = float(self.ms.sigma_data)

        def body():
            xc = _native.unet_input(g["X"], sigma_vec, copies, sigma_data)
            out = self._unet(xc, t_vec, g["CTX"], g["Y"], sigma_vec)
            if self.split is not None and self.split.fuses_tail:
                self.split.tail_dev(g["X"], out, sigma_vec, self.cfg, self.pred_type, g["OLD"], step, g["DEN"],
                                    g["XN"], sigma_data)
            else:
                out_u, out_c = self._halves(out, b)
                _native.cfg_sampler_tail_dev(g["X"], out_u, out_c, sigma_vec, self.cfg, self.pred_type, g["OLD"],
                                             step, g["DEN"], g["XN"], sigma_data)
            g["OLD"].copy_(g["DEN"])                     # the next step's old_denoised

        before = dict(_native.LAUNCHES)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            body()
        # capture launched nothing; remember what one replay launches
        g["counts"] = {k: v - before.get(k, 0) for k, v in _native.LAUNCHES.items() if v != before.get(k, 0)}
        for k, v in g["counts"].items():
            _native.LAUNCHES[k] -= v
        g["graph"] = graph
        g["derived"] = B.derived_state(self.model.diffusion_model)      # what the capture baked in stays alive with it
        cache = self.model.__dict__.setdefault("_cgs_step_graphs", collections.OrderedDict())
        cache[self._graph_key(x)] = g
        while len(cache) > 2:                              # a graph pins its activation pool: keep the two latest
            cache.popitem(last=False)
        return g

    def _graph_key(self, x):
        """Everything that is baked into a captured step: shapes, scalars that are kernel arguments, the identity of
        every weight tensor (a LoRA patch or a reload replaces or rewrites them) and of every patch object."""
        m = self.model
        params = list(m.diffusion_model.parameters())
        weights = hash(tuple((p.data_ptr(), B._ver(p)) for p in params))
        topt = self.model_options.get("transformer_options", {})
        return (tuple(x.shape), str(x.device), self.copies, tuple(self.cond_or_uncond), self.cfg, self.pred_type,
                float(self.ms.sigma_data), tuple(self.context.shape), None if self.y is None else tuple(self.y.shape),
                tuple(sorted((str(k), id(v)) for k, v in topt.items())), len(params), weights,
                None if self.split is None else self.split.key(),
                # the code the capture recorded: the op bindings (tools and tests rebind them)
                tuple(sorted(_native.FEATURES.items())),
                tuple(id(getattr(_native, f)) for f in ("attention", "layer_norm", "add_layer_norm_", "group_norm",
                                                        "geglu", "linear", "residual_add", "cat_channels")))

    def _cross_attention_modules(self):
        import comfy.ldm.modules.attention as A
        return [mod for mod in self.model.diffusion_model.modules() if isinstance(mod, A.CrossAttention)]

    def _prime_kv_caches(self, ctx, kv):
        """Cross-attention K/V of `ctx` for every layer that caches them. First time (kv None): fresh tensors, handed
        to the layers' caches, so that the capture bakes their addresses. Later: recomputed into the same tensors."""
        if kv is None:
            out = []
            for mod in self._cross_attention_modules():
                if B.get_kv_cache(mod) is None:
                    continue                                   # a layer that did not take the cached path
                k, v = mod.to_k(ctx), mod.to_v(ctx)
                B.set_kv_cache(mod, ctx, k, v)
                out.append((mod, k, v))
            return out
        for mod, k, v in kv:
            k.copy_(mod.to_k(ctx))
            v.copy_(mod.to_v(ctx))
            B.set_kv_cache(mod, ctx, k, v)
        return kv

    def _reuse_step_graph(self, x):
        if not self._graph_allowed():
            return None
        cache = self.model.__dict__.get("_cgs_step_graphs")
        if not cache:
            return None
        key = self._graph_key(x)
        g = cache.get(key)
        if g is None:
            return None
        cache.move_to_end(key)
        g["CTX"].copy_(self.context)
        if g["Y"] is not None:
            g["Y"].copy_(self.y)
        self._prime_kv_caches(g["CTX"], g["kv"])
        g["last_den"] = None
        return g

    def _graph_step(self, x, i, sigma_f, sampler, c0, c1, c2, c3, old_denoised):
        g = self.graph
        b, copies = x.shape[0], self.copies
        if self.stage is None:                            # pinned rows of this sample() call (not the graph's: the host
            self.stage = torch.zeros(64, g["P"].numel(), dtype=torch.float32).pin_memory()   # runs steps ahead)
        if i >= self.stage.shape[0]:
            grown = torch.zeros(2 * (i + 1), self.stage.shape[1], dtype=torch.float32).pin_memory()
            grown[:self.stage.shape[0]] = self.stage
            self.stage = grown
        _, _, t = self._sigma_and_t(i, sigma_f, x) if not (i < len(self.sig_host) and float(self.sig_host[i]) == sigma_f) \
            else (None, None, float(self.ts_host[i]))
        row = self.stage[i]                              # one pinned row per step: never rewritten while in flight
        row[:b] = sigma_f
        row[b:b + b * copies] = t
        row[b + b * copies:b + b * copies + 5] = torch.tensor([float(sampler), c0, c1, c2, c3])
        if self.split is not None and self.split.fuses_tail:       # sequence number of this pair exchange
            row[b + b * copies + 5:b + b * copies + 6].view(torch.int32)[0] = self.split.next_seq(x.numel(), x.device)
        g["P"].copy_(row, non_blocking=True)
        g["X"].copy_(x)
        if sampler == 3 and old_denoised is not None and old_denoised is not g["last_den"]:
            g["OLD"].copy_(old_denoised)                 # first replayed step: denoised of the last eager step
        g["graph"].replay()
        _native.add_replayed_launches(g["counts"])
        g["last_den"] = g["DEN"].clone()
        return g["last_den"], g["XN"].clone()


# ----------------------------------------------------------------------------------------------
# BaseModel.apply_model  (comfy/model_base.py:74-98) -- the generic path's model call
# ----------------------------------------------------------------------------------------------
def apply_model_forward(self, x, t, c_concat=None, c_crossattn=None, control=None, transformer_options={}, **kwargs):
    """The reference's semantics for any caller (hooks, wrappers, custom guiders); for fp32 CUDA latents and a bf16
    UNet the sigma scaling + cast and the eps / v -> denoised conversion are one kernel each instead of ~12 ATen
    launches. Anything else is the reference's own method."""
    from .layers import ORIGINAL
    pred = prediction_type(self.model_sampling)
    plain = (c_concat is None and c_crossattn is not None and pred is not None and self.manual_cast_dtype is None
             and self.get_dtype() == layers.ACT_DTYPE and layers.is_latent(x) and x.dim() == 4 and (x.numel() // max(x.shape[0], 1)) % 4 == 0)
    if not plain:
        return ORIGINAL["BaseModel.apply_model"](self, x, t, c_concat=c_concat, c_crossattn=c_crossattn, control=control,
                                                 transformer_options=transformer_options, **kwargs)
    sigma = t.float().contiguous()
    sigma_data = float(self.model_sampling.sigma_data)
    xc = _native.unet_input(x, sigma, 1, sigma_data)
    extra = {k: (v.to(layers.ACT_DTYPE) if hasattr(v, "dtype") and v.dtype not in (torch.int, torch.long) else v)
             for k, v in kwargs.items()}
    out = self.diffusion_model(xc, self.model_sampling.timestep(t).float(), context=c_crossattn.to(layers.ACT_DTYPE),
                               control=control, transformer_options=transformer_options, **extra)
    return _native.cfg_denoise(x, None, out.contiguous(), sigma, 1.0, pred, sigma_data)
