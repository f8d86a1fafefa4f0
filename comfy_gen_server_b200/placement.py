"""Single-box multi-GPU placement for the denoise path (north_star: "comfy/model_management.py gains ... placement
that shards the latent batch and cond/uncond pairs across up to 8 B200s"). The reference has no multi-device code at
all (device selection is torch.cuda.current_device() only, comfy/model_management.py:74-88);
integration.install() publishes these functions on comfy.model_management next to its single-device helpers.

One process per GPU (torch.distributed, NCCL over NVLink / NVSwitch), weights replicated, and two ways to split a job:

  batch      rank r denoises a contiguous slice of the images through the whole schedule (both CFG halves).
             No data-path collective; ONE all-gather of the final latents.
  cfg_split  ranks are paired (2k, 2k+1); a pair shares an image slice, rank 2k evaluates the cond half of the UNet
             batch, rank 2k+1 the uncond half, and the raw model outputs are exchanged every step so that both ranks
             run the identical fused CFG + sampler kernel and stay in lock-step. Used when there are fewer images
             than GPUs (BASELINE configs 4, 5) or on request.

Noise is generated once for the FULL batch from the CPU generator (comfy/sample.py:8-25) and sliced, so the latents
equal the 1-GPU run's whatever the world size.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist

from .reference import ensure_importable


@dataclass
class Placement:
    """Which slice of the latent batch, and which CFG half, a rank owns."""
    rank: int
    world_size: int
    mode: str                           # "batch" | "cfg_split"
    start: int
    end: int
    role: Optional[str] = None          # "cond" | "uncond" in cfg_split mode
    pair: Optional[Tuple[int, int]] = None


def shard_range(total: int, parts: int, index: int) -> Tuple[int, int]:
    """Contiguous near-equal slices; the first (total % parts) slices get one extra item."""
    base, extra = divmod(total, parts)
    start = index * base + min(index, extra)
    return start, start + base + (1 if index < extra else 0)


def plan_placement(batch: int, world_size: int, rank: int, cfg_split: Optional[bool] = None) -> Placement:
    """Batch sharding when there are at least as many images as ranks; otherwise (or on request) split the
    cond/uncond pair too (configs 4 and 5)."""
    if world_size == 1:
        return Placement(rank, 1, "batch", 0, batch)
    if cfg_split is None:
        cfg_split = batch < world_size and world_size % 2 == 0
    if not cfg_split:
        s, e = shard_range(batch, world_size, rank)
        return Placement(rank, world_size, "batch", s, e)
    if world_size % 2 != 0:
        raise ValueError("cfg_split needs an even number of ranks")
    pair_index = rank // 2
    s, e = shard_range(batch, world_size // 2, pair_index)
    return Placement(rank, world_size, "cfg_split", s, e, "cond" if rank % 2 == 0 else "uncond",
                     (2 * pair_index, 2 * pair_index + 1))


def gather_latents(local: torch.Tensor, placements: List[Placement], group=None) -> torch.Tensor:
    """The one collective of the batch-sharded path: all-gather the final latents [b_r, C, h, w] (ragged
    slices are padded to the largest) and reassemble the full batch in order on every rank."""
    world = dist.get_world_size(group)
    if world == 1:
        return local
    sizes = [p.end - p.start for p in placements]
    mx = max(sizes)
    pad = local
    if local.shape[0] < mx:
        pad = torch.cat([local, local.new_zeros((mx - local.shape[0],) + tuple(local.shape[1:]))])
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad.contiguous(), group=group)
    out = []
    seen = set()
    for p, buf in zip(placements, bufs):
        if (p.start, p.end) in seen:        # cfg_split: both ranks of a pair hold the same slice
            continue
        seen.add((p.start, p.end))
        out.append(buf[:p.end - p.start])
    return torch.cat(out)


_pair_groups = {}


def pair_group(placement: Placement):
    """Process group of a CFG pair. new_group is collective over the world, so every rank creates all pairs."""
    key = placement.world_size
    if key not in _pair_groups:
        _pair_groups[key] = [dist.new_group([2 * i, 2 * i + 1]) for i in range(placement.world_size // 2)]
    return _pair_groups[key][placement.rank // 2]


def exchange_cfg_halves(out_local: torch.Tensor, placement: Placement, group) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-step exchange in cfg_split mode as a library collective: a 2-rank all-gather of the raw model outputs, so
    both ranks hold (uncond, cond)."""
    bufs = [torch.empty_like(out_local), torch.empty_like(out_local)]
    dist.all_gather(bufs, out_local.contiguous(), group=group)
    return bufs[1], bufs[0]          # (uncond from the odd rank, cond from the even rank)


class PairExchange:
    """How the two ranks of a CFG pair trade their halves of the model output each step (guider.FusedDenoise calls
    exchange(out) -> (uncond, cond)). This one is the plain NCCL version."""
    graph_capturable = True
    fuses_tail = False

    def __init__(self, placement: Placement, group):
        self.placement, self.group = placement, group

    def exchange(self, out_local):
        return exchange_cfg_halves(out_local, self.placement, self.group)

    def agree(self, flag: bool) -> bool:
        """Both ranks of the pair must make the same replay-or-not decision (a rank that replays a graph while its
        partner launches step by step still meets it in the exchange, but a rank that probes while the other does
        not would not): logical AND over the pair."""
        t = torch.tensor([1 if flag else 0], device=torch.cuda.current_device(), dtype=torch.int32)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(int(t[0]))

    def key(self):
        return ("nccl", self.placement.role, self.placement.pair)


_peer_buffers: dict = {}


class PeerExchange(PairExchange):
    """The exchange fused into the tail kernel (cgs_cfg_sampler_tail_peer_fwd): each rank's kernel pushes its half of
    the model output straight into the partner's receive buffer over NVLink (symmetric memory: every rank allocates
    the same buffer and maps the partner's), flags it chunk by chunk, and consumes the partner's half as it lands --
    no collective launch between the UNet and the tail, and nothing in the step that a CUDA graph cannot hold.
    Bit-identical to the NCCL version (same kernel arithmetic on the same bf16 halves)."""
    fuses_tail = True
    graph_capturable = True
    FLAG_CAPACITY = 4096                      # chunks of 1024 latent elements: up to 4 Mi elements per exchange

    def __init__(self, placement: Placement, group):
        super().__init__(placement, group)
        self._buf = None
        self._ring = None

    def _buffers(self, n_elems: int, device):
        """Symmetric allocation [flags | slot 0 | slot 1] for exchanges of n_elems bf16 values, made once per
        (pair, size) and kept: the rendezvous is a collective, and the sequence number lives with the buffers."""
        key = (self.placement.pair, n_elems, str(device))
        buf = _peer_buffers.get(key)
        if buf is None:
            import torch.distributed._symmetric_memory as symm_mem
            flag_bytes = self.FLAG_CAPACITY * 4
            total = flag_bytes + 2 * n_elems * 2
            t = symm_mem.empty(total, dtype=torch.uint8, device=device)
            t.zero_()
            hdl = symm_mem.rendezvous(t, self.group)
            me = dist.get_rank(self.group)
            ptrs = list(hdl.buffer_ptrs)
            torch.cuda.synchronize(device)
            dist.barrier(group=self.group)              # both sides zeroed before anyone raises a flag
            buf = {"tensor": t, "handle": hdl, "seq": 0, "flag_bytes": flag_bytes,
                   "local": ptrs[me], "peer": ptrs[1 - me]}
            _peer_buffers[key] = buf
        return buf

    def next_seq(self, n_elems: int, device) -> int:
        buf = self._buffers(n_elems, device)
        s = buf["seq"]
        buf["seq"] = s + 1
        return s

    def tail_dev(self, x, out_local, sigma_vec, cfg, pred_type, old, step, denoised, x_next, sigma_data):
        """step: device fp32 row (mode, c0..c3, sequence-number bits) -- see next_seq()."""
        from . import _native
        buf = self._buffers(x.numel(), x.device)
        fb = buf["flag_bytes"]
        return _native.cfg_sampler_tail_peer(
            x, out_local, self.placement.role == "cond", buf["local"] + fb, buf["peer"] + fb, buf["local"], buf["peer"],
            self.FLAG_CAPACITY, sigma_vec, cfg, pred_type, old, step, denoised, x_next, sigma_data)

    def tail_eager(self, x, out_local, sigma_vec, cfg, pred_type, old, sampler, c0, c1, c2, c3, sigma_data):
        if self._ring is None:
            self._ring = (torch.zeros(256, 8, dtype=torch.float32).pin_memory(), [0])
        rows, pos = self._ring
        row = rows[pos[0] % rows.shape[0]]               # a pinned row per step: not rewritten while in flight
        pos[0] += 1
        row[:5] = torch.tensor([float(sampler), c0, c1, c2, c3])
        row[5:6].view(torch.int32)[0] = self.next_seq(x.numel(), x.device)
        step = row.to(x.device, non_blocking=True)
        den, xn = torch.empty_like(x), torch.empty_like(x)
        return self.tail_dev(x, out_local, sigma_vec, cfg, pred_type, x if old is None else old, step, den, xn,
                             sigma_data)

    def key(self):
        return ("peer", self.placement.role, self.placement.pair)


_split_cls = None


def split_guider_class():
    """SplitCFGGuider: a comfy.samplers.CFGGuider whose rank evaluates only its half (cond or uncond) of the UNet
    batch. Only the fused fast path is supported (plain conds, no sampler hooks); anything else must run unsplit.
    (Created on first use: the class derives from the reference's guider, which is imported only then.)"""
    global _split_cls
    if _split_cls is None:
        ensure_importable()
        import comfy.samplers

        class SplitCFGGuider(comfy.samplers.CFGGuider):
            def __init__(self, model_patcher, placement: Placement, group, exchange: Optional[PairExchange] = None):
                super().__init__(model_patcher)
                self.cgs_split = exchange if exchange is not None else PairExchange(placement, group)

            def predict_noise(self, x, timestep, model_options={}, seed=None):
                raise RuntimeError("SplitCFGGuider supports only the fused fast path (plain conds, no sampler hooks, "
                                   "euler / dpmpp_2m with integration.install() active)")
        _split_cls = SplitCFGGuider
    return _split_cls


def distributed_sample(sample_fn: Callable, noise: torch.Tensor, latent_image: torch.Tensor, *, cfg_split=None,
                       group=None, **kwargs) -> torch.Tensor:
    """Run `sample_fn(noise_slice, latent_slice, placement=..., **kwargs)` on this rank's slice and all-gather
    the final latents. Noise must be the FULL-batch tensor (generated once from the CPU generator,
    comfy/sample.py:13-15) so results are identical to the 1-GPU run for any world size."""
    if not dist.is_initialized():
        return sample_fn(noise, latent_image, placement=Placement(0, 1, "batch", 0, noise.shape[0]), **kwargs)
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    placements = [plan_placement(noise.shape[0], world, r, cfg_split) for r in range(world)]
    me = placements[rank]
    local = sample_fn(noise[me.start:me.end], latent_image[me.start:me.end], placement=me, **kwargs)
    return gather_latents(local, placements, group)


def sample_distributed(model, noise, steps, cfg, sampler_name, scheduler, positive, negative, latent_image,
                       cfg_split=None, group=None, callback=None, disable_pbar=True, seed=None, exchange_factory=None):
    """Multi-GPU form of comfy.sample.sample (one process per GPU, torch.distributed initialised): this rank
    denoises its slice of the batch -- and, in cfg_split placement, its half of the cond/uncond pair -- and every
    rank returns the full batch of final latents after one all-gather. `noise` and `latent_image` are FULL-batch
    tensors, identical on every rank. Without an initialised process group this is comfy.sample.sample."""
    ensure_importable()
    import comfy.model_management
    import comfy.sample
    import comfy.samplers

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return comfy.sample.sample(model, noise, steps, cfg, sampler_name, scheduler, positive, negative, latent_image,
                                   callback=callback, disable_pbar=disable_pbar, seed=seed)

    def run(noise_slice, latent_slice, placement):
        ks = comfy.samplers.KSampler(model, steps=steps, device=model.load_device, sampler=sampler_name,
                                     scheduler=scheduler, model_options=model.model_options)
        if placement.mode == "cfg_split":
            grp = pair_group(placement)
            ex = exchange_factory(placement, grp) if exchange_factory is not None else None
            guider = split_guider_class()(model, placement, grp, ex)
        else:
            guider = comfy.samplers.CFGGuider(model)
        guider.set_conds(positive, negative)
        guider.set_cfg(cfg)
        cb = callback if placement.rank == 0 else None
        return guider.sample(noise_slice, latent_slice, comfy.samplers.sampler_object(sampler_name), ks.sigmas,
                             callback=cb, disable_pbar=disable_pbar, seed=seed)

    out = distributed_sample(run, noise, latent_image, cfg_split=cfg_split, group=group)
    return out.to(comfy.model_management.intermediate_device())
