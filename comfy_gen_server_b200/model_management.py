"""Device placement for the denoise path.

Single-GPU pieces mirror the reference's helpers (comfy/model_management.py:74-88, 382-449, 696-701)
as far as the hot path touches them. New here, as north_star asks: single-box multi-GPU placement --
one process per GPU (torch.distributed, NCCL over NVLink/NVSwitch), model weights replicated, the
latent batch and optionally the cond/uncond pair sharded across ranks, and ONE all-gather of the
final latents (plus a per-step pair exchange of raw model outputs when the CFG pair is split).
The reference has no multi-device code at all.
"""
from __future__ import annotations

import copy
from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


class InterruptProcessingException(Exception):
    pass


def get_torch_device():
    if not torch.cuda.is_available():
        raise RuntimeError("the B200 denoise path needs a CUDA device (there is no CPU path)")
    return torch.device("cuda", torch.cuda.current_device())


def intermediate_device():
    return torch.device("cpu")


def get_free_memory(dev=None, torch_free_too=False):
    """comfy/model_management.py:684-709 for CUDA devices."""
    if dev is None:
        dev = get_torch_device()
    dev = torch.device(dev)
    if dev.type != "cuda":
        import psutil
        free = psutil.virtual_memory().available
        return (free, free) if torch_free_too else free
    stats = torch.cuda.memory_stats(dev)
    mem_free_cuda, _ = torch.cuda.mem_get_info(dev)
    mem_free_torch = stats['reserved_bytes.all.current'] - stats['active_bytes.all.current']
    total = mem_free_cuda + mem_free_torch
    return (total, mem_free_torch) if torch_free_too else total


def prepare_mask(noise_mask, shape, device):
    """comfy/sampler_helpers.py:5-13."""
    noise_mask = torch.nn.functional.interpolate(noise_mask.reshape((-1, 1, noise_mask.shape[-2], noise_mask.shape[-1])),
                                                 size=(shape[2], shape[3]), mode="bilinear")
    noise_mask = torch.cat([noise_mask] * shape[1], dim=1)
    if noise_mask.shape[0] < shape[0]:
        noise_mask = noise_mask.repeat((shape[0] + noise_mask.shape[0] - 1) // noise_mask.shape[0], 1, 1, 1)[:shape[0]]
    return noise_mask.to(device)


class ModelPatcher:
    """The slice of comfy/model_patcher.py the hot path reads: the wrapped model, its devices, and
    `model_options` with the transformer_options patch registry (model_patcher.py:71-159)."""
    def __init__(self, model, load_device, offload_device=None, size=0):
        self.model = model
        self.load_device = torch.device(load_device)
        self.offload_device = torch.device(offload_device) if offload_device is not None else self.load_device
        self.model_options = {"transformer_options": {}}
        self.object_patches = {}

    def clone(self):
        n = ModelPatcher(self.model, self.load_device, self.offload_device)
        n.model_options = copy.deepcopy(self.model_options)
        n.object_patches = self.object_patches.copy()
        return n

    def model_dtype(self):
        return self.model.get_dtype()

    def get_model_object(self, name):
        if name in self.object_patches:
            return self.object_patches[name]
        obj = self.model
        for part in name.split("."):
            obj = getattr(obj, part)
        return obj

    # patch registry (model_patcher.py:106-159)
    def set_model_sampler_cfg_function(self, sampler_cfg_function, disable_cfg1_optimization=False):
        self.model_options["sampler_cfg_function"] = sampler_cfg_function
        if disable_cfg1_optimization:
            self.model_options["disable_cfg1_optimization"] = True

    def set_model_sampler_post_cfg_function(self, post_cfg_function, disable_cfg1_optimization=False):
        self.model_options["sampler_post_cfg_function"] = \
            self.model_options.get("sampler_post_cfg_function", []) + [post_cfg_function]
        if disable_cfg1_optimization:
            self.model_options["disable_cfg1_optimization"] = True

    def set_model_unet_function_wrapper(self, unet_wrapper_function):
        self.model_options["model_function_wrapper"] = unet_wrapper_function

    def set_model_denoise_mask_function(self, denoise_mask_function):
        self.model_options["denoise_mask_function"] = denoise_mask_function

    def set_model_patch(self, patch, name):
        to = self.model_options["transformer_options"]
        to.setdefault("patches", {})
        to["patches"][name] = to["patches"].get(name, []) + [patch]

    def set_model_patch_replace(self, patch, name, block_name, number, transformer_index=None):
        to = self.model_options["transformer_options"]
        to.setdefault("patches_replace", {}).setdefault(name, {})
        block = (block_name, number, transformer_index) if transformer_index is not None else (block_name, number)
        to["patches_replace"][name][block] = patch

    def set_model_attn1_patch(self, patch):
        self.set_model_patch(patch, "attn1_patch")

    def set_model_attn2_patch(self, patch):
        self.set_model_patch(patch, "attn2_patch")

    def set_model_attn1_replace(self, patch, block_name, number, transformer_index=None):
        self.set_model_patch_replace(patch, "attn1", block_name, number, transformer_index)

    def set_model_attn2_replace(self, patch, block_name, number, transformer_index=None):
        self.set_model_patch_replace(patch, "attn2", block_name, number, transformer_index)

    def set_model_attn1_output_patch(self, patch):
        self.set_model_patch(patch, "attn1_output_patch")

    def set_model_attn2_output_patch(self, patch):
        self.set_model_patch(patch, "attn2_output_patch")


def prepare_sampling(model_patcher: ModelPatcher, noise_shape):
    """comfy/sampler_helpers.py:60-70: make sure the model is resident on its load device."""
    model = model_patcher.model
    p = next(model.diffusion_model.parameters())
    if p.device != model_patcher.load_device:
        model.diffusion_model.to(model_patcher.load_device)
    return model, [model_patcher]


# ----------------------------------------------------------------------------------------------
# multi-GPU placement
# ----------------------------------------------------------------------------------------------
@dataclass
class Placement:
    """Which slice of the latent batch, and which CFG half, a rank owns.

    mode "batch": every rank runs both CFG halves of its own images.
    mode "cfg_split": ranks are paired (2k, 2k+1); the pair shares one image slice, rank 2k runs the
    cond half, rank 2k+1 the uncond half, and they exchange raw model outputs every step.
    """
    rank: int
    world_size: int
    mode: str
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
    """Per-step exchange in cfg_split mode: a 2-rank all-gather of the raw model outputs, so both ranks hold
    (uncond, cond) and run the identical fused tail kernel, staying in lock-step without a broadcast."""
    bufs = [torch.empty_like(out_local), torch.empty_like(out_local)]
    dist.all_gather(bufs, out_local.contiguous(), group=group)
    return bufs[1], bufs[0]          # (uncond from the odd rank, cond from the even rank)


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
