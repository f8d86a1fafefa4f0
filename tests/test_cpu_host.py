"""CPU-only checks of the host side: the C-ABI library loads and exports every symbol declared in
include/cgs_b200.h; the product refuses CPU tensors and never touches the oracle; install() rebinds the reference
tree and uninstall() restores it; and -- with tests/fake_native.py standing in for the kernels -- the whole drop-in
(comfy.sample.sample with install() active) reproduces the reference's golden latents, falls through to the
reference's composition for hooks / non-plain conds, and shards across two gloo ranks. No compute kernels run."""
import os
import re
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, load_golden, randn, cosine


def test_cabi_library_exports_every_declared_symbol():
    from comfy_gen_server_b200 import _native
    header = open(os.path.join(ROOT, "include", "cgs_b200.h")).read()
    declared = set(re.findall(r"\b(cgs_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_native.EXPORTED_SYMBOLS)
    lib = _native.load_library()
    for sym in declared:
        assert hasattr(lib, sym), sym
    assert lib.cgs_abi_version() == 1


def test_product_refuses_cpu_tensors():
    from comfy_gen_server_b200 import _native, layers
    q = torch.zeros(1, 8, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        _native.attention(q, q, q, 1)
    with pytest.raises(RuntimeError):
        layers.attention_b200(q, q, q, 1)
    with pytest.raises(RuntimeError):
        _native.layer_norm(q, q[0, 0], q[0, 0])
    with pytest.raises(RuntimeError):
        _native.geglu(q, torch.zeros(128, 64, dtype=torch.bfloat16), None)
    with pytest.raises(RuntimeError):
        _native.cfg_denoise(torch.zeros(1, 4, 8, 8), None, torch.zeros(1, 4, 8, 8), torch.ones(1), 1.0, 0)
    with pytest.raises(_native.NativeLibraryError):
        _native.load_library("/nonexistent/libcgs_b200.so")


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "comfy_gen_server_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


def test_product_declares_no_copy_of_a_reference_class():
    """The host side is a plug-in: it binds functions onto the reference's classes and must not re-declare them."""
    pkg = os.path.join(ROOT, "comfy_gen_server_b200")
    forbidden = ("CrossAttention", "BasicTransformerBlock", "SpatialTransformer", "UNetModel", "ResBlock", "KSampler",
                 "KSAMPLER", "ModelPatcher", "BaseModel", "ModelSamplingDiscrete", "CONDRegular", "CONDCrossAttn")
    for f in os.listdir(pkg):
        if f.endswith(".py"):
            src = open(os.path.join(pkg, f)).read()
            for name in forbidden:
                assert not re.search(rf"^class\s+{name}\b", src, re.M), (f, name)


def test_dpmpp2m_coefficients_reproduce_reference_update():
    """x' = c_x x - c_d (c_cur D - c_old D_old) with host coefficients == the reference loop, bit for bit."""
    from comfy_gen_server_b200 import sampling as ks
    g = load_golden("sampler_kat")
    key = "dpmpp_2m_karras_30"
    sig = g[key + ".sigmas"]
    x = g[key + ".x0"].clone()
    old = None
    for i in range(len(sig) - 1):
        d = x / (1.0 + sig[i] ** 2)
        mode, cx, cd, cc, co = ks.dpmpp_2m_coefficients(sig, i, old is not None)
        f = lambda v: torch.tensor(v, dtype=torch.float32)
        dd = d if mode == 2 else f(cc) * d - f(co) * old
        x = f(cx) * x - f(cd) * dd
        old = d
    assert torch.equal(x, g[key + ".final"])


def test_placement_plans():
    from comfy_gen_server_b200 import placement as mm
    # config 3: SDXL batch 8 on 1/2/4/8 GPUs -> pure batch sharding
    for world in (1, 2, 4, 8):
        ps = [mm.plan_placement(8, world, r) for r in range(world)]
        assert all(p.mode == "batch" for p in ps)
        assert [p.start for p in ps] == [r * 8 // world for r in range(world)] and ps[-1].end == 8
    # config 5: batch 4 on 8 GPUs -> 4 image shards x 2 CFG halves
    ps = [mm.plan_placement(4, 8, r) for r in range(8)]
    assert all(p.mode == "cfg_split" for p in ps)
    assert [(p.start, p.end, p.role) for p in ps[:4]] == [(0, 1, "cond"), (0, 1, "uncond"), (1, 2, "cond"), (1, 2, "uncond")]
    # config 4 on 2 GPUs with the CFG pair split on request: both ranks hold all 8 images
    ps = [mm.plan_placement(8, 2, r, cfg_split=True) for r in range(2)]
    assert [(p.start, p.end, p.role) for p in ps] == [(0, 8, "cond"), (0, 8, "uncond")]
    assert mm.shard_range(10, 4, 0) == (0, 3) and mm.shard_range(10, 4, 3) == (8, 10)
    with pytest.raises(ValueError):
        mm.plan_placement(4, 3, 0, cfg_split=True)


# ----------------------------------------------------------------------------------------------
# the drop-in, end to end, with torch stand-ins for the kernels
# ----------------------------------------------------------------------------------------------
def test_install_rebinds_the_reference_tree_and_uninstall_restores_it(reference_tree):
    import comfy.cli_args
    import comfy.k_diffusion.sampling as rk
    import comfy.ldm.modules.attention as ra
    import comfy.ldm.modules.diffusionmodules.openaimodel as ru
    import comfy.model_base as rb
    import comfy.model_management as rm
    import comfy.ops as ro
    import comfy.samplers as rs
    from comfy_gen_server_b200 import integration as b200, layers
    targets = lambda: (ra.optimized_attention, ra.optimized_attention_masked, ra.GEGLU.forward,
                       ro.disable_weight_init.LayerNorm.forward, ro.disable_weight_init.GroupNorm.forward,
                       ra.BasicTransformerBlock._forward, ra.SpatialTransformer.forward, ru.ResBlock._forward,
                       ru.Upsample.forward, ru.UNetModel.forward, rb.BaseModel.apply_model, rk.sample_euler,
                       rk.sample_dpmpp_2m, comfy.cli_args.args.bf16_unet, hasattr(rm, "plan_placement"))
    orig = targets()
    log = b200.install()
    try:
        assert b200.installed() and len(log) >= 14
        assert ra.optimized_attention is layers.attention_b200 and ra.optimized_attention_masked is layers.attention_b200
        assert all(a is not b for a, b in zip(orig[:13], targets()[:13]))
        assert comfy.cli_args.args.bf16_unet is True and rm.plan_placement(8, 2, 1).start == 4
        assert rs.ksampler("dpmpp_2m").sampler_function is rk.sample_dpmpp_2m    # getattr lookup sees the new function
        assert rs.ksampler("euler").sampler_function.__module__ == "comfy_gen_server_b200.sampling"
        # manual_cast's classes inherit the rebound methods; the CPU-capable selector is untouched
        assert ro.manual_cast.LayerNorm.forward is ro.disable_weight_init.LayerNorm.forward
        assert ra.optimized_attention_for_device(torch.device("cpu")) is ra.attention_sub_quad
        assert b200.install() == []                                               # idempotent
    finally:
        b200.uninstall()
    assert targets() == orig and not b200.installed() and not layers.ORIGINAL


def test_installed_ops_leave_other_callers_alone(reference_tree, installed):
    """With install() active a CPU / fp32 caller of the reference's op classes (the text encoder, the VAE) still gets
    the reference's own computation."""
    import comfy.ops
    ln = comfy.ops.disable_weight_init.LayerNorm(32)
    gn = comfy.ops.manual_cast.GroupNorm(4, 32)
    with torch.no_grad():
        for p in list(ln.parameters()) + list(gn.parameters()):
            p.copy_(randn(tuple(p.shape), 3) * 0.1 + 1.0)
    x = randn((2, 5, 32), 1)
    assert torch.equal(ln(x), torch.nn.functional.layer_norm(x, (32,), ln.weight, ln.bias, ln.eps))
    img = randn((2, 32, 4, 4), 2)
    assert torch.allclose(gn(img), torch.nn.functional.group_norm(img, 4, gn.weight, gn.bias, gn.eps))


E2E = [("tiny_sdxl", "e2e_tiny_sdxl_dpmpp2m", 16, 30, "dpmpp_2m", "karras"),
       ("tiny_sd15", "e2e_tiny_sd15_euler", 16, 20, "euler", "normal"),
       ("tiny_sd21v", "e2e_tiny_sd21v_euler", 12, 25, "euler", "normal")]


def _cpu_case(family, golden, hw):
    import comfy.sample
    from comfy_gen_server_b200 import models
    g = load_golden(golden)
    seed = int(g["seed"])
    patcher, model = models.build_model(family, "cpu", seed=seed, dtype=torch.float32)
    patcher.model_options["cuda_graph"] = False
    latent = torch.zeros(2, 4, hw, hw)
    noise = comfy.sample.prepare_noise(latent, seed)
    extra = lambda k: {"pooled_output": g[k]} if k in g else {}
    return g, seed, patcher, latent, noise, [[g["positive"], extra("pooled_pos")]], [[g["negative"], extra("pooled_neg")]]


@pytest.mark.parametrize("family,golden,hw,steps,sampler,sched", E2E)
def test_drop_in_reproduces_reference_latents_with_torch_stand_ins(reference_tree, family, golden, hw, steps, sampler,
                                                                   sched):
    """comfy.sample.sample with install() active -- fused blocks, NHWC-capable UNet path, fused CFG + sampler step --
    on fp32 CPU tensors with torch stand-ins for the kernels: equal to the reference's golden latents to fp32
    rounding, i.e. the host-side plumbing (cond order, K/V caching, bias folding, coefficients) changes no result."""
    import comfy.sample
    import fake_native
    from comfy_gen_server_b200 import integration
    g, seed, patcher, latent, noise, pos, neg = _cpu_case(family, golden, hw)
    with fake_native.patched() as calls, torch.inference_mode():
        integration.install()
        try:
            out = comfy.sample.sample(patcher, noise, steps, 7.0, sampler, sched, pos, neg, latent, seed=seed,
                                      disable_pbar=True)
        finally:
            integration.uninstall()
    assert calls["cfg_sampler_tail"] == steps and calls["add_layer_norm_"] > 0 and calls["geglu"] > 0
    assert (out - g["final"]).abs().max() < 2e-3 and cosine(out, g["final"]) > 0.999999


def test_drop_in_with_every_linear_on_the_library_gemm(reference_tree):
    """The large-batch routing of the transformer block (fused QKV, to_out + residual and ff.net.2 + residual on the
    library GEMM, plain LayerNorm after them) forced on for the tiny model: same latents as the reference's golden."""
    import comfy.sample
    import fake_native
    from comfy_gen_server_b200 import block, integration
    g, seed, patcher, latent, noise, pos, neg = _cpu_case("tiny_sdxl", "e2e_tiny_sdxl_dpmpp2m", 16)
    saved = block.NATIVE_LINEAR_MIN_ROWS
    with fake_native.patched() as calls, torch.inference_mode():
        integration.install()
        block.NATIVE_LINEAR_MIN_ROWS = 0
        try:
            out = comfy.sample.sample(patcher, noise, 30, 7.0, "dpmpp_2m", "karras", pos, neg, latent, seed=seed,
                                      disable_pbar=True)
        finally:
            block.NATIVE_LINEAR_MIN_ROWS = saved
            integration.uninstall()
    assert calls.get("add_layer_norm_", 0) == 0 and calls["linear"] > 0 and calls["layer_norm"] > 0
    assert (out - g["final"]).abs().max() < 2e-3 and cosine(out, g["final"]) > 0.999999


def test_hooks_and_non_plain_conds_fall_through_to_the_reference_composition(reference_tree):
    """(i) a post-CFG hook, (ii) an attention output patch, (iii) two positive conds with an area and a strength: the
    fused step / fused block must stand aside for EVERY step and the result must equal the unmodified reference's."""
    import comfy.sample
    import fake_native
    from comfy_gen_server_b200 import integration
    g, seed, patcher, latent, noise, pos, neg = _cpu_case("tiny_sd15", "e2e_tiny_sd15_euler", 16)
    steps = 3
    run = lambda p, po=pos: comfy.sample.sample(p, noise, steps, 7.0, "euler", "normal", po, neg, latent, seed=seed,
                                                disable_pbar=True)
    hooked = patcher.clone()
    post_calls = []
    hooked.set_model_sampler_post_cfg_function(lambda a: (post_calls.append(1), a["denoised"] * 0.98)[1])
    patched_attn = patcher.clone()
    attn_calls = []
    patched_attn.set_model_attn1_output_patch(lambda n, extra: (attn_calls.append(extra["block"]), n * 1.01)[1])
    area_pos = [pos[0], [randn(tuple(g["positive"].shape), 77), {"area": (8, 16, 4, 0), "strength": 0.7}]]
    with torch.inference_mode():
        want = [run(hooked), run(patched_attn), run(patcher, area_pos)]
        n_post, n_attn = len(post_calls), len(attn_calls)
        with fake_native.patched() as calls:
            integration.install()
            try:
                got = [run(hooked), run(patched_attn), run(patcher, area_pos)]
                tails = calls.get("cfg_sampler_tail", 0)
            finally:
                integration.uninstall()
    assert len(post_calls) == 2 * n_post == 2 * steps and len(attn_calls) == 2 * n_attn
    # (ii) keeps the fused CFG + update step (the patch lives inside the UNet); (i) and (iii) must not use it
    assert tails == steps
    for a, b in zip(got, want):
        assert (a - b).abs().max() < 2e-3, float((a - b).abs().max())


# ----------------------------------------------------------------------------------------------
# two gloo ranks
# ----------------------------------------------------------------------------------------------
def _dist_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from comfy_gen_server_b200 import placement as mm
    try:
        full_noise = torch.arange(5 * 4 * 2 * 2, dtype=torch.float32).view(5, 4, 2, 2)
        latent = torch.zeros_like(full_noise)

        # batch sharding (ragged: 5 images over 2 ranks) + the single all-gather of final latents
        def fake_sample(n, l, placement):
            return n * 2.0 + placement.start * 0.0          # sample-independent function of the slice
        out = mm.distributed_sample(fake_sample, full_noise, latent)
        ok1 = torch.equal(out, full_noise * 2.0)

        # CFG split: both ranks hold all images, exchange halves every "step", run the same combine
        def split_sample(n, l, placement):
            grp = mm.pair_group(placement)
            x = n.clone()
            for step in range(3):
                local = x * (1.0 if placement.role == "cond" else -1.0) + step
                out_u, out_c = mm.exchange_cfg_halves(local, placement, grp)
                x = out_u + (out_c - out_u) * 7.0
            return x
        out2 = mm.distributed_sample(split_sample, full_noise, latent, cfg_split=True)
        x = full_noise.clone()
        for step in range(3):
            u, c = -x + step, x + step
            x = u + (c - u) * 7.0
        ok2 = torch.equal(out2, x)

        # the real thing on two ranks: sample_distributed through the reference's guider + install(), batch-sharded
        # and CFG-split, against the 1-process comfy.sample.sample (torch stand-ins for the kernels)
        from comfy_gen_server_b200.reference import ensure_importable
        ensure_importable(cpu=True)
        import comfy.sample
        import comfy.utils
        import fake_native
        from comfy_gen_server_b200 import integration, models
        comfy.utils.PROGRESS_BAR_ENABLED = False
        torch.set_num_threads(2)
        seed = 44
        patcher, _ = models.build_model("tiny_sdxl", "cpu", seed=seed, dtype=torch.float32)
        patcher.model_options["cuda_graph"] = False
        g = lambda shape, s: torch.randn(shape, generator=torch.Generator().manual_seed(s))
        lat = torch.zeros(3, 4, 16, 16)
        noise = comfy.sample.prepare_noise(lat, seed)
        pos = [[g((1, 77, 128), 1), {"pooled_output": g((1, 1280), 3)}]]
        neg = [[g((1, 154, 128), 2), {"pooled_output": g((1, 1280), 4)}]]
        with fake_native.patched(), torch.inference_mode():
            integration.install()
            ref = comfy.sample.sample(patcher, noise, 3, 7.0, "dpmpp_2m", "karras", pos, neg, lat, seed=seed,
                                      disable_pbar=True)
            a = mm.sample_distributed(patcher, noise, 3, 7.0, "dpmpp_2m", "karras", pos, neg, lat, seed=seed)
            b = mm.sample_distributed(patcher, noise, 3, 7.0, "dpmpp_2m", "karras", pos, neg, lat, seed=seed,
                                      cfg_split=True)
            integration.uninstall()
        # (a rank with ONE image takes other BLAS kernels than the batched run: fp32 rounding, amplified by CFG 7)
        tol = 1e-3 * float(ref.abs().max())
        ok3 = float((a - ref).abs().max()) < tol and a.shape == ref.shape
        ok4 = float((b - ref).abs().max()) < tol and b.shape == ref.shape
        q.put((rank, ok1, ok2, ok3, ok4))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharding_gather_and_cfg_exchange():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_dist_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=600) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True, True, True, True), (1, True, True, True, True)]


def test_reference_param_shapes_match_unet_config(reference_tree):
    """unet_config.param_shapes (the seeded-weight layout shared by the oracle, the goldens and the GPU tests) equals
    the state dict of the reference's UNetModel for the tiny families (models.build_model asserts it for any family)."""
    from comfy_gen_server_b200 import models
    for fam in ("tiny_sd15", "tiny_sd21v", "tiny_sdxl"):
        models.build_model(fam, "cpu", seed=1, dtype=torch.float32)


def test_upsample_conv_phase_weights_are_the_same_convolution():
    """The algebra behind cgs_upsample2x_conv3x3_nhwc_fwd, checked on the CPU in float64: conv3x3(nearest 2x(x)) ==
    the four 2x2 phase convolutions of x with the pre-summed weights, interleaved."""
    import torch.nn.functional as F
    from comfy_gen_server_b200 import _native
    g = torch.Generator().manual_seed(3)
    x = torch.randn(2, 5, 6, 7, generator=g, dtype=torch.float64)
    w = torch.randn(4, 5, 3, 3, generator=g, dtype=torch.float64)
    ref = F.conv2d(F.interpolate(x, scale_factor=2, mode="nearest"), w, padding=1)
    w4 = _native.upsample2x_conv_weights(w)                      # [4, Cout, 2, 2, Cin]
    assert w4.shape == (4, 4, 2, 2, 5)
    out = torch.zeros_like(ref)
    for a in range(2):
        for b in range(2):
            k = w4[a * 2 + b].permute(0, 3, 1, 2).double()       # [Cout, Cin, 2, 2]
            # tap (dr, ds) reads input pixel (i + a - 1 + dr, j + b - 1 + ds): pad so that index 0 is offset a - 1
            xp = F.pad(x, (1 - b, b, 1 - a, a))
            out[:, :, a::2, b::2] = F.conv2d(xp, k)
    assert (out - ref).abs().max() < 1e-5          # (the weights are summed in fp32)


@pytest.mark.parametrize("created_in_inference", [False, True])
@pytest.mark.parametrize("converted_in_inference", [False, True])
@pytest.mark.parametrize("used_in_inference", [False, True])
def test_channels_last_weight_conversion_keeps_parameters_usable(created_in_inference, converted_in_inference,
                                                                 used_in_inference):
    """The in-place channels-last conversion of convolution weights must leave a parameter as it found it: the
    reference's executor loads checkpoints and samples under torch.inference_mode() (execution.py:537), tests and
    scripts do not."""
    from comfy_gen_server_b200 import unet_path, vae_path
    for convert in (unet_path._ensure_channels_last_weights, lambda m: vae_path._channels_last(m, torch.zeros(1, 8, 4, 4))):
        with torch.inference_mode(created_in_inference):
            conv = torch.nn.Conv2d(8, 8, 3, padding=1)
        with torch.inference_mode(converted_in_inference):
            convert(conv)
        assert conv.weight.is_contiguous(memory_format=torch.channels_last)
        assert conv.weight.is_inference() == created_in_inference
        with torch.inference_mode(used_in_inference), torch.no_grad():
            assert conv(torch.randn(1, 8, 4, 4)).shape == (1, 8, 4, 4)


def _fill(module, seed):
    """Seeded weights for a module built with the reference's no-init operator classes."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in module.named_parameters():
            if p.dim() == 1 and ("norm" in name or name.endswith("in_layers.0.weight") or name.endswith("out_layers.0.weight")
                                 or name.endswith("out.0.weight")) and name.endswith("weight"):
                p.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
            else:
                p.copy_(torch.randn(p.shape, generator=g) * (0.1 if p.dim() == 1 else p[0].numel() ** -0.5))


def test_channels_last_unet_path_on_the_cpu_equals_the_reference(reference_tree):
    """The wide-UNet host logic that the tiny golden models never reach (model_channels / 32 >= 8): channels-last
    pipeline, 3x3 convolutions with bias / skip image folded into them, interpolate + conv, deferred up-sampling conv
    bias in the skip concatenation -- on fp32 CPU tensors with torch stand-ins for the kernels, against the reference's
    own UNetModel.forward of the same module."""
    import comfy.ops
    import comfy.ldm.modules.diffusionmodules.openaimodel as ref_unet
    import fake_native
    from comfy_gen_server_b200 import integration
    cfg = {'use_checkpoint': False, 'image_size': 32, 'out_channels': 4, 'use_spatial_transformer': True, 'legacy': False,
           'adm_in_channels': None, 'in_channels': 4, 'model_channels': 256, 'num_res_blocks': [1, 1],
           'transformer_depth': [1, 1], 'channel_mult': [1, 2], 'transformer_depth_middle': 1,
           'use_linear_in_transformer': True, 'context_dim': 64, 'num_head_channels': 64,
           'transformer_depth_output': [1, 1, 1, 1], 'use_temporal_attention': False, 'use_temporal_resblock': False}
    unet = ref_unet.UNetModel(**cfg, device="cpu", dtype=torch.float32, operations=comfy.ops.disable_weight_init).eval()
    _fill(unet, 21)
    x, t, ctx = randn((2, 4, 32, 32), 1), torch.tensor([500.0, 20.0]), randn((2, 77, 64), 2)
    with torch.inference_mode():
        want = unet(x, t, context=ctx, transformer_options={})
        with fake_native.patched() as calls:
            integration.install()
            try:
                got = unet(x, t, context=ctx, transformer_options={})
                again = unet(x, t, context=ctx, transformer_options={})      # weights now channels-last, caches warm
            finally:
                integration.uninstall()
        after = unet(x, t, context=ctx, transformer_options={})            # the reference again, on the converted weights
    assert calls["conv3x3"] >= 8 and calls.get("upsample2x_conv3x3", 0) >= 2 and calls["cat_channels"] >= 8, dict(calls)
    assert unet.input_blocks[1][0].in_layers[2].weight.is_contiguous(memory_format=torch.channels_last)
    scale = float(want.abs().max())
    assert (got - want).abs().max() < 2e-4 * scale and (again - want).abs().max() < 2e-4 * scale
    assert (after - want).abs().max() < 2e-4 * scale


def test_vae_decoder_path_on_the_cpu_equals_the_reference(reference_tree):
    """vae_path's host logic (channels-last conversion, bias / `x + h` folded into the convolutions, interpolate + conv,
    the mid-block attention as q k^T -> row softmax -> . v on the GEMM entry points) with torch stand-ins on the CPU,
    against the reference's own Decoder.forward."""
    import comfy.ldm.modules.diffusionmodules.model as vm
    import fake_native
    from comfy_gen_server_b200 import integration
    dd = {'double_z': True, 'z_channels': 4, 'resolution': 32, 'in_channels': 3, 'out_ch': 3, 'ch': 32,
          'ch_mult': [1, 2], 'num_res_blocks': 1, 'attn_resolutions': [], 'dropout': 0.0}
    dec = vm.Decoder(**dd).eval()
    _fill(dec, 22)
    z = randn((2, 4, 16, 16), 3)
    with torch.inference_mode():
        want = dec(z)
        with fake_native.patched() as calls:
            integration.install()
            try:
                got = dec(z)
            finally:
                integration.uninstall()
    assert calls["conv3x3"] >= 6 and calls["upsample2x_conv3x3"] == 1 and calls["softmax_rows"] == 2, dict(calls)
    assert calls["linear"] == 1 + 2 * 3, dict(calls)              # one fused q / k / v; per image: scores, . v, proj_out (+ x)
    assert (got - want).abs().max() < 2e-4 * float(want.abs().max())
