"""The allocation-free GroupNorm entry point (cgs_groupnorm_nhwc_noalloc_fwd), called straight through the C ABI.

It exists for callers that are not allowed to allocate device memory themselves (code running inside a host
framework's memory budget, C callers without an allocator, latency paths that cannot afford cudaMalloc's
synchronisation): its scratch lives in a fixed array inside the library. The test feeds it and the ordinary entry
point (caller-provided scratch) the same inputs and expects identical outputs, twice in a row with different
problems, so that whatever the first call left in the shared array is shown not to matter. Every buffer the calls
touch is created up front; the entry point itself gets only pointers.

Collected last on purpose (file name): if it fails under `-x`, everything else has already run.
"""
import ctypes

import pytest
import torch

from conftest import randn

pytestmark = pytest.mark.gpu
CGS_SCRATCH_TOO_SMALL = 5


def _problem(n, c, h, w, seed, with_addend):
    x = (randn((n, c, h, w), seed, 2.0) + 0.25).to("cuda", torch.bfloat16).contiguous(memory_format=torch.channels_last)
    e = randn((n, c), seed + 1).to("cuda", torch.bfloat16) if with_addend else None
    gamma = (randn((c,), seed + 2) * 0.2 + 1).to("cuda", torch.bfloat16)
    beta = (randn((c,), seed + 3) * 0.2).to("cuda", torch.bfloat16)
    return x, e, gamma, beta


def _ptr(t):
    return None if t is None else t.data_ptr()


def test_noalloc_groupnorm_matches_the_ordinary_entry_point_twice():
    from comfy_gen_server_b200 import _native
    lib = _native.load_library()
    stream = torch.cuda.current_stream().cuda_stream
    # two different problems back to back: the second one reuses (and partly overlaps) the first one's scratch slots
    for (n, c, h, w, silu, addend) in [(4, 640, 32, 32, 1, True), (2, 1280, 16, 16, 0, False)]:
        x, e, gamma, beta = _problem(n, c, h, w, 11 * n + c, addend)
        groups, eps = 32, 1e-5
        need = lib.cgs_groupnorm_nhwc_scratch_bytes(n, c, h * w, groups)
        scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
        y_ref = torch.empty_like(x)
        y_new = torch.empty_like(x)
        rc = lib.cgs_groupnorm_nhwc_fwd(x.data_ptr(), _ptr(e), gamma.data_ptr(), beta.data_ptr(), y_ref.data_ptr(),
                                        scratch.data_ptr(), need, n, c, h * w, groups, eps, silu, stream)
        assert rc == 0, lib.cgs_last_error()
        rc = lib.cgs_groupnorm_nhwc_noalloc_fwd(x.data_ptr(), _ptr(e), gamma.data_ptr(), beta.data_ptr(),
                                                y_new.data_ptr(), n, c, h * w, groups, eps, silu, stream)
        assert rc == 0, lib.cgs_last_error()
        torch.cuda.synchronize()
        # Same kernels, same fixed reduction order behind both entry points: the outputs are equal bit for bit.
        assert torch.equal(y_new, y_ref)
        tol = float(y_ref.float().abs().max()) * 2 ** -7 + 1e-3      # library-wide bf16 bar (tests/test_gpu_kernels.py)
        # and against a plain fp32 computation, so that "identical" cannot mean "identically wrong"
        hh = x if e is None else x + e[..., None, None]
        ref = torch.nn.functional.group_norm(hh.float(), groups, gamma.float(), beta.float(), eps)
        if silu:
            ref = torch.nn.functional.silu(ref)
        assert (y_new.float() - ref).abs().max() <= tol


def test_noalloc_groupnorm_refuses_problems_beyond_its_built_in_buffer():
    from comfy_gen_server_b200 import _native
    lib = _native.load_library()
    n, c, hw, groups = 16384, 320, 64, 32          # 16384 samples: 4 MiB of partial statistics, the buffer holds 1 MiB
    assert lib.cgs_groupnorm_nhwc_scratch_bytes(n, c, hw, groups) > (1 << 20)
    dummy = torch.zeros(16, dtype=torch.bfloat16, device="cuda")   # never dereferenced: the call must fail first
    rc = lib.cgs_groupnorm_nhwc_noalloc_fwd(dummy.data_ptr(), None, dummy.data_ptr(), dummy.data_ptr(),
                                            dummy.data_ptr(), n, c, hw, groups, 1e-5, 1,
                                            torch.cuda.current_stream().cuda_stream)
    assert rc == CGS_SCRATCH_TOO_SMALL
    assert b"built-in" in lib.cgs_last_error()
    torch.cuda.synchronize()
