"""Timeline of CTA 0 of the convolution GEMM (clock64 stamps per work item; tracing build only):
   python -c "from comfy_gen_server_b200 import build as b; b.build_variant('build_variants/libcgs_gemm_trace.so', ['-DCGS_GEMM_TRACE'], ['geglu_gemm.cu'])"
   CGS_LIB=build_variants/libcgs_gemm_trace.so python tools/gemm_trace.py"""
import ctypes, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from comfy_gen_server_b200 import _native as N
N._lib = N.load_library(os.path.join(ROOT, os.environ["CGS_LIB"]))
lib = N._lib
cl = torch.channels_last
buf = torch.zeros(32 * 16, dtype=torch.int64, device="cuda")
lib.cgs_gemm_set_trace.argtypes = [ctypes.c_void_p]
for n, hw, cin, cout in [(16, 128, 320, 320), (16, 64, 640, 640), (16, 32, 1280, 1280)]:
    x = torch.randn(n, cin, hw, hw, device="cuda").bfloat16().contiguous(memory_format=cl)
    w = (torch.randn(cout, cin, 3, 3, device="cuda") * (9 * cin) ** -0.5).bfloat16().contiguous(memory_format=cl)
    b = torch.randn(cout, device="cuda").bfloat16()
    for _ in range(3):
        buf.zero_()
        assert lib.cgs_gemm_set_trace(buf.data_ptr()) == 0
        N.conv3x3(x, w, b)
    torch.cuda.synchronize()
    t = buf.cpu().view(32, 16)
    t0 = int(t[0, 0])
    print(f"conv {cin}->{cout} @ {hw}^2: item | mma: wait_tempty tempty_ok first_full last_commit | epi: wait_tfull tfull_ok ld0 ld1 ld2 item_end   (clocks since the first stamp)")
    for i in range(0, 6):
        row = [(int(v) - t0) if v > 0 else -1 for v in t[i, :11]]
        print(f"  {i:2d} | " + " ".join(f"{v:8d}" for v in row[:4]) + " | " + " ".join(f"{v:8d}" for v in row[4:11]))
        p0 = [(int(v) - int(t[i, 6])) if v > 0 else -1 for v in t[i, 11:16]]
        print("       pass 0 after its tcgen05.ld: computed, store-read waited, barrier 1, tile written, barrier 2: " + " ".join(f"{v:6d}" for v in p0))

for M, K, Nn, res in [(16384, 1280, 1280, True), (16384, 1280, 1280, False), (16384, 5120, 1280, True), (65536, 640, 640, True)]:
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(Nn, K, device="cuda") * K ** -0.5).bfloat16()
    b = torch.randn(Nn, device="cuda").bfloat16()
    r = torch.randn(M, Nn, device="cuda").bfloat16() if res else None
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for _ in range(3):
        buf.zero_()
        assert lib.cgs_gemm_set_trace(buf.data_ptr()) == 0
        ev[0].record()
        N.linear(x, w, b, r)
        ev[1].record()
    torch.cuda.synchronize()
    t = buf.cpu().view(32, 16)
    t0 = int(t[0, 0])
    print(f"linear M={M} K={K} N={Nn} residual={res}: {ev[0].elapsed_time(ev[1]) * 1e3:.1f} us; item | mma: wait_tempty tempty_ok first_full last_commit | epi: wait_tfull tfull_ok ld0 ld1 ld2 item_end")
    for i in range(0, 7):
        row = [(int(v) - t0) if v > 0 else -1 for v in t[i, :11]]
        print(f"  {i:2d} | " + " ".join(f"{v:8d}" for v in row[:4]) + " | " + " ".join(f"{v:8d}" for v in row[4:11]))
        p0 = [(int(v) - int(t[i, 6])) if v > 0 else -1 for v in t[i, 11:16]]
        print("       pass 0 after its tcgen05.ld: computed, store-read waited, barrier 1, tile written, barrier 2: " + " ".join(f"{v:6d}" for v in p0))
