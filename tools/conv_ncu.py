"""A few convolution launches for ncu (tools: ncu --metrics ... python tools/conv_ncu.py): the SDXL shapes, one launch each."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from comfy_gen_server_b200 import _native as N
if os.environ.get("CGS_LIB"):          # a private variant build (tools/gemm_dbg.sh)
    N._lib = N.load_library(os.path.join(ROOT, os.environ["CGS_LIB"]))
if os.environ.get("CONV_VARIANT"):      # cgs_gemm_set_variant code (11 tiles of 256, 12 of 320, 13 of 192 columns)
    assert N.load_library().cgs_gemm_set_variant(int(os.environ["CONV_VARIANT"])) == 0
cl = torch.channels_last
cases = [(16, 128, 320, 320), (16, 128, 960, 320), (16, 64, 640, 640), (16, 64, 1920, 640), (16, 32, 1280, 1280)]
for n, hw, cin, cout in cases:
    x = torch.randn(n, cin, hw, hw, device="cuda").bfloat16().contiguous(memory_format=cl)
    w = (torch.randn(cout, cin, 3, 3, device="cuda") * (9 * cin) ** -0.5).bfloat16().contiguous(memory_format=cl)
    b = torch.randn(cout, device="cuda").bfloat16()
    r = torch.randn(n, cout, hw, hw, device="cuda").bfloat16().contiguous(memory_format=cl)
    for _ in range(2):
        N.conv3x3(x, w, b)
        N.conv3x3(x, w, b, r)
    torch.cuda.synchronize()
