"""Build attention.cu with different -D tuning overrides into build/variants/*.so (here, no GPU needed), or time
them (on the GPU box).   usage: attn_variants.py build TAG=-DX=1,-DY=2 ... | attn_variants.py bench"""
import glob, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VDIR = os.path.join(ROOT, "comfy_gen_server_b200", "build", "variants")


def build(specs):
    from comfy_gen_server_b200 import build as b
    b.build()
    os.makedirs(VDIR, exist_ok=True)
    for old in glob.glob(os.path.join(VDIR, "*.so")):
        os.remove(old)
    objs = [os.path.join(b.PKG_DIR, "build", s.replace(".cu", ".o")) for s in b.SOURCES if s != "attention.cu"]
    procs = []
    for spec in specs:
        tag, _, defs = spec.partition("=")
        obj = os.path.join(VDIR, f"attention_{tag}.o")
        cmd = [b._nvcc(), *b.NVCC_FLAGS, *[d for d in defs.split(",") if d], "-I", b.CSRC, "-c",
               os.path.join(b.CSRC, "attention.cu"), "-o", obj]
        procs.append((tag, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for tag, obj, p in procs:
        out, _ = p.communicate()
        if p.returncode:
            print(out)
            raise SystemExit(f"variant {tag} failed")
        spills = [l.strip() for l in out.splitlines() if "spill" in l and "0 bytes spill stores" not in l]
        print(tag, "spilling variants:", len(spills))
        subprocess.check_call([b._nvcc(), "-shared", "-o", os.path.join(VDIR, f"libcgs_{tag}.so"), obj, *objs, "-lcudart"])


def bench():
    import torch
    from comfy_gen_server_b200 import _native as N
    from tools import kernel_bench as kb
    cases = [("L1_self_4096", 16, 10, 4096, 4096, 64), ("L2_self_1024", 16, 20, 1024, 1024, 64),
             ("cfg5_16384", 2, 10, 16384, 16384, 64), ("sd15_d40_4096", 32, 8, 4096, 4096, 40),
             ("sd15_d80_1024", 32, 8, 1024, 1024, 80), ("sd15_d160_256", 32, 8, 256, 256, 160)]
    data = {c[0]: tuple(kb.rnd((c[1], n, c[2] * c[5])) for n in (c[3], c[4], c[4])) for c in cases}
    res = {}
    for lib in sorted(glob.glob(os.path.join(VDIR, "libcgs_*.so"))):
        tag = os.path.basename(lib)[7:-3]
        N._lib = N.load_library(lib)
        row = {}
        for label, B, H, Nq, Nk, d in cases:
            q, k, v = data[label]
            ms = kb.timeit(lambda: N.attention(q, k, v, H))
            row[label] = round(4.0 * B * H * Nq * Nk * d / ms / 1e9, 1)
        res[tag] = row
        print(tag, json.dumps(row), flush=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "attn_variants.json"), "w"), indent=1)


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        bench()
