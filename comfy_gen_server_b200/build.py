"""Build the C-ABI CUDA library (libcgs_b200.so) in-tree with nvcc for sm_100a.

The library has no dependency on torch or libcuda at link time: launchers take raw pointers,
sizes, strides and a cudaStream_t (as void*), and return an int status (include/cgs_b200.h).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libcgs_b200.so")
STAMP_PATH = os.path.join(PKG_DIR, ".libcgs_b200.stamp")

SOURCES = ["common.cu", "layernorm.cu", "sampler_tail.cu", "geglu_gemm.cu", "attention.cu", "groupnorm.cu", "transpose.cu",
           "softmax.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]
if os.environ.get("CGS_BUILD_TRACE") == "1":      # attention timeline stamps (tools/attn_trace.py); slower kernel
    NVCC_FLAGS.append("-DCGS_ATTN_TRACING=1")


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA extension cannot be built")


def _sources() -> list[str]:
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _digest() -> str:
    h = hashlib.sha256()
    for name in sorted(os.listdir(CSRC)):
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(CSRC, name), "rb") as f:
                h.update(name.encode())
                h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH) or not os.path.exists(STAMP_PATH):
        return True
    with open(STAMP_PATH) as f:
        return f.read().strip() != _digest()


def build_variant(lib_path: str, defines: list[str], sources: list[str] | None = None) -> str:
    """A private copy of the library with extra -D switches (tuning / timing experiments: tools/attn_variants.py,
    tools/gemm_dbg.sh). Never the product library: _native.load_library(path) loads it explicitly."""
    nvcc = _nvcc()
    obj_dir = lib_path + ".obj"
    os.makedirs(obj_dir, exist_ok=True)
    procs, objs = [], []
    for src in _sources():
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        objs.append(obj)
        flags = [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")]
        extra = defines if (sources is None or src in sources) else []
        base_obj = os.path.join(PKG_DIR, "build", src.replace(".cu", ".o"))
        if not extra and os.path.exists(base_obj) and not needs_build():
            objs[-1] = base_obj                                   # unchanged translation unit: reuse the product object
            continue
        procs.append((src, subprocess.Popen([nvcc, *flags, *extra, "-I", CSRC, "-c", os.path.join(CSRC, src), "-o", obj],
                                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed for variant {lib_path} ({src})")
    r = subprocess.run([nvcc, "-shared", "-o", lib_path, *objs, "-lcudart"], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed")
    return lib_path


def build(force: bool = False, verbose: bool = True) -> str:
    """Compile every .cu into objects (in parallel) and link the shared library."""
    if not force and not needs_build():
        return LIB_PATH
    nvcc = _nvcc()
    obj_dir = os.path.join(PKG_DIR, "build")
    os.makedirs(obj_dir, exist_ok=True)
    procs = []
    objs = []
    for src in _sources():
        obj = os.path.join(obj_dir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc, *NVCC_FLAGS, "-I", CSRC, "-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log_lines = []
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        log_lines.append(f"==== {src} (exit {p.returncode})\n{out}")
        failed |= p.returncode != 0
    log = "\n".join(log_lines)
    with open(os.path.join(obj_dir, "nvcc.log"), "w") as f:
        f.write(log)
    if failed:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed; see output above")
    link = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-lcudart"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("link failed")
    with open(STAMP_PATH, "w") as f:
        f.write(_digest())
    if verbose:
        print(f"built {LIB_PATH}")
    return LIB_PATH


if __name__ == "__main__":
    build(force="--force" in sys.argv)
