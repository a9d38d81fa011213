"""In-tree build of the two shared objects of the package (no JIT cache, no pip):

  libb200_w8a8.so      nvcc, sm_100a only: kernels + the C ABI of include/b200_w8a8.h
  b200_w8a8_torch.so   g++: TORCH_LIBRARY_FRAGMENT(sgl_kernel, ...) shim over the C ABI

Both land next to this file so that they travel with the repo snapshot to the GPU box.
nvcc cross-compiles without a GPU.  Run: python sglang-diffusion_b200/build.py [--force] [--verbose]
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
import time

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
INCLUDE = os.path.join(ROOT, "include")
OBJ_DIR = os.path.join(PKG, "_obj")
LIB_SO = os.path.join(PKG, "libb200_w8a8.so")
TORCH_SO = os.path.join(PKG, "b200_w8a8_torch.so")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")
GXX = "/usr/bin/g++"

CU_SOURCES = ["capi.cu", "quant.cu", "quant_fused.cu", "moe.cu", "gemm_sm100.cu", "gemm_splitk_sm100.cu", "gemm_block128_sm100.cu", "allreduce_symm.cu"]
CU_HEADERS = ["ptx.cuh", "quant_common.cuh", "gemm_common.cuh", "b200_w8a8_internal.h"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
    "-I" + INCLUDE, "-I" + CSRC,
]


def _mtime(p):
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def _newest(paths):
    return max(_mtime(p) for p in paths)


def build_cuda(force=False, verbose=False) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    deps = [os.path.join(CSRC, h) for h in CU_HEADERS] + [os.path.join(INCLUDE, "b200_w8a8.h"), os.path.abspath(__file__)]
    dep_t = _newest(deps)
    procs, objs = [], []
    for s in CU_SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ_DIR, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _mtime(obj) < max(_mtime(src), dep_t):
            cmd = [NVCC, "-c", src, "-o", obj] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
            procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"---- nvcc {s} FAILED ----\n{out}\n")
        elif verbose or "warning" in out:
            sys.stderr.write(f"---- nvcc {s} ----\n{out}\n")
    if failed:
        raise RuntimeError("nvcc compilation failed")
    if force or procs or _mtime(LIB_SO) < _newest(objs):
        # only the extern "C" b2_* API is exported (version script hides everything else)
        vs = os.path.join(OBJ_DIR, "exports.map")
        with open(vs, "w") as f:
            f.write("{ global: b2_*; local: *; };\n")
        cmd = [NVCC, "-shared", "-o", LIB_SO] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
               "-Xlinker", "--version-script=" + vs]
        subprocess.check_call(cmd)
    return LIB_SO


def build_torch_shim(force=False, verbose=False) -> str:
    import torch  # noqa: F401
    from torch.utils import cpp_extension as ce

    src = os.path.join(CSRC, "torch_ops.cpp")
    deps = [src, os.path.join(INCLUDE, "b200_w8a8.h"), LIB_SO, os.path.abspath(__file__)]
    if not force and _mtime(TORCH_SO) >= _newest(deps):
        return TORCH_SO
    tlib = os.path.join(os.path.dirname(torch.__file__), "lib")
    inc = ["-I" + p for p in ce.include_paths()] + ["-I" + INCLUDE, "-I" + os.path.join(CUDA_HOME, "include"),
                                                    "-I" + sysconfig.get_paths()["include"]]
    cmd = [GXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-D_GLIBCXX_USE_CXX11_ABI=1", src, "-o", TORCH_SO] + inc + [
        "-L" + PKG, "-l:libb200_w8a8.so", "-L" + tlib, "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
        "-Wl,-rpath,$ORIGIN", "-Wl,-rpath," + tlib, "-Wl,--no-as-needed",
    ]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return TORCH_SO


def build_all(force=False, verbose=False):
    t0 = time.time()
    a = build_cuda(force, verbose)
    b = build_torch_shim(force, verbose)
    if verbose:
        print(f"[b200_w8a8] built {a}, {b} in {time.time() - t0:.0f}s", file=sys.stderr)
    return a, b


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv)
