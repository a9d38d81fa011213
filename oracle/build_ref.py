"""TEST INFRASTRUCTURE ONLY -- builds the *reference's own* CPU path as the parity oracle.

Compiles, where they lie under /root/reference, the three files that hold the
reference CPU implementation of the W8A8 linear path

    sgl-kernel/csrc/cpu/gemm.cpp       (convert_weight_packed, pack_vnni, s8s8 compensation)
    sgl-kernel/csrc/cpu/gemm_int8.cpp  (per_token_quant_int8_cpu, int8_scaled_mm_cpu, int8_scaled_mm_with_quant)
    sgl-kernel/csrc/cpu/gemm_fp8.cpp   (fp8_scaled_mm_cpu)

together with oracle/ref_shim.cpp (our own 60-line registration file) into
oracle/_ref/sgl_cpu_oracle.so.  No reference source is copied into this repo and
the reference's own build system (CMake + libnuma) is not run.  Flags follow
sgl-kernel/csrc/cpu/CMakeLists.txt:77-91 (-O3 -fopenmp, SGLANG_CPU_FP8_CVT_FTZ);
`-march=native` is replaced by `-march=sapphirerapids` so that the object does
not depend on which AVX512/AMX host compiled it (vec.h:3-5 needs AVX512F +
AVX512BF16 + AMX_BF16 or every kernel body is a TORCH_CHECK(false) stub).

The .so only travels to the GPU box as a prebuilt file (oracle/_ref/ is
git-ignored but not gpurun-ignored); /root/reference does not exist there.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_CPU = "/root/reference/sgl-kernel/csrc/cpu"
OUT_DIR = os.path.join(HERE, "_ref")
OUT_SO = os.path.join(OUT_DIR, "sgl_cpu_oracle.so")
GXX = "/usr/bin/g++"  # /opt/gcc/bin/g++ cannot find libgomp.spec
REF_SOURCES = ["gemm.cpp", "gemm_int8.cpp", "gemm_fp8.cpp"]
REQUIRED_CPU_FLAGS = ("avx512f", "avx512_bf16", "avx512_vnni", "amx_bf16", "amx_int8", "amx_tile")


def host_can_run_ref() -> bool:
    """True when this host's CPU has every ISA extension the reference kernels use."""
    try:
        with open("/proc/cpuinfo") as f:
            txt = f.read()
    except OSError:
        return False
    for line in txt.splitlines():
        if line.startswith("flags"):
            flags = set(line.split(":", 1)[1].split())
            return all(x in flags for x in REQUIRED_CPU_FLAGS)
    return False


def reference_present() -> bool:
    return all(os.path.isfile(os.path.join(REF_CPU, s)) for s in REF_SOURCES)


def _torch_paths():
    import torch
    from torch.utils import cpp_extension as ce

    inc = ce.include_paths()
    lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    return inc, lib


def build(force: bool = False, verbose: bool = True) -> str | None:
    """Build oracle/_ref/sgl_cpu_oracle.so.  Returns its path, or None when the
    reference sources are not on this machine (then a prebuilt file is used if present)."""
    if not reference_present():
        return OUT_SO if os.path.isfile(OUT_SO) else None
    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = [os.path.join(REF_CPU, s) for s in REF_SOURCES] + [os.path.join(HERE, "ref_shim.cpp")]
    if not force and os.path.isfile(OUT_SO):
        newest = max(os.path.getmtime(s) for s in srcs + [os.path.abspath(__file__)])
        if os.path.getmtime(OUT_SO) >= newest:
            return OUT_SO
    inc, lib = _torch_paths()
    cflags = [
        "-O3", "-std=c++17", "-fPIC", "-fopenmp", "-march=sapphirerapids", "-Wno-unknown-pragmas",
        "-DSGLANG_CPU_FP8_CVT_FTZ", "-D_GLIBCXX_USE_CXX11_ABI=1",
        "-I" + REF_CPU, "-I" + sysconfig.get_paths()["include"],
    ] + ["-I" + p for p in inc]
    objs = []
    procs = []
    t0 = time.time()
    for s in srcs:
        o = os.path.join(OUT_DIR, os.path.basename(s).replace(".cpp", ".o"))
        objs.append(o)
        procs.append((s, subprocess.Popen([GXX, "-c", s, "-o", o] + cflags)))
    for s, p in procs:
        if p.wait() != 0:
            raise RuntimeError(f"oracle: compiling {s} failed")
    link = [GXX, "-shared", "-o", OUT_SO] + objs + [
        "-L" + lib, "-lc10", "-ltorch_cpu", "-ltorch", "-Wl,-rpath," + lib,
        "-L/usr/lib/gcc/x86_64-linux-gnu/13", "-lgomp",
    ]
    subprocess.check_call(link)
    for o in objs:
        os.remove(o)
    if verbose:
        print(f"[oracle] built {OUT_SO} in {time.time() - t0:.0f}s", file=sys.stderr)
    return OUT_SO


_LOADED = False


def load() -> bool:
    """Load the prebuilt reference .so into this process (torch.ops.sgl_kernel_ref.*).
    Returns False when it is missing or this host cannot execute it."""
    global _LOADED
    if _LOADED:
        return True
    if not os.path.isfile(OUT_SO) or not host_can_run_ref():
        return False
    import torch

    torch.ops.load_library(OUT_SO)
    _LOADED = True
    return True


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
