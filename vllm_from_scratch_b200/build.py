"""Builds libpa_b200.so (the C-ABI library, include/pa_b200.h) in-tree with nvcc for sm_100a.

    python -m vllm_from_scratch_b200.build [--force] [--verbose]

The library has no torch / Python dependency (static cudart, driver entry points resolved at
run time), so it cross-compiles on a CPU-only box and loads anywhere; kernels obviously need a
B200.  The built .so is git-ignored but travels to the GPU box with the tree.
"""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
BUILD = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libpa_b200.so")

SOURCES = ["pa_api.cu", "pa_decode_tma.cu", "pa_decode_generic.cu", "pa_decode_stream.cu", "pa_decode_static.cu", "pa_reduce.cu", "pa_cache_ops.cu", "pa_prefill_tc.cu", "pa_prefill_tc2.cu", "pa_rope_cache.cu"]
HEADERS = [os.path.join(CSRC, "pa_common.cuh"), os.path.join(CSRC, "pa_tc_common.cuh"), os.path.join(INCLUDE, "pa_b200.h"), os.path.join(INCLUDE, "pa_b200_debug.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
    "-I", INCLUDE,
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libpa_b200.so")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src: str, verbose: bool) -> str:
    obj = os.path.join(BUILD, src.replace(".cu", ".o"))
    path = os.path.join(CSRC, src)
    if _stale(obj, [path] + HEADERS):
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", path, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + log)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{log}")
        if verbose:
            print(log)
    return obj


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            os.remove(os.path.join(BUILD, f))
    with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), SOURCES))
    if _stale(LIB, objs):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB


# ---- the thin torch C++ extension (csrc/torch_shim.cpp) -----------------------------------------------
SHIM_SRC = os.path.join(CSRC, "torch_shim.cpp")
SHIM = os.path.join(HERE, "_pa_torch.so")


def build_torch_shim(force: bool = False) -> str:
    """Compile csrc/torch_shim.cpp in-tree against this interpreter's torch (g++ only: the shim holds no device
    code) and link it to libpa_b200.so through an $ORIGIN rpath.  Torch is needed HERE, not in libpa_b200.so."""
    build_library()
    if not force and not _stale(SHIM, [SHIM_SRC, os.path.join(INCLUDE, "pa_b200.h"), LIB]):
        return SHIM
    import sysconfig

    import torch
    from torch.utils import cpp_extension as ce
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    inc = []
    for d in ce.include_paths("cuda") + [sysconfig.get_paths()["include"]]:
        inc += ["-isystem", d]
    libdirs = ce.library_paths("cuda")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_pa_torch", "-DTORCH_API_INCLUDE_EXTENSION_H",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}", "-Wno-attributes"] + inc + [SHIM_SRC, "-o", SHIM]
    for d in libdirs:
        cmd += ["-L", d, f"-Wl,-rpath,{d}"]
    cmd += ["-L", HERE, "-l:libpa_b200.so", "-Wl,-rpath,$ORIGIN", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
            "-ltorch_python"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(os.path.join(BUILD, "torch_shim.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("torch shim build failed:\n" + (r.stdout + r.stderr)[-4000:])
    return SHIM


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
    print(build_torch_shim(force="--force" in sys.argv))
