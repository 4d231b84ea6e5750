"""In-tree build of the two native artefacts (no pip install, no JIT cache):

  das3r_b200/libdas3r_raster.so   CUDA kernels + the C ABI of include/das3r_raster.h (no torch dependency)
  das3r_b200/_C<ext-suffix>.so    thin torch binding over that C ABI (g++ only)

Both are compiled for sm_100a only (`-gencode arch=compute_100a,code=sm_100a -lineinfo`). nvcc cross-compiles
without a GPU, so this runs in the CPU-only build container; the .so files travel to the GPU box with the tree.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
BUILD_DIR = os.path.join(HERE, "build")

LIB = os.path.join(HERE, "libdas3r_raster.so")
EXT = os.path.join(HERE, "_C" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))

CU_SOURCES = ["forward.cu", "backward.cu", "knn.cu", "frontend.cu", "loss.cu", "adam.cu", "comm.cu", "init.cu", "api.cu"]
HEADERS = ["common.cuh", "kernels.cuh", os.path.join(ROOT, "include", "das3r_raster.h"), "blend_common.cuh",
           os.path.join(ROOT, "include", "das3r_frontend.h"), os.path.join(ROOT, "include", "das3r_loss.h"), os.path.join(ROOT, "include", "das3r_adam.h"),
           os.path.join(ROOT, "include", "das3r_comm.h"), os.path.join(ROOT, "include", "das3r_init.h")]


def _mtime(p: str) -> float:
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError(f"build step failed: {cmd[0]} ... {cmd[-1]}")


def _newest_header() -> float:
    return max(_mtime(h if os.path.isabs(h) else os.path.join(CSRC, h)) for h in HEADERS)


def build_lib(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(BUILD_DIR, exist_ok=True)
    hdr = _newest_header()
    objs, jobs = [], []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(BUILD_DIR, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _mtime(o) < max(_mtime(s), hdr):
            cmd = [NVCC, *ARCH, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)
    if jobs:
        with ThreadPoolExecutor(max_workers=len(jobs)) as ex:
            list(ex.map(_run, jobs))
    if force or jobs or not os.path.exists(LIB):
        _run([NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "static"])
    return LIB


def build_ext(force: bool = False) -> str:
    src = os.path.join(CSRC, "torch_ext.cpp")
    if not force and _mtime(EXT) >= max(_mtime(src), _mtime(HEADERS[2])) and os.path.exists(EXT):
        return EXT
    import torch
    from torch.utils import cpp_extension as ce

    inc = [f"-I{p}" for p in ce.include_paths()] + [f"-I{sysconfig.get_paths()['include']}", "-I/usr/local/cuda/include"]
    torch_lib = os.path.join(os.path.dirname(torch.__file__), "lib")
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=_C",
           "-DTORCH_API_INCLUDE_EXTENSION_H", f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           *inc, src, "-o", EXT, f"-L{HERE}", "-ldas3r_raster", f"-L{torch_lib}", "-ltorch", "-ltorch_cpu",
           "-ltorch_cuda", "-ltorch_python", "-lc10", "-lc10_cuda", "-Wl,-rpath,$ORIGIN", f"-Wl,-rpath,{torch_lib}"]
    _run(cmd)
    return EXT


def build_all(force: bool = False) -> None:
    build_lib(force)
    build_ext(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    print(LIB)
    print(EXT)
