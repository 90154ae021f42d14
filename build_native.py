"""Builds the in-tree native artefacts for sm_100a.

  distributed-pytorch-training_b200/_native/libdptb200.so   CUDA kernels + C ABI (no torch dependency)
  distributed-pytorch-training_b200/_native/_dptb200_C.so   torch extension: C++ reducer / bucket planner

Both are compiled with explicit nvcc / g++ command lines (no JIT cache): the .so files travel
with the repository snapshot to the GPU box.  Incremental: a source is recompiled only when it
(or a header) is newer than its object file.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG = ROOT / "distributed-pytorch-training_b200"
OUT = PKG / "_native"
OBJ = ROOT / "build" / "obj"
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
NVCC = os.path.join(CUDA_HOME, "bin", "nvcc")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]

KERNEL_SOURCES = [
    "csrc/ops/igemm_tcgen05.cu",
    "csrc/ops/gemm2sm_tcgen05.cu",
    "csrc/ops/norm_pool_loss.cu",
    "csrc/ops/optim.cu",
    "csrc/comm/p2p_collectives.cu",
]
EXT_SOURCES = [
    "csrc/reducer/reducer.cpp",
    "csrc/reducer/bindings.cpp",
]


def _headers_mtime() -> float:
    m = 0.0
    for pat in ("csrc/**/*.cuh", "csrc/**/*.h", "csrc/**/*.hpp", "include/*.h"):
        for h in ROOT.glob(pat):
            m = max(m, h.stat().st_mtime)
    return m


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError(f"native build failed: {cmd[0]} {cmd[-1]}")


def _stale(src: Path, obj: Path, hdr_m: float) -> bool:
    return (not obj.exists()) or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_m)


def _extra_defines() -> list[str]:
    """Build variants selected through the environment (each changes the object names, so switching
    back and forth never links stale objects).  None at the moment."""
    return []


def build_kernels(verbose: bool = False) -> Path:
    OUT.mkdir(parents=True, exist_ok=True)
    OBJ.mkdir(parents=True, exist_ok=True)
    hdr_m = _headers_mtime()
    extra = _extra_defines()
    tag = "".join("." + d[2:].split("=")[0].lower() for d in extra)
    stamp = OBJ / "variant.txt"
    relink = (stamp.read_text() if stamp.exists() else "") != tag
    jobs, objs = [], []
    for s in KERNEL_SOURCES:
        src = ROOT / s
        obj = OBJ / (src.stem + tag + ".o")
        objs.append(obj)
        if _stale(src, obj, hdr_m):
            jobs.append([NVCC, *NVCC_FLAGS, *extra, "-c", str(src), "-o", str(obj)])
    if verbose and jobs:
        print(f"[build] compiling {len(jobs)} CUDA sources for sm_100a")
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(_run, jobs))
    lib = OUT / "libdptb200.so"
    if jobs or relink or not lib.exists():
        _run([NVCC, "-shared", "-o", str(lib), *map(str, objs), "-lcudart"])
        stamp.write_text(tag)
    return lib


def build_extension(verbose: bool = False) -> Path | None:
    srcs = [ROOT / s for s in EXT_SOURCES if (ROOT / s).exists()]
    if not srcs:
        return None
    import torch
    from torch.utils import cpp_extension

    OUT.mkdir(parents=True, exist_ok=True)
    OBJ.mkdir(parents=True, exist_ok=True)
    hdr_m = _headers_mtime()
    inc = [f"-I{p}" for p in cpp_extension.include_paths("cuda")] + [
        f"-I{sysconfig.get_paths()['include']}", f"-I{ROOT / 'include'}"]
    cxx11 = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    flags = ["-O2", "-std=c++17", "-fPIC", f"-D_GLIBCXX_USE_CXX11_ABI={cxx11}",
             "-DTORCH_EXTENSION_NAME=_dptb200_C", "-DTORCH_API_INCLUDE_EXTENSION_H", "-w"]
    jobs, objs = [], []
    for src in srcs:
        obj = OBJ / (src.stem + ".ext.o")
        objs.append(obj)
        if _stale(src, obj, hdr_m):
            jobs.append(["g++", *flags, *inc, "-c", str(src), "-o", str(obj)])
    if verbose and jobs:
        print(f"[build] compiling {len(jobs)} C++ extension sources")
    with ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(_run, jobs))
    lib = OUT / "_dptb200_C.so"
    if jobs or not lib.exists():
        tlib = Path(torch.__file__).parent / "lib"
        _run(["g++", "-shared", "-o", str(lib), *map(str, objs), f"-L{tlib}", f"-L{OUT}",
              "-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda",
              f"-L{CUDA_HOME}/lib64", "-lcudart", "-ldptb200",
              f"-Wl,-rpath,{tlib}", "-Wl,-rpath,$ORIGIN"])
    return lib


def build_all(verbose: bool = False) -> None:
    build_kernels(verbose)
    build_extension(verbose)


if __name__ == "__main__":
    build_all(verbose=True)
    print("ok")
