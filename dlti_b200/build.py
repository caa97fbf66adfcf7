"""In-tree build of the native library (CUDA kernels for sm_100a + C++ host runtime).

`python -m dlti_b200.build` (or `build_all()`) compiles every `csrc/*.cu` and `csrc/*.cpp`
with nvcc/g++ into `dlti_b200/lib/libdlti_b200.so` (GPU kernels, C ABI declared in
`include/dlti_b200.h`) and `dlti_b200/lib/libdlti_cpu.so` (host-only: the AVX-512 / AVX2 + OpenMP CPU Adam
of the ZeRO-Offload path). The .so files are git-ignored but travel with the gpurun snapshot.

The reference has no build system at all; DeepSpeed JIT-builds its ops
with TORCH_CUDA_ARCH_LIST=7.0 (reference training/train_deepspeed_zero2.py:125-129) — here
everything is built ahead of time for exactly one target: sm_100a.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "lib"
OBJDIR = PKG.parent / "build" / "obj"
INCLUDE = PKG.parent / "include"

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
# the image exports CXX=/opt/gcc/bin/g++, a trimmed toolchain without libgomp: prefer the distro g++
CXX = (os.environ.get("DLTI_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else None)
       or os.environ.get("CXX") or shutil.which("g++") or "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-I", str(CSRC), "-I", str(INCLUDE),
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-march=x86-64-v3", "-mf16c", "-I", str(CSRC),
             "-I", str(INCLUDE)]

GPU_LIB = LIBDIR / "libdlti_b200.so"
CPU_LIB = LIBDIR / "libdlti_cpu.so"


def _newer(src: Path, dst: Path, extra: list[Path]) -> bool:
    if not dst.exists():
        return True
    t = dst.stat().st_mtime
    return any(s.stat().st_mtime > t for s in [src, *extra])


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    if r.stderr.strip() and os.environ.get("DLTI_BUILD_VERBOSE"):
        print(r.stderr, file=sys.stderr)


def build_gpu(force: bool = False, verbose: bool = False) -> Path:
    OBJDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    srcs = sorted(CSRC.glob("*.cu"))
    jobs = []
    objs = []
    for s in srcs:
        o = OBJDIR / (s.stem + ".o")
        objs.append(o)
        if force or _newer(s, o, headers):
            cmd = [NVCC, *NVCC_FLAGS, "-c", str(s), "-o", str(o)]
            if verbose:
                cmd.insert(1, "-Xptxas")
                cmd.insert(2, "-v")
            jobs.append(cmd)
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if jobs or not GPU_LIB.exists():
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(GPU_LIB),
              *map(str, objs), "-lcudart"])
    return GPU_LIB


def build_debug() -> Path:
    """Debug variant of the GPU library (-DDLTI_DEBUG_CHECKS: signal-pad sequence checks in the peer-memory
    barriers; see ops/_debug.py for the host-side half). Written next to the product library, never loaded unless
    DLTI_GPU_LIB points at it."""
    out = LIBDIR / "variants" / "libdlti_b200_debug.so"
    objdir = OBJDIR.parent / "obj_debug"
    objdir.mkdir(parents=True, exist_ok=True)
    out.parent.mkdir(parents=True, exist_ok=True)
    objs, jobs = [], []
    for s in sorted(CSRC.glob("*.cu")):
        o = objdir / (s.stem + ".o")
        objs.append(o)
        jobs.append([NVCC, *NVCC_FLAGS, "-DDLTI_DEBUG_CHECKS", "-c", str(s), "-o", str(o)])
    with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
        list(ex.map(_run, jobs))
    _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(out), *map(str, objs), "-lcudart"])
    return out


def build_cpu(force: bool = False) -> Path:
    OBJDIR.mkdir(parents=True, exist_ok=True)
    LIBDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cpp"))
    if not srcs:
        return CPU_LIB
    if force or any(_newer(s, CPU_LIB, []) for s in srcs):
        _run([CXX, *CXX_FLAGS, "-shared", "-o", str(CPU_LIB), *map(str, srcs), "-lrt", "-lpthread"])
    return CPU_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_cpu(force)
    build_gpu(force, verbose)


if __name__ == "__main__":
    if "--debug-checks" in sys.argv:
        print("built:", build_debug())
        sys.exit(0)
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built:", GPU_LIB, CPU_LIB if CPU_LIB.exists() else "")
