"""In-tree build of ``_lib/libdlion.so`` (sm_100a only).

``python -m distributed_lion_pytorch_b200.build`` cross-compiles every ``csrc/*.cu`` with
``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo`` (no GPU needed) and links one
shared library with a plain C ABI (``include/dlion.h``).  The library links cudart
statically and depends on nothing else, so it is loadable with ``ctypes`` from any process —
including ones that never import torch.  The ``.so`` is git-ignored but travels to the GPU
box with the working tree.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
LIBDIR = PKG / "_lib"
OBJDIR = LIBDIR / "obj"
LIB = LIBDIR / "libdlion.so"
INCLUDE = PKG.parent / "include"

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v",
              "--expt-relaxed-constexpr", "-I", str(INCLUDE)]


def find_nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC=/path/to/nvcc")
    return cand


def _sources():
    return sorted(CSRC.glob("*.cu"))


def _stamp(srcs) -> str:
    h = hashlib.sha256()
    for f in sorted(list(srcs) + list(CSRC.glob("*.cuh")) + [INCLUDE / "dlion.h"]):
        h.update(f.name.encode())
        h.update(f.read_bytes())
    h.update(" ".join(ARCH_FLAGS + NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> Path:
    srcs = _sources()
    stamp_file = LIBDIR / "stamp.txt"
    stamp = _stamp(srcs)
    if not force and LIB.exists() and stamp_file.exists() and stamp_file.read_text().strip() == stamp:
        return LIB
    nvcc = find_nvcc()
    OBJDIR.mkdir(parents=True, exist_ok=True)
    log_lines = []

    def compile_one(src: Path):
        obj = OBJDIR / (src.stem + ".o")
        cmd = [nvcc, *ARCH_FLAGS, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(compile_one, srcs))
    objs = [str(o) for o, _ in results]
    for (o, err), s in zip(results, srcs):
        log_lines.append(f"==== {s.name}\n{err}")
    link = [nvcc, *ARCH_FLAGS, "-shared", "-o", str(LIB), *objs, "-cudart", "static", "-ldl"]
    r = subprocess.run(link, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    (LIBDIR / "ptxas.log").write_text("\n".join(log_lines))
    stamp_file.write_text(stamp)
    if verbose:
        print("\n".join(log_lines))
    return LIB


CPU_SRC = PKG / "csrc_cpu" / "cpu_ops.cpp"
CPU_LIB = LIBDIR / "libdlion_cpu.so"


def build_cpu(force: bool = False) -> Path:
    """g++ -O3 -fopenmp build of the host-side codec/vote library (no CUDA involved)."""
    if not force and CPU_LIB.exists() and CPU_LIB.stat().st_mtime >= CPU_SRC.stat().st_mtime:
        return CPU_LIB
    LIBDIR.mkdir(parents=True, exist_ok=True)
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    base = [cxx, "-O3", "-std=c++17", "-fPIC", "-shared", str(CPU_SRC), "-o", str(CPU_LIB)]
    r = subprocess.run(base[:5] + ["-fopenmp"] + base[5:], capture_output=True, text=True)
    if r.returncode != 0:          # no libgomp on this image: the pragmas are then simply ignored
        r = subprocess.run(base, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed for cpu_ops.cpp:\n{r.stdout}\n{r.stderr}")
    return CPU_LIB


if __name__ == "__main__":
    build_cpu(force="--force" in sys.argv)
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
