"""Build libfsgpu.so (hand-written sm_100a CUDA + the C ABI of include/fsgpu.h) in-tree with nvcc.

The library is the product; there is no CPU fallback and no JIT: `python -m fluidsolver_b200.build`
(or `__graft_entry__.build()`) cross-compiles without a GPU, and the resulting
fluidsolver_b200/libfsgpu.so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "csrc", "build")
LIB = os.path.join(HERE, "libfsgpu.so")
SOURCES = ["stencil.cu", "bconds.cu", "project.cu", "twophase.cu", "solver_setup.cu", "solver_cycle.cu",
           "solver_pcg.cu", "comm.cu", "output.cu", "api.cu"]
NO_FMA = {"solver_setup.cu"}
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _compile(src: str, force: bool) -> str:
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    deps = [os.path.join(CSRC, src)] + [os.path.join(CSRC, h) for h in os.listdir(CSRC)
                                        if h.endswith((".h", ".cuh"))]
    deps.append(os.path.join(HERE, "..", "include", "fsgpu.h"))
    if not force and _newer(obj, deps):
        return obj
    # the multigrid setup is compiled without FMA contraction: the coarse operators then carry
    # the same roundings as HYPRE's (built without -march, i.e. no FMA), in particular the EXACT
    # zero of the singular 1x1 coarsest operator that pfmg_setup.c:349-357 tests for
    extra = ["-fmad=false"] if src in NO_FMA else []
    cmd = [NVCC] + FLAGS + extra + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(obj.replace(".o", ".ptxas.log"), "w") as f:
        f.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(lambda s: _compile(s, force), SOURCES))
    if force or not _newer(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                     "-cudart", "static"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


def build_examples() -> list:
    """the C++ host layer (include/shim) + example drivers, plain g++ against libfsgpu.so"""
    root = os.path.dirname(HERE)
    outs = []
    for name in ("lid_driven_cavity", "two_phase_drop", "periodic_channel"):
        src = os.path.join(root, "examples", name + ".cpp")
        out = os.path.join(root, "examples", name)
        deps = [src, LIB, os.path.join(root, "include", "fsgpu.h"),
                os.path.join(root, "include", "shim", "FluidSolverGPU.hpp")]
        outs.append(out)
        if _newer(out, deps):
            continue
        cmd = [os.environ.get("CXX", "/usr/bin/g++"), "-std=c++20", "-O2", "-Wall", "-I" + os.path.join(root, "include"),
               src, "-L" + HERE, "-lfsgpu", "-Wl,-rpath,$ORIGIN/../fluidsolver_b200", "-o", out]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"g++ failed for the shim example {name}:\n{r.stdout}\n{r.stderr}")
    return outs


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(build_examples())
