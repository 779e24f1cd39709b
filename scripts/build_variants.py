"""Build kernel-variant copies of libfsgpu.so for A/B runs on the GPU box (FSGPU_LIB=... selects one).
usage: python scripts/build_variants.py name=file.cu:-DFLAG=1,-DOTHER=2 ...  -> fluidsolver_b200/libfsgpu_<name>.so"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsolver_b200 import build as B  # noqa: E402

B.build()
for spec in sys.argv[1:]:
    name, rest = spec.split("=", 1)
    src, flags = rest.split(":", 1)
    flags = [f for f in flags.split(",") if f]
    obj = os.path.join(B.OBJ, f"{src[:-3]}_{name}.o")
    extra = ["-fmad=false"] if src in B.NO_FMA else []
    subprocess.check_call([B.NVCC] + B.FLAGS + extra + flags + ["-c", os.path.join(B.CSRC, src), "-o", obj],
                          stdout=open(obj.replace(".o", ".ptxas.log"), "w"), stderr=subprocess.STDOUT)
    objs = [obj if s == src else os.path.join(B.OBJ, s.replace(".cu", ".o")) for s in B.SOURCES]
    lib = os.path.join(B.HERE, f"libfsgpu_{name}.so")
    subprocess.check_call([B.NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                                     "-cudart", "static"])
    print(lib)
