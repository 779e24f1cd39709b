"""CPU test (-m "not gpu"): the C++ host layer (include/shim/FluidSolverGPU.hpp) compiles and links against
libfsgpu.so with every call of the reference's surface instantiated (tests/shim_compile_check.cpp);
the binary returns before its first device call when run without arguments."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shim_compiles_and_links(F, tmp_path):
    exe = str(tmp_path / "shim_compile_check")
    cmd = [os.environ.get("CXX", "g++"), "-std=c++20", "-O0", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "shim_compile_check.cpp"), "-L" + os.path.dirname(F.LIB_PATH), "-lfsgpu",
           "-Wl,-rpath," + os.path.dirname(F.LIB_PATH), "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert subprocess.run([exe]).returncode == 0
