"""File formats of the output row, CPU side: the numpy restatement in
tests/cases.py and the host half of the library's writer (fsgpu_npy_write_host, no GPU needed) against
files written by the unmodified reference (tests/golden/io_random24x10.npz: to_npy(directory, fs),
src/IO.hpp:231-269, and the drivers' VTK DataWriter, src/VTKWriter.hpp)."""
import os

import numpy as np

from cases import NPY_STATE_FILES, grid_coords, npy_bytes, random_state, vtk_bytes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KEY = {"VISC": "visc", "P": "p", "P_JUMP_U": "p_jump_u", "P_JUMP_V": "p_jump_v", "RHO_U": "rho_u", "RHO_V": "rho_v",
       "U": "U", "V": "V", "RHO_U_OLD": "rho_u_old", "RHO_V_OLD": "rho_v_old", "U_OLD": "U_old", "V_OLD": "V_old"}


def _arrays():
    st = random_state(24, 10, seed=2410)
    x, y = grid_coords(st)
    arrs = {k: st.a[KEY[f]] for k, f in NPY_STATE_FILES.items()}
    arrs.update(x=x, xm=(x[:-1] + x[1:]) / 2, y=y, ym=(y[:-1] + y[1:]) / 2)
    return st, x, y, arrs


def test_npy_restatement_and_host_writer_vs_reference_files(tmp_path):
    import fluidsolver_b200 as F
    g = np.load(os.path.join(GOLD, "io_random24x10.npz"))
    _, _, _, arrs = _arrays()
    assert sorted("npy_" + k for k in arrs) == sorted(k for k in g.files if k.startswith("npy_"))
    for k, a in arrs.items():
        ref = g["npy_" + k].tobytes()
        assert npy_bytes(a) == ref, k
        f = tmp_path / (k + ".npy")
        f.write_bytes(b"\0" * (len(ref) + 999))  # an older, longer file must not leave a tail
        F.npy_write_host(str(f), a)
        assert f.read_bytes() == ref, k
        assert np.array_equal(np.load(f), a)


def test_vtk_restatement_vs_reference_file(oracle):
    g = np.load(os.path.join(GOLD, "io_random24x10.npz"))
    st, x, y, _ = _arrays()
    div = oracle.calc_divergence(st)
    ui, vi = oracle.interpolate_UV(st)
    got = vtk_bytes(0.125, x, y, [("pressure", st.a["p"]), ("divergence", div)], [("velocity", ui, vi)])
    assert got == g["vtk"].tobytes()


MON_F = ("time", "dt", "max(U)", "momentum (x)")
MON_I = ("iter(p)", "a rather long column name")
MON_ROWS_F = [[0.0, 0.1, 0.0, 0.0], [0.1, 0.099, 1.5e-3, -2.25e7], [1.0 / 3.0, 1e-300, 12345.678, 9.999999e99],
              [2.5, 0.1, -1.0, 1e-5]]
MON_ROWS_I = [[0, 7], [12, 123456789], [-3, 0], [50, 10 ** 15]]
SAVE_DTS = [0.3, 0.3, 0.3, 0.3, 0.35, 0.45, 0.5, 0.25, 0.25]


def _reference_monitor_and_saves(path):
    """(log bytes, should_save sequence) from the unmodified reference; should_save keeps a function-static
    last-save time, so its sequence can only be taken once per process: the golden file keeps it"""
    from oracle_lib import Ref
    Ref.monitor_log(path, MON_F, MON_I, MON_ROWS_F, MON_ROWS_I)
    t, seq = 0.0, []
    for dt in SAVE_DTS:
        t += dt
        seq.append(int(Ref.should_save(t, dt, 1.0, 3.0)))
    seq.append(int(Ref.should_save(3.0, 0.25, 1.0, 3.0)))
    return open(path, "rb").read(), seq


def test_shim_monitor_should_save_vs_reference(tmp_path):
    """Monitor (src/Monitor.hpp), should_save, get_output_directory, init_output_directory (src/IO.hpp:97-175) of the
    C++ shim against the reference's: log bytes and save decisions equal tests/golden/monitor_log.txt, which
    tests/golden/make_golden_8f.py wrote with the unmodified reference"""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = tmp_path / "shim_host_check"
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-Wall", "-I" + os.path.join(root, "include"),
                           os.path.join(root, "tests", "shim_host_check.cpp"), "-L" + os.path.join(root, "fluidsolver_b200"),
                           "-lfsgpu", "-Wl,-rpath," + os.path.join(root, "fluidsolver_b200"), "-o", str(exe)])
    rows = "".join(" ".join(repr(v) for v in f) + " " + " ".join(str(v) for v in i) + "\n"
                   for f, i in zip(MON_ROWS_F, MON_ROWS_I))
    r = subprocess.run([str(exe), str(tmp_path / "mon.log"), str(tmp_path / "dirs")], input=rows, capture_output=True,
                       text=True, check=True)
    gold = open(os.path.join(GOLD, "monitor_log.txt"), "rb").read()
    log, saves = gold.split(b"#saves ")
    assert (tmp_path / "mon.log").read_bytes() == log
    out = dict(line.split(" ", 1) for line in r.stdout.strip().splitlines())
    assert out["save"] == saves.decode().strip()
    assert out["dir"] == "/base/test/output/shim_host_check/" and out["emptied"] == "1"
