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
