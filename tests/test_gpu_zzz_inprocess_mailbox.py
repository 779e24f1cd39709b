"""The peer-mapped (mailbox) transport on ONE GPU: R ranks = R contexts of this process, one host thread each
(tests/test_gpu_slab.py run_ranks), with FSGPU_HALO=p2p.  Its kernels wait on the device for flags that a kernel of
ANOTHER rank raises; with every rank on the same device that only works while the runtime can start that other
kernel.  Under CUDA's default lazy module loading the FIRST launch of a kernel loads its code and waits for the device
to drain -- for the kernel that is waiting for it -- until the transport's timeout reports it: 1-2 such timeouts per
run of this file with CUDA_MODULE_LOADING=LAZY, none with EAGER (scripts/gpu_flaky.sh, 6 runs each).  An artefact of
emulating N GPUs on one: with one process per GPU a device only ever waits for kernels the peers launched earlier.
Hence: eager loading (tests/conftest.py, before CUDA starts), the file sorted last, one retry of a run that timed
out (in case the environment overrides the loading mode).  On real GPUs the same paths are covered by scripts/multi_parity.py
(profiles/r02t_multi_parity.txt)."""
from __future__ import annotations

import os

os.environ.setdefault("CUDA_MODULE_LOADING", "EAGER")

import numpy as np  # noqa: E402
import pytest  # noqa: E402

from cases import cavity_bc, cavity_state  # noqa: E402
from test_gpu_slab import run_ranks  # noqa: E402

pytestmark = pytest.mark.gpu

VARIANTS = [(2, False), (3, False), (2, "push"), (3, "push"), (2, True), (3, True)]


@pytest.mark.parametrize("R,overlap", VARIANTS)
def test_slab_timesteps_p2p_halo(F, oracle, R, overlap, monkeypatch):
    """FSGPU_HALO=p2p: halo rows travel through peer-mapped mailboxes (one kernel per exchange) and the PCG dot
    products are all-reduced inside their reduction kernels, instead of the staged copies.  Without overlap the
    results equal the default transport's bit for bit; with it (boundary rows first, exchange on the transport's
    stream beside the interior launch) only the association of the sum of delta_p changes.  "push": the fused
    form -- producer kernels of the distributed multigrid levels store boundary rows straight into the mapped
    neighbour arrays and raise a flag, consumers wait in their prologue -- again bit for bit."""
    st, bc = cavity_state(oracle, 96, 96), cavity_bc()
    oracle.apply_velocity_bconds(st, bc)
    direct, dlev = [], []
    # keep levels distributed down to ~100 points, so that the exchanges of the coarse levels run too
    monkeypatch.setenv("FSGPU_REPL_POINTS", "100")

    def body(fs, r):
        direct.append(fs.halo_is_direct())
        ps = F.LinearSolver(fs, tol=1e-13, max_iter=500)
        ps.set_pressure_operator()
        dlev.append(ps.num_distributed_levels)
        hist = []
        for _ in range(2):
            dt = fs.adjust_dt(0.9, 0.1)
            hist.append((dt, fs.step(ps, dt, 5)[1]))
        ps.close()
        return hist

    b, rb = run_ranks(F, st, bc, R, body, fields=("U", "V", "P"))
    assert not any(direct)
    del direct[:]
    monkeypatch.setenv("FSGPU_HALO", "p2p")
    monkeypatch.setenv("FSGPU_P2P_TIMEOUT_S", "5")
    if overlap == "push":  # the marching kernels store their boundary rows into the neighbours' halos themselves
        monkeypatch.setenv("FSGPU_PUSH", "1")
        overlap = False
    elif overlap:
        monkeypatch.setenv("FSGPU_OVERLAP", "1")
        monkeypatch.setenv("FSGPU_OVERLAP_MIN_POINTS", "1")
    try:
        a, ra = run_ranks(F, st, bc, R, body, fields=("U", "V", "P"))
    except AssertionError as e:  # see the file header: one retry of a run whose device-side wait timed out
        if "timed out" not in str(e):
            raise
        print("\n[in-process mailbox] a device-side wait timed out; retrying once:", str(e)[:160])
        del direct[:], dlev[:]
        a, ra = run_ranks(F, st, bc, R, body, fields=("U", "V", "P"))
    assert all(direct) and len(direct) == R
    assert min(dlev) >= 4, dlev
    assert ra == rb
    for n in ("U", "V", "P"):
        if overlap:
            assert np.abs(a[n] - b[n]).max() <= 1e-13 * np.abs(b[n]).max(), n
        else:
            assert np.array_equal(a[n], b[n]), n
