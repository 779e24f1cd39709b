"""CPU tests of the multi-GPU (slab) host logic: partition rules through the C ABI, and a
world_size-2 / -3 gloo run of tests/_gloo_worker.py (bootstrap, halo-exchange protocol, coarse-level
ranges, reductions).  The device side of the same path is tests/test_gpu_slab.py."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_slab_range_rules(F):
    for nx in (14, 16, 37, 96, 8192, 16384):
        for R in (1, 2, 3, 4, 8):
            if (nx + 2) // R < 4:
                with pytest.raises(F.FsgpuError):
                    F.slab_range(nx, 0, R)
                continue
            e = [F.slab_range(nx, r, R) for r in range(R)]
            assert e[0][0] == 0 and e[-1][1] == nx + 2
            assert all(e[r][1] == e[r + 1][0] for r in range(R - 1))
            sizes = [hi - lo for lo, hi in e]
            assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 4
    with pytest.raises(F.FsgpuError):
        F.slab_range(64, 2, 2)


def test_slab_rows_cover_owned_and_halo(F):
    nx, ny, R = 40, 12, 3
    for name in ("P", "U", "V"):
        m = F.field_shape(name, nx, ny)[0]
        for r in range(R):
            lo, hi = F.slab_range(nx, r, R)
            r0, r1 = F.slab_rows(name, nx, ny, r, R)
            assert r0 == max(0, lo - 2) and r1 == min(m, hi + 2)
        assert F.slab_rows(name, nx, ny, R - 1, R)[1] == m  # a U-like field's extra row: last rank


def test_coarse_range_rule(F):
    for lo, hi, foff in ((0, 9, 1), (9, 18, 1), (0, 10, 0), (10, 21, 0), (5, 6, 1)):
        c0, c1 = F.slab_coarse_range(lo, hi, foff)
        assert [c for c in range(-2, 40) if lo <= 2 * c + foff < hi] == list(range(c0, c1))


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_world(world):
    env = dict(os.environ, OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + world),
           os.path.join(ROOT, "tests", "_gloo_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert r.stdout.count(" ok") == world


def test_bench_reference_arm_under_torchrun():
    """--impl reference under torchrun: rank 0 alone prints the line, the other rank exits 0"""
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libfsref.so")):
        pytest.skip("oracle/_ref/libfsref.so not built")
    env = dict(os.environ, FSBENCH_CPU_N="256")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, "bench.py"),
           "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["value"] > 0 and d["n_gpus"] == 2
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] in ("reference", "port")
