"""World-size-2 (and 3) run of the z-slab protocol on CPU over gloo: the partition the library computes
(gs_partition), the lazy two-plane ghost refresh and the norm all-reduce must reproduce the whole-grid oracle bit
for bit (Jacobi is order independent: SURVEY.md 8e "Parity under sharding")."""
import os
import socket

import numpy as np
import pytest

from cases import seeded_field
import slab_model

V, F, R = 0, 1, 2


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_numpy_sweep_is_the_oracle_sweep(oracle_mod):
    n = 14
    o = oracle_mod.Oracle(oracle_mod.make_params(n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), "port")
    shape = o.shape(0)
    v = seeded_field(shape, 5)
    f = seeded_field(shape, 6, scale=1.0)
    o.set(0, V, v)
    o.set(0, F, f)
    o.smooth(0, 1)
    got, _ = slab_model.jacobi_sweep(v, f, 1.0 / (n + 1), 0.8, 1, n + 1)
    assert np.array_equal(got, o.get(0, V))
    o.close()


@pytest.mark.parametrize("n,world", [(63, 2), (95, 3)])
def test_slab_sweeps_over_gloo_equal_whole_grid_oracle(gs, oracle_mod, tmp_path, monkeypatch, n, world):
    import torch.multiprocessing as mp
    # levels below 3e6 points are replicated by default; lower the threshold so a CPU-sized grid is split
    monkeypatch.setenv("GPUSOLVE_SLAB_MIN_POINTS", "1000")
    _, _, rep = gs.partition(gs.make_params(n), world, 0, 0)
    assert rep >= 1
    out = str(tmp_path / "slabs.npz")
    mp.spawn(slab_model.worker, args=(world, _free_port(), n, 11, out), nprocs=world, join=True)
    got = np.load(out)
    o = oracle_mod.Oracle(oracle_mod.make_params(n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), "port")
    shape = o.shape(0)
    o.set(0, V, seeded_field(shape, 11))
    o.set(0, F, seeded_field(shape, 12, scale=1.0))
    o.smooth(0, 4)
    norm4 = o.residual(0)          # ||r|| of the iterate after four sweeps
    o.smooth(0, 1)
    assert np.array_equal(got["v"], o.get(0, V)), "slab result differs from the whole-grid reference"
    assert abs(np.sqrt(float(got["sumsq"])) - norm4) <= 1e-12 * norm4
    o.close()


@pytest.mark.parametrize("n,world", [(69, 2), (63, 3), (63, 4), (25, 4)])
# 69: even-sized levels and an odd slab boundary on the deepest partitioned level; 25 on 4 ranks: the last rank holds one
# interior plane plus the halo plane of the deepest partitioned level
def test_slab_vcycles_over_gloo_equal_whole_grid_oracle(oracle_mod, tmp_path, monkeypatch, n, world):
    """whole V(2,2) cycles on z-slabs with the partition the library computes: partitioned levels with lazy two-plane
    ghosts, coarse-aligned restriction across the slab boundary, the first replicated level gathered, prolongation from
    a one-plane coarse ghost — fields equal the whole-grid oracle bit for bit after every cycle count tried"""
    import torch.multiprocessing as mp
    monkeypatch.setenv("GPUSOLVE_SLAB_MIN_POINTS", "1000")
    out = str(tmp_path / "mg.npz")
    cycles = 2
    mp.spawn(slab_model.vcycle_worker, args=(world, _free_port(), n, 41, cycles, out), nprocs=world, join=True)
    got = np.load(out)
    assert int(got["rep"]) >= 2, "the test wants at least two partitioned levels"
    o = oracle_mod.Oracle(oracle_mod.make_params(n, mode=0, pre=2, post=2, maxiter=1, tol=1e-30), "port")
    shape = o.shape(0)
    o.set(0, V, seeded_field(shape, 41))
    o.set(0, F, seeded_field(shape, 42, scale=1.0))
    norms = [o.vcycle() for _ in range(cycles)]
    assert np.array_equal(got["v"], o.get(0, V)), "slab V-cycle differs from the whole-grid reference"
    for s, want in zip(got["sums"], norms):
        assert abs(np.sqrt(s) - want) <= 1e-12 * want
    o.close()


@pytest.mark.parametrize("mode", [1, 2])
def test_slab_vcycles_in_the_exp_modes(oracle_mod, tmp_path, monkeypatch, mode):
    """FAS (two-output restriction, coarse right-hand side from restV, v -= restV before the prolongation) and the Newton
    inner cycle on three slabs — a rank with neighbours on both sides.  numpy's exp is not glibc's, so bit equality is
    asked between the slab run and the same model on one rank (what sharding must not change), and 1e-12 against the
    oracle (north_star's tolerance for the exp() modes)"""
    import torch.multiprocessing as mp
    monkeypatch.setenv("GPUSOLVE_SLAB_MIN_POINTS", "1000")
    n, cycles = 63, 2
    outs = []
    for world in (3, 1):
        out = str(tmp_path / f"mg{world}.npz")
        mp.spawn(slab_model.vcycle_worker, args=(world, _free_port(), n, 41, cycles, out, mode), nprocs=world, join=True)
        outs.append(np.load(out))
    assert int(outs[0]["rep"]) >= 2 and int(outs[1]["rep"]) == 0
    assert np.array_equal(outs[0]["v"], outs[1]["v"]), "sharding changed the bits"
    o = oracle_mod.Oracle(oracle_mod.make_params(n, mode=mode, pre=2, post=2, maxiter=1, tol=1e-30), "port")
    o.set(0, V, seeded_field(o.shape(0), 41))
    o.set(0, F, seeded_field(o.shape(0), 42, scale=1.0))
    if mode == 2:
        for l in range(o.num_levels):
            o.set(l, 5, seeded_field(o.shape(l), 51 + l))
    norms = [o.vcycle() for _ in range(cycles)]
    want = o.get(0, V)
    assert np.abs(outs[1]["v"] - want).max() <= 1e-12 * np.abs(want).max()
    assert abs(np.sqrt(outs[1]["sums"][-1]) - norms[-1]) <= 1e-11 * norms[-1]
    o.close()
