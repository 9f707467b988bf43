"""GPU suite, machines with several GPUs only (skipped on a one-GPU box): the z-slab solver under torchrun against the
one-GPU solver — fields bit for bit, norms to 1e-12 — with several partitioned levels and, from world = 3 on, ranks that have
a neighbour on both sides; both transports (peer-mapped memory, NCCL send/recv).
And two handles on two devices in one process (per-device kernel setup state)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from cases import ROOT, digest

pytestmark = pytest.mark.gpu


def _ngpu(gs):
    return gs.load().gs_device_count()


def _torchrun(world, args, env):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29530 + world), os.path.join(ROOT, "tools", "check_multigpu.py")] + args
    e = dict(os.environ)
    e.update(env)
    return subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=e)


@pytest.mark.parametrize("transport", ["p2p", "nccl"])
@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_slab_solver_equals_one_gpu(gs, world, mode, transport):
    if _ngpu(gs) < world:
        pytest.skip(f"needs {world} GPUs")
    only = os.environ.get("GPUSOLVE_TEST_WORLDS")  # e.g. "3,4": restrict an expensive multi-GPU visit
    if only and str(world) not in only.split(","):
        pytest.skip("world not selected by GPUSOLVE_TEST_WORLDS")
    only_t = os.environ.get("GPUSOLVE_TEST_TRANSPORTS")
    if only_t and transport not in only_t.split(","):
        pytest.skip("transport not selected by GPUSOLVE_TEST_TRANSPORTS")
    env = {"GPUSOLVE_SLAB_MIN_POINTS": "0", "GPUSOLVE_P2P": "1" if transport == "p2p" else "0"}
    pre = "3" if mode == 2 else "2"
    r = _torchrun(world, ["--grid", "255", "--mode", str(mode), "--cycles", "2", "--pre", pre, "--post", pre], env)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    assert "fields bitwise OK" in r.stdout


def test_two_devices_in_one_process(gs):
    """cudaFuncSetAttribute / occupancy are per device: a second handle on another device must launch the > 48 KB
    dynamic-shared-memory kernels as well as the first"""
    if _ngpu(gs) < 2:
        pytest.skip("needs 2 GPUs")
    p = gs.make_params(127, mode=0, pre=2, post=2, maxiter=1, tol=1e-30)
    a = gs.GpuSolve(p, device=0)
    b = gs.GpuSolve(p, device=1)
    na = [a.vcycle() for _ in range(2)]
    nb = [b.vcycle() for _ in range(2)]
    assert na == nb
    assert digest(a.get(0, gs.V)) == digest(b.get(0, gs.V))
    a.close()
    b.close()
