"""Real multi-GPU parity (one process per GPU, CUDA-IPC mailboxes over NVLink).  Skipped on boxes
with a single GPU; the same kernels are covered there by tests/test_gpu_multirank.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_solve_over_nvlink(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world),
           os.path.join(ROOT, "tests", "dist_parity_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "DIST_PARITY_OK %d" % world in r.stdout


def _table(stdout):
    rows = [ln for ln in stdout.splitlines() if ln.startswith("|") and "Primal" in ln]
    return [[c.strip() for c in ln.strip("|").split("|")] for ln in rows]


def test_driver_sharded_over_two_gpus_prints_the_same_table():
    """convex-optimization_b200/main.py (the reference's code/main.py) under torchrun: metrics are combined
    across the column shards on the device side and the table matches the single-GPU one."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    main_py = os.path.join(ROOT, "convex-optimization_b200", "main.py")
    args = ["--rows", "256", "--cols", "2048", "--dtype", "f64", "--solvers", "FProxGD,FGD", "--maxit", "60",
            "--alpha0", "-1", "--log-level", "WARNING"]
    one = subprocess.run([sys.executable, main_py] + args, capture_output=True, text=True, timeout=600)
    assert one.returncode == 0, one.stdout[-2000:] + one.stderr[-2000:]
    two = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", main_py] + args,
                         capture_output=True, text=True, timeout=600)
    assert two.returncode == 0, two.stdout[-2000:] + two.stderr[-2000:]
    t1, t2 = _table(one.stdout), _table(two.stdout)
    assert len(t1) == 2 and len(t2) == 2
    for a, b in zip(t1, t2):
        assert a[0] == b[0] and a[2:] == b[2:], (a, b)   # all columns but the wall-clock one
