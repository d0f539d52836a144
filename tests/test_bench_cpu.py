"""CPU tests of the checker-side pieces bench.py relies on: the generator restatements, the fp64 adapter,
the staged-reference loader and the reference arm itself (which must never map the product library)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from oracle import gen
from oracle import glprox_oracle as orc


def test_c_and_numpy_generators_agree_bit_for_bit():
    for dt in (np.float32, np.float64):
        a = gen.randn_block_numpy(97006855, 257, 1031, 5, 77, dt)
        b = gen.randn_block(97006855, 257, 1031, 5, 77, dt)
        assert a.dtype == dt and np.array_equal(a, b)
    # sharding invariance: a column block drawn on its own equals the slice of the whole
    whole = gen.randn_block(3, 64, 768, 0, 0, np.float64)
    assert np.array_equal(gen.randn_block(3, 64, 256, 0, 256, np.float64), whole[:, 256:512])
    assert np.array_equal(gen.randn_block(3, 16, 768, 48, 0, np.float64), whole[48:])


def test_generator_is_standard_normal():
    x = gen.randn_block(11, 2048, 4096, 0, 0, np.float64).ravel()
    assert abs(x.mean()) < 2e-3 and abs(x.std() - 1) < 2e-3
    assert abs((x ** 3).mean()) < 5e-3 and abs((x ** 4).mean() - 3) < 2e-2
    assert abs((np.abs(x) > 3).mean() - 0.0026998) < 1e-4
    # the polynomial log / cos behind it are accurate to a few ulp
    u = np.random.default_rng(0).random(50000)
    u[:2] = (1.0, 2.0 ** -53)
    assert np.max(np.abs(gen._gen_log(u) - np.log(u)) / np.maximum(np.abs(np.log(u)), 1e-300)) < 1e-15
    N = (np.random.default_rng(1).random(50000) * 2.0 ** 53).astype(np.uint64)
    assert np.max(np.abs(gen._gen_cos2pi(N) - np.cos(2 * np.pi * (N.astype(np.float64) * 2.0 ** -53)))) < 2e-15


def test_gen_problem_follows_the_recipe():
    A, b, x0, u, mu = gen.gen_problem(300, 1100, 3, 9, np.float32)
    assert A.shape == (300, 1100) and b.shape == (300, 3) and x0.shape == u.shape == (1100, 3) and mu == 1e-2
    assert int((np.abs(u).sum(axis=1) > 0).sum()) == round(1100 * 0.1)          # code/main.py:41-44
    assert np.array_equal(b, (A.astype(np.float64) @ u.astype(np.float64)).astype(np.float32)) or \
        np.max(np.abs(b - (A.astype(np.float64) @ u.astype(np.float64)))) <= 1e-6 * np.max(np.abs(b))
    As, bs, _, _, _ = gen.gen_problem(300, 1100, 3, 9, np.float32, rows=64)   # a row sample is a consistent sub-problem
    assert np.array_equal(As, A[:64]) and np.array_equal(bs, b[:64])


def test_blocked_f64_adapter_equals_astype():
    rng = np.random.default_rng(4)
    A = rng.standard_normal((700, 900)).astype(np.float32)
    x, r = rng.standard_normal((900, 2)), rng.standard_normal((700, 2))
    B = gen.BlockedF64(A, block_rows=96, threads=3)
    A64 = A.astype(np.float64)
    assert B.shape == (700, 900) and B.T.shape == (900, 700)
    assert np.max(np.abs(B @ x - A64 @ x)) <= 1e-11 and np.max(np.abs(B.T @ r - A64.T @ r)) <= 1e-11
    # the oracle runs on it unchanged
    b = (A64 @ rng.standard_normal((900, 2))).astype(np.float32).astype(np.float64)
    opts = {"maxit": 4, "alpha0": float(1.0 / (np.sqrt(700) + np.sqrt(900)) ** 2)}
    x0 = rng.standard_normal((900, 2))
    for algo in ("ProxGD", "FProxGD"):
        x1, k1, o1 = orc.solve(algo, x0, B, b, 1e-2, opts)
        x2, k2, o2 = orc.solve(algo, x0, A64, b, 1e-2, opts)
        assert k1 == k2 and abs(o1["fval"] - o2["fval"]) <= 1e-12 * abs(o2["fval"])


def test_staged_reference_loader():
    from oracle import make_ref
    if make_ref.stage(verbose=False) is None and make_ref.load() is None:
        pytest.skip("no /root/reference here and nothing staged")
    prox, fprox = make_ref.load()
    A, b, x0, u, mu = orc.gen_data(97006855, 64, 128, 2)
    for name, fn in (("ProxGD", prox), ("FProxGD", fprox)):
        x, k, out = fn(x0, A, b, mu, {"maxit": 6})
        xr, kr, outr = orc.solve(name, x0, A, b, mu, {"maxit": 6})
        assert k == kr and out["fval"] == outr["fval"] and np.array_equal(x, xr)


def _run_bench(extra, env_extra):
    env = dict(os.environ, **env_extra)
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + extra, env=env, capture_output=True,
                         text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


def test_reference_arm_is_clean_and_threaded():
    """--impl reference: no product library mapped, full host threads even under torchrun's OMP_NUM_THREADS=1,
    steps / warmup honoured, config identical to what our arm prints for the same command line."""
    flags = ["--rows", "1024", "--cols", "4096", "--steps", "2", "--warmup", "1", "--gpus", "1"]
    line = json.loads(_run_bench(["--impl", "reference"] + flags, {"OMP_NUM_THREADS": "1", "RANK": "0"}).strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["product_library_mapped"] is False
    assert line["steps"] == 2 and line["warmup"] == 1 and line["gpu_launches"] == 0
    assert line["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0)) == int(line["omp_num_threads_env"])
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["value"] > 0
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0
    sys.path.insert(0, ROOT)
    import bench
    import argparse
    ns = argparse.Namespace(algo="FProxGD", m=1024, n=4096, l=2, dtype="f32", seed=97006855, gpus=1)
    assert line["config"] == bench.bench_config(ns)
    # other ranks of a torchrun launch print nothing and exit 0
    assert _run_bench(["--impl", "reference"] + flags, {"RANK": "1"}).strip() == ""
