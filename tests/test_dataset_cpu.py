"""Host logic either side of the solvers: the fixture generator, the GLPSHARD
on-disk format (any stored shard count -> any world size), the metric algebra and the log / table
formats of the reference driver (code/main.py:37-51,94-130).  No GPU, no CUDA calls."""
import os
import struct
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "convex-optimization_b200"))
sys.path.insert(0, os.path.join(HERE, ".."))

from glprox import combine_metrics, dataset, shard_bounds  # noqa: E402
import main as driver  # noqa: E402  (convex-optimization_b200/main.py)
from oracle import glprox_oracle as oracle  # noqa: E402


def test_gen_data_is_the_reference_fixture():
    # known answers harvested from the unmodified reference
    n, m, l, mu, A, b, u, x0 = dataset.gen_data(97006855)
    assert (n, m, l, mu) == (512, 256, 2, 1e-2)
    assert A.shape == (256, 512) and A.dtype == np.float64 and A.flags.c_contiguous
    np.testing.assert_allclose(A[0, :3], [0.22040485, -1.50563052, 1.83398592], atol=5e-9)
    assert abs(A.sum() - 311.912367670878) < 1e-9
    assert abs(b.sum() - 136.01022273437763) < 1e-9
    assert abs(x0.sum() - 43.342126701375136) < 1e-10
    assert np.count_nonzero(np.abs(u).sum(axis=1)) == 51
    # and the same arrays as the oracle's restatement, bit for bit
    oA, ob, ox0, ou, _ = oracle.gen_data(97006855)
    for mine, theirs in ((A, oA), (b, ob), (u, ou), (x0, ox0)):
        assert np.array_equal(mine, theirs)


def _numpy_metrics(x, u, ref=None):
    """The lambdas of code/main.py:48-50, restated."""
    fro = lambda z: np.sqrt(np.sum(z * z))
    return {"errfun": (lambda x1, x2: fro(x1 - x2) / (1 + fro(x1))),
            "err_exact": fro(x - u) / (1 + fro(x)),
            "sparsity": np.sum(np.abs(x) > 1e-6 * np.max(np.abs(x))) / x.size}


@pytest.mark.parametrize("world", [1, 2, 3])
def test_combine_metrics_matches_the_reference_lambdas(world):
    rng = np.random.default_rng(5)
    n, l = 640, 3
    x = rng.standard_normal((n, l)) * (rng.random((n, 1)) < 0.2)
    x[7] *= 1e-9  # below the sparsity cut
    u = rng.standard_normal((n, l)) * (rng.random((n, 1)) < 0.1)
    want = _numpy_metrics(x, u)
    gmax = np.max(np.abs(x))
    tot = np.zeros(6)
    for c0, c1 in (shard_bounds(n, world, 64) if world > 1 else [(0, n)]):
        xs, us = x[c0:c1], u[c0:c1]
        tot += [np.sum(xs * xs), np.sum(us * us), np.sum((xs - us) ** 2), np.sum(np.sqrt(np.sum(xs * xs, axis=1))),
                np.sum(np.abs(xs) > 1e-6 * gmax), xs.size]
    got = combine_metrics(list(tot), gmax, 2.5, 1e-2, True)
    assert got["sparsity"] == want["sparsity"]
    assert abs(got["err_exact"] - want["err_exact"]) <= 1e-14
    assert abs(got["err_ref"] - want["errfun"](u, x)) <= 1e-14
    assert abs(got["objective"] - (1.25 + 1e-2 * np.sum(np.linalg.norm(x, axis=1)))) <= 1e-12
    assert np.isnan(combine_metrics(list(tot), gmax, None, 1e-2, False)["objective"])
    assert np.isnan(combine_metrics(list(tot), gmax, None, 1e-2, False)["err_exact"])


def _problem(m, n, l, dtype, seed=3):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n)).astype(dtype)
    u = (rng.standard_normal((n, l)) * (rng.random((n, 1)) < 0.1)).astype(dtype)
    x0 = rng.standard_normal((n, l)).astype(dtype)
    b = (A.astype(np.float64) @ u).astype(dtype)
    return A, b, x0, u


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("stored,world", [(1, 1), (1, 2), (3, 1), (3, 2), (2, 3)])
def test_shard_round_trip_any_stored_count_to_any_world(tmp_path, dtype, stored, world):
    m, n, l = 37, 448, 2
    A, b, x0, u = _problem(m, n, l, dtype)
    prefix = str(tmp_path / "prob")
    align = 128 if dtype == np.float32 else 64
    cuts = shard_bounds(n, stored, align) if stored > 1 else [(0, n)]
    for k, (c0, c1) in enumerate(cuts):
        path = dataset.write_shard(prefix, k, stored, A[:, c0:c1], b, x0[c0:c1], u[c0:c1], n_total=n, col0=c0,
                                   mu=0.02, seed=11)
        assert os.path.getsize(path) % dataset.HEADER_BYTES == 0
    shards = dataset.open_shards(prefix)
    assert [s.k for s in shards] == list(range(stored))
    assert all(s.seed == 11 and s.mu == 0.02 and s.dtype == np.dtype(dtype) for s in shards)
    for c0, c1 in (shard_bounds(n, world, align) if world > 1 else [(0, n)]):
        pieces, bb, xx0, uu, mu = dataset.read_columns(shards, c0, c1)
        got = np.empty((m, c1 - c0), dtype)
        for dst, block in pieces:
            got[:, dst:dst + block.shape[1]] = block
        assert np.array_equal(got, A[:, c0:c1])
        assert np.array_equal(bb, b) and np.array_equal(xx0, x0[c0:c1]) and np.array_equal(uu, u[c0:c1])
        assert mu == 0.02


def test_shard_without_truth_and_ragged_or_empty_ranges(tmp_path):
    A, b, x0, _ = _problem(5, 70, 1, np.float64)
    prefix = str(tmp_path / "nou")
    dataset.write_shard(prefix, 0, 1, A, b, x0, None, n_total=70, col0=0)
    shards = dataset.open_shards(prefix)
    assert shards[0].u is None
    pieces, bb, xx0, uu, _ = dataset.read_columns(shards, 3, 64)   # ragged, unaligned range
    assert uu is None and np.array_equal(pieces[0][1], A[:, 3:64]) and np.array_equal(xx0, x0[3:64])
    pieces, _, xx0, _, _ = dataset.read_columns(shards, 10, 10)     # empty range
    assert pieces == [] and xx0.shape == (0, 1)
    with pytest.raises(ValueError):
        dataset.read_columns(shards, 0, 71)


def test_shard_files_are_validated(tmp_path):
    A, b, x0, u = _problem(4, 128, 2, np.float32)
    prefix = str(tmp_path / "bad")
    with pytest.raises(FileNotFoundError):
        dataset.open_shards(prefix)
    with pytest.raises(ValueError):
        dataset.write_shard(prefix, 0, 1, A.astype(np.float16), b, x0, u, n_total=128, col0=0)
    # one of two shards missing
    dataset.write_shard(prefix, 0, 2, A[:, :64], b, x0[:64], u[:64], n_total=128, col0=0)
    with pytest.raises(ValueError, match="found 1 shard files"):
        dataset.open_shards(prefix)
    # a gap in the tiling
    dataset.write_shard(prefix, 1, 2, A[:, 96:], b, x0[96:], u[96:], n_total=128, col0=96)
    with pytest.raises(ValueError, match="does not continue the tiling"):
        dataset.open_shards(prefix)
    path = dataset.write_shard(prefix, 1, 2, A[:, 64:], b, x0[64:], u[64:], n_total=128, col0=64)
    assert len(dataset.open_shards(prefix)) == 2
    raw = open(path, "rb").read()
    # wrong magic / future version / truncated body
    open(path, "wb").write(b"NOTSHARD" + raw[8:])
    with pytest.raises(ValueError, match="not a GLPSHARD"):
        dataset.Shard(path)
    open(path, "wb").write(raw[:8] + struct.pack("<I", 9) + raw[12:])
    with pytest.raises(ValueError, match="unsupported version"):
        dataset.Shard(path)
    open(path, "wb").write(raw[:dataset.HEADER_BYTES + 16])
    with pytest.raises(ValueError, match="shorter than its header"):
        dataset.Shard(path)
    open(path, "wb").write(raw[:40])
    with pytest.raises(ValueError, match="truncated header"):
        dataset.Shard(path)
    assert not [f for f in os.listdir(tmp_path) if f.endswith(".part")]


def test_log_line_and_table_follow_the_reference_formats():
    # the ProxGD row of doc/report.md:446 from the reference's own numbers
    met = {"sparsity": 0.099609375, "err_exact": 3.790643790049492e-05}
    row = driver.format_row("ProxGD Primal", 1.53, 1768, 6.103773850792892e-01, met, 4.38e-06)
    assert row == {"solver": "ProxGD Primal", "cpu": " 1.53", "iter": " 1768", "optval": "6.10377E-01",
                   "sparsity": "0.0996", "err-to-exact": "3.79E-05", "err-to-best": "4.38E-06"}
    assert driver.log_line(row) == ("[ProxGD Primal]: cpu:  1.53, iter:  1768, optval: 6.10377E-01, "
                                    "sparsity: 0.0996, err-to-exact: 3.79E-05, err-to-best: 4.38E-06")
    assert driver.format_row("x", 0.0, None, 1.0, met, None)["iter"] == "   -1"   # main.py:118
    table = driver.markdown_table([row]).splitlines()
    assert table[0] == "# Statistics"
    assert [c.strip() for c in table[1].strip("|").split("|")] == driver.HEADERS
    assert set(table[2]) == {"|", "-"}
    assert [c.strip() for c in table[3].strip("|").split("|")] == [row[h].strip() for h in driver.HEADERS]
    assert len({len(t) for t in table[1:]}) == 1  # aligned columns
