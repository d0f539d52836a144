"""CPU side of the synthetic problems at BASELINE sizes (TEST INFRASTRUCTURE ONLY).

* ``randn_block_numpy`` -- NumPy restatement of the counter-based N(0,1) stream of the product's device
  generator (convex-optimization_b200/csrc/glp_misc.cuh: randn_at); same IEEE operations in the same order,
  hence the same bits.  Small blocks only (about 60 array passes per value).
* ``randn_block``       -- the same function from oracle/glprox_gen.c (OpenMP), for the 17 GB problems.
* ``gen_problem``       -- the recipe of code/main.py:37-51 at scale, exactly as
  glprox.dataset.gen_data_device lays it out: A from the stream, u / x0 from MT19937 (O(n l) host draws),
  b = A u accumulated in fp64 and rounded to A's dtype.
* ``BlockedF64``        -- lets the oracle run in fp64 on an fp32 A (the "fp32 oracle" of SURVEY.md 8(c): fp64
  arithmetic on the fp32-rounded inputs) without a 2x host copy of A: ``A @ x`` and ``A.T @ r`` convert one
  row block at a time, row blocks spread over a thread pool.

Only tests/, __graft_entry__ and the CPU legs of bench.py import this; it never imports the product.
"""
import ctypes as C
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(HERE, "_build")
LIB_PATH = os.path.join(LIB_DIR, "libglprox_oracle.so")
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def build(force=False):
    """gcc -O2 -fopenmp -ffp-contract=off (no FMA contraction: the device code uses __dmul_rn/__dadd_rn)."""
    src = os.path.join(HERE, "glprox_gen.c")
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= os.path.getmtime(src):
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared", "-o", LIB_PATH, src, "-lm"])
    return LIB_PATH


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH)
        lib.glp_oracle_randn.restype = C.c_int
        lib.glp_oracle_randn.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_uint64, C.c_int64, C.c_int64]
        lib.glp_oracle_matmul.restype = C.c_int
        lib.glp_oracle_matmul.argtypes = [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_void_p]
        _lib = lib
    return _lib


# ---- NumPy restatement ------------------------------------------------------------------------------
def _mix64(z):
    z = z + np.uint64(0x9E3779B97F4A7C15)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


_LOG_C = [1.0, 0.3333333333333333, 0.2, 0.14285714285714285, 0.1111111111111111, 0.09090909090909091,
          0.07692307692307693, 0.06666666666666667, 0.058823529411764705, 0.05263157894736842, 0.047619047619047616]
_COS_C = [1.0, -0.5, 0.041666666666666664, -0.001388888888888889, 2.48015873015873e-05, -2.755731922398589e-07,
          2.08767569878681e-09, -1.1470745597729725e-11, 4.779477332387385e-14, -1.5619206968586225e-16]
_SIN_C = [1.0, -0.16666666666666666, 0.008333333333333333, -0.0001984126984126984, 2.7557319223985893e-06,
          -2.505210838544172e-08, 1.6059043836821613e-10, -7.647163731819816e-13, 2.8114572543455206e-15]


def _horner(coef, z):
    p = np.full_like(z, coef[-1])
    for c in coef[-2::-1]:
        p = p * z + c      # NumPy never contracts: one rounded multiply, one rounded add
    return p


def _gen_log(u):
    bits = u.view(np.uint64)
    e = (bits >> np.uint64(52)).astype(np.int64) - 1023
    mant = ((bits & np.uint64(0x000FFFFFFFFFFFFF)) | np.uint64(0x3FF0000000000000)).view(np.float64)
    big = mant > 1.4142135623730951
    mant = np.where(big, mant * 0.5, mant)
    e = e + big
    s = (mant + -1.0) / (mant + 1.0)
    p = _horner(_LOG_C, s * s)
    return e.astype(np.float64) * 0.6931471805599453 + (2.0 * s) * p


def _gen_cos2pi(N):
    kq = (N + np.uint64(1 << 50)) >> np.uint64(51)
    rem = N.astype(np.int64) - (kq << np.uint64(51)).astype(np.int64)
    th = (rem.astype(np.float64) * 1.1102230246251565e-16) * 6.283185307179586
    t2 = th * th
    v = np.where((kq & np.uint64(1)) == 0, _horner(_COS_C, t2), th * _horner(_SIN_C, t2))
    return np.where((kq == 1) | (kq == 2), -v, v)


def randn_block_numpy(seed, rows, cols, row0=0, col0=0, dtype=np.float64):
    """rows x cols block at (row0, col0) of the logical N(0,1) matrix of `seed` (glp_misc.cuh: randn_at)."""
    with np.errstate(over="ignore"):
        r = (np.arange(rows, dtype=np.uint64) + np.uint64(row0)).reshape(-1, 1)
        c = (np.arange(cols, dtype=np.uint64) + np.uint64(col0)).reshape(1, -1)
        h = _mix64(_mix64(np.uint64(seed) ^ (r * np.uint64(0xD1342543DE82EF95))) + c)
        h2 = _mix64(h ^ np.uint64(0xA0761D6478BD642F))
    u1 = ((h >> np.uint64(11)) + np.uint64(1)).astype(np.float64) * 1.1102230246251565e-16
    radius = np.sqrt(-2.0 * _gen_log(u1))
    return (radius * _gen_cos2pi(h2 >> np.uint64(11))).astype(dtype)


def randn_block(seed, rows, cols, row0=0, col0=0, dtype=np.float32, out=None):
    """The same block from the C restatement (all host cores)."""
    dtype = np.dtype(dtype)
    if out is None:
        out = np.empty((rows, cols), dtype=dtype)
    assert out.dtype == dtype and out.shape == (rows, cols) and out.strides[1] == dtype.itemsize
    rc = load().glp_oracle_randn(out.ctypes.data, 0 if dtype == np.float32 else 1, rows, cols,
                                 out.strides[0] // dtype.itemsize, int(seed) & (2 ** 64 - 1), row0, col0)
    if rc != 0:
        raise RuntimeError("glp_oracle_randn failed: %d" % rc)
    return out


def sparse_truth(seed, n, l):
    """u and x0 of the scaled recipe: the reference's own draws (code/main.py:41-46) without the A draw."""
    g = np.random.Generator(np.random.MT19937(seed=seed))
    k = round(n * 0.1)
    picked = g.permutation(n)[:k]
    u = np.zeros((n, l))
    u[picked, :] = g.standard_normal(size=(k, l))
    x0 = g.standard_normal(size=(n, l))
    return u, x0


def matmul_f64(A, x):
    """A @ x with fp64 accumulation in ascending column order (zero rows of x skipped), no copy of A."""
    A = np.asarray(A)
    assert A.ndim == 2 and A.strides[1] == A.itemsize and A.dtype in (np.float32, np.float64)
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty((A.shape[0], x.shape[1]), dtype=np.float64)
    rc = load().glp_oracle_matmul(A.ctypes.data, 0 if A.dtype == np.float32 else 1, A.shape[0], A.shape[1],
                                  A.strides[0] // A.itemsize, x.ctypes.data, x.shape[1], y.ctypes.data)
    if rc != 0:
        raise RuntimeError("glp_oracle_matmul failed: %d" % rc)
    return y


def gen_problem(m, n, l, seed, dtype=np.float32, mu=1e-2, rows=None):
    """(A, b, x0, u, mu) as glprox.dataset.gen_data_device builds them on one rank: A from the counter-based
    stream, u / x0 rounded to `dtype`, b = A u (fp64 accumulation) rounded to `dtype`.  `rows` limits the
    problem to its first rows (a bounded sample of the workload for the CPU timing legs)."""
    dtype = np.dtype(dtype)
    mm = m if rows is None else min(m, rows)
    A = randn_block(seed, mm, n, 0, 0, dtype)
    u, x0 = sparse_truth(seed, n, l)
    u, x0 = u.astype(dtype), x0.astype(dtype)
    b = matmul_f64(A, u.astype(np.float64)).astype(dtype)
    return A, b, x0, u, mu


# ---- fp64 oracle arithmetic on an fp32 matrix -------------------------------------------------------
class BlockedF64:
    """Matrix adapter for oracle.solve: behaves like A.astype(float64) in `A @ x` and `A.T @ r`."""

    def __init__(self, A, block_rows=256, threads=None, _t=False):
        self.A = A
        self.block_rows = block_rows
        self.threads = threads or len(os.sched_getaffinity(0))
        self._t = _t
        self.dtype = np.dtype(np.float64)

    @property
    def shape(self):
        return self.A.shape[::-1] if self._t else self.A.shape

    @property
    def T(self):
        return BlockedF64(self.A, self.block_rows, self.threads, not self._t)

    def _blocks(self):
        m = self.A.shape[0]
        return [(r, min(m, r + self.block_rows)) for r in range(0, m, self.block_rows)]

    def __matmul__(self, x):
        x = np.asarray(x, dtype=np.float64)
        from threadpoolctl import threadpool_limits
        blocks = self._blocks()
        with threadpool_limits(limits=1), ThreadPoolExecutor(self.threads) as pool:
            if not self._t:
                out = np.empty((self.A.shape[0], x.shape[1]))

                def fwd(rr):
                    out[rr[0]:rr[1]] = self.A[rr[0]:rr[1]].astype(np.float64) @ x
                list(pool.map(fwd, blocks))
                return out
            parts = list(pool.map(lambda rr: self.A[rr[0]:rr[1]].astype(np.float64).T @ x[rr[0]:rr[1]], blocks))
        acc = parts[0]
        for p in parts[1:]:    # fixed block order: run-to-run identical
            acc = acc + p
        return acc
