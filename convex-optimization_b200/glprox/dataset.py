"""Problem data either side of the solvers: generation and an on-disk shard format.

The reference regenerates (A, b, u, x0) from a seed on every run (code/main.py:37-51) and has no
on-disk or wire format.  At 17 GB per problem that does not scale, so this module provides

* ``gen_data(seed)``            -- the reference's fixture, value for value (NumPy MT19937; host side);
* ``gen_data_device(...)``      -- the same recipe at scale: A ~ N(0,1) from the library's counter-based
                                   device generator (``glp_randn``: any column block of the same logical
                                   matrix can be produced independently on any rank), 10 % non-zero row
                                   groups in u, b = A u accumulated in fp64 (``glp_rhs_exact``); every value is
                                   bit-identical to the CPU restatement in oracle/gen.py;
* ``write_shard / open_shard / read_columns`` -- a flat binary format, one file per column block of A
                                   (the sharding of SURVEY.md 8(e)), readable by any world size:
                                   a rank's columns are gathered from whichever stored shards hold them.

File layout (little endian), one file ``<prefix>.<k>-of-<K>.glps`` per stored shard k of K:

    offset 0     char[8]  magic "GLPSHARD"
           8     u32      version (1)         12  u32  dtype (0 = float32, 1 = float64)
          16     i64      m                   24  i64  n_total          32  i64  l
          40     i64      col0                48  i64  n_local          56  i64  lda (elements, >= n_local)
          64     f64      mu                  72  u64  seed
          80     u64      off_A   88 u64 off_b   96 u64 off_x0   104 u64 off_u (0 = absent)
         112     u32      shard index k      116  u32  shard count K
    offset 4096  A[m][lda] row-major, then b[m][l] (replicated in every shard), x0[n_local][l], u[n_local][l],
                 each section starting on a 4096-byte boundary (page-aligned for mmap / direct IO).
"""
import os
import struct

import numpy as np

MAGIC = b"GLPSHARD"
VERSION = 1
HEADER_BYTES = 4096
_HDR = struct.Struct("<8sIIqqqqqqdQQQQQII")
_DTYPES = {0: np.float32, 1: np.float64}
_DTYPE_ID = {np.dtype(np.float32): 0, np.dtype(np.float64): 1}


def gen_data(seed=97006855, n=512, m=256, l=2, mu=1e-2):
    """code/main.py:37-51 (same draws in the same order, so the arrays are identical to the reference's)."""
    g = np.random.Generator(np.random.MT19937(seed=seed))
    A = g.standard_normal(size=(m, n))
    k = round(n * 0.1)
    p = g.permutation(n)[:k]
    u = np.zeros(shape=(n, l))
    u[p, :] = g.standard_normal(size=(k, l))
    b = np.matmul(A, u)
    x0 = g.standard_normal(size=(n, l))
    return n, m, l, mu, A, b, u, x0


def sparse_truth(seed, n, l):
    """u and x0 of the scaled recipe (host draws: O(n l), independent of the sharding)."""
    g = np.random.Generator(np.random.MT19937(seed=seed))
    k = round(n * 0.1)
    picked = g.permutation(n)[:k]
    u = np.zeros((n, l))
    u[picked, :] = g.standard_normal(size=(k, l))
    x0 = g.standard_normal(size=(n, l))
    return u, x0


def gen_data_device(m, n, l, seed, dtype="f32", rank=0, world=1, mu=1e-2, group=None):
    """This rank's shard of the synthetic problem, generated in HBM.
    Returns (A[m, n_local], b[m, l], x0[n_local, l], u[n_local, l], mu, (col0, col1))."""
    import torch
    from . import _lib
    from .solver import shard_bounds
    lib = _lib.load()
    f32 = dtype in ("f32", np.float32, torch.float32)
    tdt = torch.float32 if f32 else torch.float64
    col0, col1 = shard_bounds(n, world, 128 if f32 else 64)[rank] if world > 1 else (0, n)
    n_loc = col1 - col0
    A = torch.empty((m, n_loc), dtype=tdt, device="cuda")
    _lib.check(lib.glp_randn(A.data_ptr(), _lib.GLP_F32 if f32 else _lib.GLP_F64, m, n_loc, A.stride(0), seed, 0, col0,
                             torch.cuda.current_stream().cuda_stream))
    u, x0 = sparse_truth(seed, n, l)
    u_loc = torch.from_numpy(u[col0:col1]).to("cuda", tdt)
    x0_loc = torch.from_numpy(x0[col0:col1]).to("cuda", tdt)
    # b = A u with fp64 accumulation in ascending column order (glp_rhs_exact: the same order and rounding as the
    # CPU restatement oracle/glprox_gen.c, so both build the same b); partial product per rank, summed over ranks
    u_rounded = u[col0:col1].astype(np.float32 if f32 else np.float64).astype(np.float64)
    u_loc64 = torch.from_numpy(u_rounded).to("cuda")
    nz = torch.from_numpy(np.flatnonzero(np.any(u_rounded != 0.0, axis=1)).astype(np.int64)).to("cuda")
    Au = torch.empty((m, l), dtype=torch.float64, device="cuda")
    _lib.check(lib.glp_rhs_exact(A.data_ptr(), _lib.GLP_F32 if f32 else _lib.GLP_F64, m, A.stride(0), u_loc64.data_ptr(), l,
                                 nz.data_ptr() if nz.numel() else None, nz.numel(), Au.data_ptr(),
                                 torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(Au, group=group)
    return A, Au.to(tdt), x0_loc, u_loc, mu, (col0, col1)


# ---- on-disk shards -------------------------------------------------------------------------------
def shard_path(prefix, k, K):
    return "%s.%03d-of-%03d.glps" % (prefix, k, K)


def _align(off):
    return (off + HEADER_BYTES - 1) // HEADER_BYTES * HEADER_BYTES


def _rows_to_file(f, arr, chunk_bytes=256 << 20):
    """Write a 2-D NumPy array or (CUDA) torch tensor row block by row block (bounded host memory)."""
    if isinstance(arr, np.ndarray):
        a = np.ascontiguousarray(arr)
        rows = max(1, chunk_bytes // max(1, a.strides[0])) if a.ndim == 2 and a.shape[0] else 1
        for r in range(0, a.shape[0], rows):
            f.write(memoryview(a[r:r + rows]).cast("B"))
        return
    rows = max(1, chunk_bytes // max(1, arr.shape[1] * arr.element_size()))
    for r in range(0, arr.shape[0], rows):
        f.write(memoryview(arr[r:r + rows].contiguous().cpu().numpy()).cast("B"))


def write_shard(prefix, k, K, A, b, x0, u=None, *, n_total, col0, mu=1e-2, seed=0):
    """Store shard k of K: A = A_full[:, col0:col0+n_local], x0/u the matching row groups, b in full."""
    m, n_loc = int(A.shape[0]), int(A.shape[1])
    l = int(b.shape[1])
    np_dt = np.dtype(str(A.dtype).replace("torch.", ""))
    if np_dt not in _DTYPE_ID:
        raise ValueError("A must be float32 or float64")
    es = np_dt.itemsize
    conv = lambda z: z.astype(np_dt) if isinstance(z, np.ndarray) else z.to(dtype=A.dtype) if hasattr(z, "to") else np.asarray(z, np_dt)
    off_A = HEADER_BYTES
    off_b = _align(off_A + m * n_loc * es)
    off_x0 = _align(off_b + m * l * es)
    off_u = _align(off_x0 + n_loc * l * es) if u is not None else 0
    hdr = _HDR.pack(MAGIC, VERSION, _DTYPE_ID[np_dt], m, int(n_total), l, int(col0), n_loc, n_loc, float(mu),
                    int(seed) & (2 ** 64 - 1), off_A, off_b, off_x0, off_u, int(k), int(K))
    path = shard_path(prefix, k, K)
    tmp = path + ".part"
    with open(tmp, "wb") as f:
        f.write(hdr.ljust(HEADER_BYTES, b"\0"))
        for off, arr in ((off_A, A), (off_b, conv(b)), (off_x0, conv(x0))) + (((off_u, conv(u)),) if u is not None else ()):
            f.seek(off)
            _rows_to_file(f, arr)
        f.truncate(_align(f.tell()))
    os.replace(tmp, path)  # a reader never sees a half-written shard
    return path


class Shard:
    """Header + memory-mapped sections of one stored shard (nothing is read until sliced)."""

    def __init__(self, path):
        self.path = path
        with open(path, "rb") as f:
            raw = f.read(_HDR.size)
        if len(raw) < _HDR.size:
            raise ValueError("%s: truncated header" % path)
        (magic, ver, dt, self.m, self.n_total, self.l, self.col0, self.n_local, self.lda, self.mu, self.seed,
         off_A, off_b, off_x0, off_u, self.k, self.K) = _HDR.unpack(raw)
        if magic != MAGIC:
            raise ValueError("%s: not a GLPSHARD file" % path)
        if ver != VERSION:
            raise ValueError("%s: unsupported version %d" % (path, ver))
        if dt not in _DTYPES:
            raise ValueError("%s: unknown dtype id %d" % (path, dt))
        self.dtype = np.dtype(_DTYPES[dt])
        need = (off_u or off_x0) + self.n_local * self.l * self.dtype.itemsize
        if os.path.getsize(path) < need:
            raise ValueError("%s: file shorter than its header says (%d < %d)" % (path, os.path.getsize(path), need))
        mm = lambda off, shape: np.memmap(path, dtype=self.dtype, mode="r", offset=off, shape=shape)
        self.A = mm(off_A, (self.m, self.lda))[:, :self.n_local] if self.m and self.n_local else np.empty((self.m, self.n_local), self.dtype)
        self.b = mm(off_b, (self.m, self.l)) if self.m else np.empty((0, self.l), self.dtype)
        self.x0 = mm(off_x0, (self.n_local, self.l)) if self.n_local else np.empty((0, self.l), self.dtype)
        self.u = (mm(off_u, (self.n_local, self.l)) if self.n_local else np.empty((0, self.l), self.dtype)) if off_u else None
        self.col1 = self.col0 + self.n_local


def open_shards(prefix):
    """All stored shards of a problem, ordered by column; checks that they tile [0, n_total)."""
    import glob
    paths = sorted(glob.glob(glob.escape(prefix) + ".*-of-*.glps"))
    if not paths:
        raise FileNotFoundError("no shards match %s.*-of-*.glps" % prefix)
    shards = sorted((Shard(p) for p in paths), key=lambda s: s.col0)
    s0 = shards[0]
    if len(shards) != s0.K:
        raise ValueError("%s: found %d shard files, headers say %d" % (prefix, len(shards), s0.K))
    pos = 0
    for s in shards:
        if (s.m, s.n_total, s.l, s.dtype, s.K) != (s0.m, s0.n_total, s0.l, s0.dtype, s0.K) or s.col0 != pos:
            raise ValueError("%s: shard %s does not continue the tiling at column %d" % (prefix, s.path, pos))
        pos = s.col1
    if pos != s0.n_total:
        raise ValueError("%s: shards cover %d of %d columns" % (prefix, pos, s0.n_total))
    return shards


def read_columns(shards, c0, c1):
    """(pieces, b, x0, u, mu): the columns [c0, c1) of A as a list of (dst_col, memmap block) and the
    matching row groups, gathered from the stored shards that hold them (any stored K -> any world)."""
    s0 = shards[0]
    if not 0 <= c0 <= c1 <= s0.n_total:
        raise ValueError("column range [%d, %d) outside [0, %d)" % (c0, c1, s0.n_total))
    pieces, x0s, us = [], [], []
    for s in shards:
        lo, hi = max(c0, s.col0), min(c1, s.col1)
        if lo >= hi:
            continue
        pieces.append((lo - c0, s.A[:, lo - s.col0:hi - s.col0]))
        x0s.append(s.x0[lo - s.col0:hi - s.col0])
        us.append(None if s.u is None else s.u[lo - s.col0:hi - s.col0])
    cat = lambda parts: np.concatenate(parts, axis=0) if parts else np.empty((0, s0.l), s0.dtype)
    u = None if any(p is None for p in us) or not us and s0.u is None else cat(us)
    return pieces, np.asarray(s0.b), cat(x0s), u, s0.mu


def load_problem_device(prefix, rank=0, world=1, chunk_bytes=256 << 20):
    """This rank's column block of a stored problem, uploaded to HBM through two pinned staging
    buffers (the file is read one row block at a time; host memory use is bounded by 2 chunks).
    Returns (A, b, x0, u, mu, (col0, col1))."""
    import torch
    from .solver import shard_bounds
    shards = open_shards(prefix)
    s0 = shards[0]
    f32 = s0.dtype == np.float32
    c0, c1 = shard_bounds(s0.n_total, world, 128 if f32 else 64)[rank] if world > 1 else (0, s0.n_total)
    pieces, b, x0, u, mu = read_columns(shards, c0, c1)
    tdt = torch.float32 if f32 else torch.float64
    A = torch.empty((s0.m, c1 - c0), dtype=tdt, device="cuda")
    es = s0.dtype.itemsize
    stage = [torch.empty(chunk_bytes // es, dtype=tdt).pin_memory() for _ in range(2)]
    done = [torch.cuda.Event(), torch.cuda.Event()]
    slot = 0
    for dst, block in pieces:
        w = block.shape[1]
        rows = max(1, (chunk_bytes // es) // max(1, w))
        for r in range(0, s0.m, rows):
            nr = min(rows, s0.m - r)
            done[slot].synchronize()  # the previous copy out of this buffer has finished
            host = stage[slot][:nr * w].view(nr, w)
            np.copyto(host.numpy(), block[r:r + nr])
            A[r:r + nr, dst:dst + w].copy_(host, non_blocking=True)
            done[slot].record()
            slot ^= 1
    torch.cuda.synchronize()
    dev = lambda z: None if z is None else torch.from_numpy(np.array(z)).to("cuda", tdt)  # copy: memmaps are read-only
    return A, dev(b), dev(x0), dev(u), mu, (c0, c1)
