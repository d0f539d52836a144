"""Host side of the drop-in solvers: option semantics, marshalling, handle lifetime.

Mirrors the reference interface  (x, iter, out) = gl_X(x0, A, b, mu_0, opts)
(code/gl_ProxGD_primal.py:9,146 ; code/gl_FProxGD_primal.py:9,161 ; caller code/main.py:113-116).
All numerics happen in libglprox.so (hand-written sm_100a kernels); this module only merges
options, moves buffers and formats the reference's `out` dictionary.
"""
import ctypes as C
import logging
import time

import numpy as np

from . import _lib

logger = logging.getLogger("opt")  # same logger name as the reference (gl_ProxGD_primal.py:6)

# gl_ProxGD_primal.py:10-19 / gl_FProxGD_primal.py:10-19
DEFAULT_OPTS = {
    "ProxGD": {"maxit": 2500, "thres": 1e-3, "step_type": "line_search", "alpha0": 2e-3, "ftol": 1e-6,
               "stable_len_threshold": 70, "line_search_attenuation_coeffi": 0.9,
               "maxit_line_search_iter": 5},
    "FProxGD": {"maxit": 1500, "thres": 1e-3, "step_type": "line_search", "alpha0": 1e-3, "ftol": 1e-6,
                "stable_len_threshold": 70, "line_search_attenuation_coeffi": 0.98,
                "maxit_line_search_iter": 5},
}
# gl_FGD_primal.py:10-21, gl_GD_primal.py:10-19, gl_SGD_primal.py:10-18
DEFAULT_OPTS.update({
    "FGD": {"maxit": 1500, "thres": 1e-3, "step_type": "line_search", "alpha0": 1e-3, "ftol": 1e-6,
            "stable_len_threshold": 70, "line_search_attenuation_coeffi": 0.98, "maxit_line_search_iter": 5,
            "delta": 1e-6},
    "GD": {"maxit": 2500, "thres": 1e-3, "step_type": "diminishing", "alpha0": 1e-3, "ftol": 1e-5,
           "stable_len_threshold": 100, "continuous_subgradient_flag": False, "delta": 1e-3},
    "SGD": {"maxit": 2100, "thres": 1e-3, "step_type": "diminishing", "alpha0": 1e-3, "ftol": 1e-5,
            "stable_len_threshold": 100, "continuous_subgradient_flag": False},
})
# gl_ADMM_dual.py:11-17
ADMM_DUAL_DEFAULT_OPTS = {"maxit": 100, "thres": 1e-3, "tau": (1 + 5 ** 0.5) * 0.5, "rho": 1e2, "converge_len": 20}
ADMM_PRIMAL_DEFAULT_OPTS = {"maxit": 100, "thres": 1e-3, "tau": (1 + 5 ** 0.5) * 0.5, "rho": 1e-2, "eta_0": 100,
                            "converge_len": 10, "converge_thres": 1e-5, "step_type": "fixed"}   # gl_ADMM_primal.py:11-20
ALM_DUAL_DEFAULT_OPTS = dict(ADMM_DUAL_DEFAULT_OPTS)                                            # gl_ALM_dual.py:67-73
ALGO_ID = {"ProxGD": _lib.GLP_PROXGD, "FProxGD": _lib.GLP_FPROXGD, "FGD": _lib.GLP_FGD, "GD": _lib.GLP_GD,
           "SGD": _lib.GLP_SGD}
# additive keys (not in the reference): dtype 'fp32'|'fp64', backend 'auto'|'stream'|'persistent',
# device (int), poll_interval (int), dist (bool), tensor_cores (wide l >= 16 mode) 'auto' (= fp16 pairs) | 'f16x2' |
# '3xtf32' | 'tf32+bf16' | 'tf32' | 'off'
TENSOR_MODES = {"auto": 0, "off": 1, "tf32": 2, "tf32+bf16": 3, "f16x2": 4, "3xtf32": 5}


def _tensor_mode(merged):
    mode = merged.get("tensor_cores", "auto")
    if mode not in TENSOR_MODES:
        raise ValueError("opts['tensor_cores'] must be one of %s, got %r" % (sorted(TENSOR_MODES), mode))
    return TENSOR_MODES[mode]


def merge_opts(algo, opts):
    """`{**default_opts, **opts}` (gl_*_primal.py:21): caller keys win, unknown keys are ignored."""
    merged = dict(DEFAULT_OPTS[algo])
    merged.update(opts or {})
    return merged


def pack_opts(algo, merged):
    st = merged["step_type"]
    if st not in _lib.STEP_TYPES:
        # the reference logs "Unsupported type." and then dies with TypeError (gl_ProxGD_primal.py:100-101)
        logger.error("Unsupported type.")
        raise ValueError("unsupported step_type %r (expected one of %s)" % (st, sorted(_lib.STEP_TYPES)))
    if st == "bb" and algo != "ProxGD":
        raise ValueError("step_type 'bb' (Barzilai-Borwein first trial step, an additive opt-in) exists for gl_ProxGD_primal only")
    if algo in ("GD", "SGD"):
        if st == "line_search":  # the reference returns None as the step and dies with TypeError (gl_GD_primal.py:75-76)
            logger.error("Unsupported type.")
            raise ValueError("gl_%s_primal has no line search: step_type must be fixed, diminishing or diminishing2" % algo)
        # continuous_subgradient_flag (gl_GD_primal.py:43-45, gl_SGD_primal.py:35-37): the caller has replaced
        # alpha0 by 1 / lambda_max(A^T A) (GroupLassoProblem.lambda_max: Lanczos on the two stream kernels)
    backend = merged.get("backend", "auto")
    if backend not in _lib.BACKENDS:
        raise ValueError("unknown backend %r" % (backend,))
    o = _lib.GlpOpts()
    o.maxit = int(merged["maxit"])
    o.step_type = _lib.STEP_TYPES[st]
    o.stable_len_threshold = int(merged["stable_len_threshold"])
    o.maxit_line_search_iter = int(merged.get("maxit_line_search_iter", 0))
    o.thres = float(merged["thres"])
    o.alpha0 = float(merged["alpha0"])
    o.ftol = float(merged["ftol"])
    o.line_search_attenuation_coeffi = float(merged.get("line_search_attenuation_coeffi", 1.0))
    o.delta = float(merged.get("delta", 0.0))
    o.backend = _lib.BACKENDS[backend]
    o.poll_interval = int(merged.get("poll_interval", 0))
    return o


def _running_min(f):
    """f_hist_best with Python's min(f_best, f_now) semantics (NaN never replaces the best)."""
    f = np.asarray(f, dtype=np.float64)
    if f.size == 0 or not np.isnan(f).any():
        return list(np.minimum.accumulate(f)) if f.size else []
    best, out = np.inf, []
    for v in f:
        best = v if v < best else best
        out.append(np.float64(best))
    return out


def _format_out(res, f_hist, tt):
    fh = [np.float64(v) for v in f_hist]
    return {
        "tt": float(tt),
        "fval": np.float64(res.fval),
        "f_hist": fh,
        "f_hist_best": _running_min(f_hist),
        "glprox": {"device_ms": res.device_ms, "n_fwd": int(res.n_fwd), "n_tpass": int(res.n_tpass),
                   "n_launches": int(res.n_launches), "sparsity": res.sparsity,
                   "stage_end": [int(v) for v in res.stage_end],
                   "backend": {1: "stream", 2: "persistent"}.get(res.backend_used, "?")},
    }


def _raise_for(rc, res=None):
    if rc == _lib.GLP_ERR_UNBOUND:
        # gl_FProxGD_primal.py:156,161 with maxit == 0
        raise UnboundLocalError("cannot access local variable 'x' where it is not associated with a value")
    _lib.check(rc)
    if res is not None and res.status != _lib.GLP_OK:
        # the library returned before the device-side state machine reached `done`: never hand a truncated
        # solve back as a normal (x, k, out)
        raise RuntimeError("libglprox: solve ended with status %d before the stop rule or the iteration budget was "
                           "reached (k=%d)" % (res.status, res.iters))


def _mode_dtype(A_dtype, merged):
    want = merged.get("dtype")
    if want is None:
        return np.float32 if A_dtype == np.float32 else np.float64
    table = {"fp32": np.float32, "float32": np.float32, "f32": np.float32,
             "fp64": np.float64, "float64": np.float64, "f64": np.float64}
    if want not in table:
        raise ValueError("opts['dtype'] must be 'fp32' or 'fp64', got %r" % (want,))
    return table[want]


def _log_history(f_hist, res, merged=None, mu_0=None, s_hist=None):
    """The reference's DEBUG lines (gl_ProxGD_primal.py:45,54,134-136 / gl_FProxGD_primal.py:45,56,149-151), in the
    reference's order, reproduced after the solve from what the device recorded: `alpha0=` once, `new mu=` at
    the start of every continuation stage, and every 100th iteration `iter / objective / sparsity`, where the
    objective is the one recorded AT iteration k and the sparsity is that of the iterate PRODUCED by iteration
    k (the reference prints sparsity_func(x) of the new x): s_hist[k] when the next iteration recorded it, the
    returned iterate's sparsity when k is the last iteration."""
    if not logger.isEnabledFor(logging.DEBUG):
        return
    if merged is not None:
        logger.debug("alpha0= {:10E}".format(float(merged["alpha0"])))
    stage_end = [int(v) for v in res.stage_end]
    k_total = len(f_hist)
    start = 0
    for stage, factor in enumerate((100.0, 10.0, 1.0)):
        if mu_0 is not None:
            logger.debug("new mu= {:10E}".format(factor * mu_0))
        end = stage_end[stage] if stage_end[stage] > 0 or stage == 0 else start
        end = min(max(end, start), k_total)
        # a stage shorter than maxit ended on the stop rule: the reference breaks before it reaches the log line
        # of that last iteration (:123-125 / :132-134)
        broke = merged is not None and (end - start) < int(merged["maxit"])
        for k in range(start + 1, end + 1):
            if k % 100 == 0 and not (broke and k == end):
                if s_hist is not None and k < len(s_hist):
                    sp = float(s_hist[k])
                elif k == k_total:
                    sp = float(res.sparsity)
                else:
                    sp = float("nan")
                logger.debug("iter= {:5}, objective= {:10E}, sparsity= {:3f}".format(k, float(f_hist[k - 1]), sp))
        start = end


def solve(algo, x0, A, b, mu_0, opts):
    """Drop-in body shared by gl_ProxGD_primal / gl_FProxGD_primal."""
    merged = merge_opts(algo, opts)
    o = pack_opts(algo, merged)
    try:
        import torch
        is_torch = isinstance(A, torch.Tensor)
    except ImportError:  # torch is only plumbing for the device-tensor path
        is_torch = False
    if is_torch:
        return _solve_torch(algo, x0, A, b, mu_0, merged, o)
    mode = _tensor_mode(merged)
    if merged.get("dist"):
        return _solve_numpy_sharded(algo, x0, A, b, mu_0, merged, o)
    if mode != 0 or logger.isEnabledFor(logging.DEBUG) or (algo in ("GD", "SGD") and merged.get("continuous_subgradient_flag")):
        # the one-shot host entry point always uses the default tensor mode and keeps no handle to read the
        # sparsity history from (DEBUG log lines): explicit modes and DEBUG logging go through a handle
        import torch
        dt = _mode_dtype(np.asarray(A).dtype, merged)
        dev = torch.device("cuda", int(merged.get("device", 0)))
        to = lambda z: torch.from_numpy(np.ascontiguousarray(z, dtype=dt)).to(dev)
        x, k, out = _solve_torch(algo, to(x0), to(A), to(b), mu_0, merged, o)
        return x.cpu().numpy(), k, out
    return _solve_numpy(algo, x0, A, b, mu_0, merged, o)


def _solve_numpy(algo, x0, A, b, mu_0, merged, o):
    lib = _lib.load()
    A = np.asarray(A)
    dt = _mode_dtype(A.dtype, merged)
    if A.ndim != 2:
        raise ValueError("A must be 2-D")
    m, n = A.shape
    x0 = np.asarray(x0)
    b = np.asarray(b)
    if x0.ndim != 2 or b.ndim != 2 or x0.shape[0] != n or b.shape[0] != m or x0.shape[1] != b.shape[1]:
        raise ValueError("shape mismatch: A %s, x0 %s, b %s" % (A.shape, x0.shape, b.shape))
    l = x0.shape[1]
    t0 = time.perf_counter()
    Ac = np.ascontiguousarray(A, dtype=dt)
    bc = np.ascontiguousarray(b, dtype=dt)
    x0c = np.ascontiguousarray(x0, dtype=dt)
    x = np.empty((n, l), dtype=dt)
    cap = 3 * int(o.maxit) + 8
    f_hist = np.empty(cap, dtype=np.float64)
    res = _lib.GlpResult()
    rc = lib.glp_solve_host(int(merged.get("device", 0)), ALGO_ID[algo],
                            _lib.GLP_F32 if dt == np.float32 else _lib.GLP_F64, m, n, l,
                            Ac.ctypes.data, Ac.strides[0] // Ac.itemsize, bc.ctypes.data, x0c.ctypes.data,
                            float(mu_0), C.byref(o), x.ctypes.data, C.byref(res), f_hist.ctypes.data, cap)
    _raise_for(rc, res)
    tt = time.perf_counter() - t0
    fh = f_hist[:res.iters]
    _log_history(fh, res, merged, float(mu_0))
    return x, int(res.iters), _format_out(res, fh, tt)


def _solve_numpy_sharded(algo, x0, A, b, mu_0, merged, o):
    """opts['dist'] = True (additive key): one process per GPU under torch.distributed; A is THIS
    rank's column block A[:, J_rank] and x0 its row groups x0[J_rank, :] (host arrays, ideally
    pinned); b is the full right-hand side on every rank.  Returns this rank's rows of x."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        raise RuntimeError("opts['dist'] needs an initialised torch.distributed process group")
    dt = _mode_dtype(np.asarray(A).dtype, merged)
    tdt = torch.float32 if dt == np.float32 else torch.float64
    dev = torch.device("cuda", torch.cuda.current_device())
    t0 = time.perf_counter()
    Ad = torch.from_numpy(np.ascontiguousarray(A, dtype=dt)).to(dev, non_blocking=True)
    bd = torch.from_numpy(np.ascontiguousarray(b, dtype=dt)).to(dev, non_blocking=True)
    x0d = torch.from_numpy(np.ascontiguousarray(x0, dtype=dt)).to(dev, non_blocking=True)
    n_tot = torch.tensor([Ad.shape[1]], device=dev, dtype=torch.int64)
    dist.all_reduce(n_tot)
    n_total = int(n_tot.item())          # also waits for the three uploads (same stream)
    t1 = time.perf_counter()
    prob = GroupLassoProblem(Ad, bd)
    try:
        attach_distributed(prob, n_total)
        t2 = time.perf_counter()
        x, k, out = prob.solve(algo, x0d, mu_0, merged)
        t3 = time.perf_counter()
        xh = x.cpu().numpy()
    finally:
        prob.close()
    t4 = time.perf_counter()
    out["tt"] = t4 - t0
    # where the wall time of a sharded host-array call goes (ms): upload of this rank's shard, handle + mailbox mapping,
    # the solve, download + teardown
    out["glprox"].update(upload_ms=1e3 * (t1 - t0), attach_ms=1e3 * (t2 - t1), solve_ms=1e3 * (t3 - t2),
                         teardown_ms=1e3 * (t4 - t3))
    return xh, k, out


class GroupLassoProblem:
    """A and b resident in HBM behind a libglprox handle (torch provides memory and the stream).

    Lets a caller solve repeatedly without re-uploading A -- the reference re-reads A from host
    memory for every product (gl_ProxGD_primal.py:24-27,60-61)."""

    def __init__(self, A, b, comm=None, n_total=None):
        import torch
        lib = _lib.load()
        if not (isinstance(A, torch.Tensor) and A.is_cuda and A.dim() == 2):
            raise ValueError("A must be a 2-D CUDA tensor")
        if A.dtype not in (torch.float32, torch.float64):
            raise ValueError("A must be float32 or float64")
        if A.stride(1) != 1:
            A = A.contiguous()
        es = A.element_size()
        if (A.stride(0) * es) % 16 != 0 or A.data_ptr() % 16 != 0:
            pad = (-A.shape[1]) % (16 // es)
            Ap = torch.zeros((A.shape[0], A.shape[1] + pad), dtype=A.dtype, device=A.device)
            Ap[:, :A.shape[1]] = A
            A = Ap[:, :A.shape[1]]
        self.A = A
        if not isinstance(b, torch.Tensor) or b.dim() != 2 or b.shape[0] != A.shape[0] or b.shape[1] < 1:
            # the reference fails with NumPy's shape error in A @ x - b (gl_ProxGD_primal.py:25); the C ABI takes raw
            # pointers, so the check has to happen here
            raise ValueError("b must be a 2-D tensor of shape (%d, l), got %s" % (
                A.shape[0], tuple(b.shape) if isinstance(b, torch.Tensor) else type(b).__name__))
        self.b = b.to(device=A.device, dtype=A.dtype).contiguous()
        self.m, self.n = A.shape
        self.l = self.b.shape[1]
        self.dtype = A.dtype
        self.device = A.device
        self.stream = torch.cuda.current_stream(A.device)
        self._h = C.c_void_p()
        with torch.cuda.device(A.device):
            _lib.check(lib.glp_create(C.byref(self._h), A.device.index, self.m, self.n, self.l,
                                      _lib.GLP_F32 if A.dtype == torch.float32 else _lib.GLP_F64,
                                      A.data_ptr(), A.stride(0), self.b.data_ptr(),
                                      comm, self.stream.cuda_stream))

    @property
    def handle(self):
        return self._h

    # ---- multi-GPU: this problem is the column shard A[:, J_rank] of a larger one ------------------
    def export_mailbox(self, world):
        """(64-byte CUDA IPC handle, raw device pointer) of this rank's peer mailbox."""
        lib = _lib.load()
        blob = (C.c_char * 64)()
        ptr = C.c_void_p()
        with self._device_ctx():
            _lib.check(lib.glp_comm_export(self._h, int(world), blob, C.byref(ptr)))
        return bytes(blob.raw), ptr.value

    def attach_peers(self, rank, world, n_total, peers, kind=0):
        """peers: kind 0 -> world*64 bytes of IPC handles in rank order; kind 1 -> list of device pointers."""
        lib = _lib.load()
        if kind == 0:
            if len(peers) != 64 * world:
                raise ValueError("expected %d bytes of IPC handles, got %d" % (64 * world, len(peers)))
            buf = C.create_string_buffer(bytes(peers), 64 * world)
        else:
            buf = (C.c_void_p * world)(*[int(p) for p in peers])
        with self._device_ctx():
            _lib.check(lib.glp_comm_attach(self._h, int(rank), int(world), int(n_total), buf, int(kind)))
        self.rank, self.world, self.n_total = int(rank), int(world), int(n_total)

    def _device_ctx(self):
        import torch
        return torch.cuda.device(self.device)

    def _iterate(self, x, what="x0"):
        """Validate an (n_local, l) iterate and bring it to the handle's device / dtype.  The library reads n*l
        elements from the raw pointer whatever the tensor holds, so a wrong shape must be refused here (the
        reference raises NumPy's broadcasting error at its first product)."""
        import torch
        if not isinstance(x, torch.Tensor):
            raise ValueError("%s must be a torch tensor on the handle path (NumPy arrays take the gl_* entry points)" % what)
        if x.dim() != 2 or tuple(x.shape) != (self.n, self.l):
            raise ValueError("%s must have shape (%d, %d), got %s" % (what, self.n, self.l, tuple(x.shape)))
        return x.to(device=self.device, dtype=self.dtype).contiguous()

    def _order_after_caller(self):
        """Casts / copies of the arguments ran on the caller's current stream; the library launches on the stream
        captured at construction.  Make the latter wait for the former (no-op when they are the same stream)."""
        import torch
        cur = torch.cuda.current_stream(self.device)
        if cur.cuda_stream != self.stream.cuda_stream:
            self.stream.wait_stream(cur)

    def _order_before_caller(self):
        import torch
        cur = torch.cuda.current_stream(self.device)
        if cur.cuda_stream != self.stream.cuda_stream:
            cur.wait_stream(self.stream)

    def close(self):
        if self._h:
            _lib.load().glp_destroy(self._h)
            self._h = C.c_void_p()
        comm = getattr(self, "_nccl_comm", None)
        if comm:
            _lib.load().glp_nccl_comm_destroy(comm)
            self._nccl_comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def lambda_max(self, max_steps=400, tol=1e-14, group=None):
        """Largest eigenvalue of A^T A (= of A A^T), what `np.max(LA.eigvals(A.T @ A))` returns in
        gl_GD_primal.py:44 / gl_SGD_primal.py:36 -- without the dense n x n eigendecomposition: symmetric Lanczos
        on A A^T with full re-orthogonalisation.  The two O(mn) products of a step are the stream kernels
        (glp_tpass_prox: y = A^T u, glp_fwd: A y; partial products of a column shard are summed over `group`), the
        O(m) vector updates run in torch on the device, the k x k tridiagonal eigenvalue problem on the host."""
        import torch
        lib = _lib.load()
        m, n, l = self.m, self.n, self.l
        dev = self.device
        sharded = getattr(self, "world", 1) > 1
        null = C.c_void_p()
        U = torch.zeros((m, l), dtype=torch.float64, device=dev)
        Y = torch.empty((n, l), dtype=torch.float64, device=dev)
        Z = torch.empty((m, l), dtype=torch.float64, device=dev)

        def apply(u):   # A A^T u through the two streams (column 0 of the l-column operands)
            U.zero_()
            U[:, 0] = u
            self._order_after_caller()
            _lib.check(lib.glp_tpass_prox(self._h, None, U.data_ptr(), Y.data_ptr(), null, null, null, null, null, null))
            self._order_before_caller()
            Yt = Y.to(self.dtype).contiguous()
            self._order_after_caller()
            _lib.check(lib.glp_fwd(self._h, Yt.data_ptr(), l, Z.data_ptr(), None))
            self._order_before_caller()
            w = Z[:, 0].clone()
            if sharded:
                import torch.distributed as dist
                dist.all_reduce(w, group=group)
            return w

        with torch.cuda.device(dev):
            v = torch.full((m,), 1.0 / np.sqrt(m), dtype=torch.float64, device=dev)
            v[::2] *= 1.0 + 1e-3            # a fixed, generic start (not orthogonal to the leading eigenvector)
            v /= v.norm()
            V = [v]
            alphas, betas = [], []
            lam_prev = 0.0
            for k in range(min(max_steps, m)):
                w = apply(V[-1])
                a_k = float(torch.dot(w, V[-1]))
                alphas.append(a_k)
                Vm = torch.stack(V, 0)
                w = w - Vm.t() @ (Vm @ w)                 # full re-orthogonalisation (twice is enough)
                w = w - Vm.t() @ (Vm @ w)
                b_k = float(w.norm())
                T = np.diag(alphas) + np.diag(betas, 1) + np.diag(betas, -1)
                lam = float(np.linalg.eigvalsh(T)[-1])
                if k > 0 and (abs(lam - lam_prev) <= tol * abs(lam) or b_k <= 1e-300):
                    break
                lam_prev = lam
                betas.append(b_k)
                V.append(w / b_k)
            return lam

    def solve(self, algo, x0, mu_0, opts=None):
        import torch
        lib = _lib.load()
        merged = merge_opts(algo, opts)
        if algo in ("GD", "SGD") and merged.get("continuous_subgradient_flag"):
            merged["alpha0"] = 1.0 / self.lambda_max()       # gl_GD_primal.py:43-45 / gl_SGD_primal.py:35-37
        o = pack_opts(algo, merged)
        mode = _tensor_mode(merged)
        x0d = self._iterate(x0)
        x = torch.empty_like(x0d)
        cap = 3 * int(o.maxit) + 8
        f_hist = np.empty(cap, dtype=np.float64)
        res = _lib.GlpResult()
        t0 = time.perf_counter()
        with torch.cuda.device(self.device):
            self._order_after_caller()
            _lib.check(lib.glp_set_tensor_mode(self._h, mode))
            rc = lib.glp_solve(self._h, ALGO_ID[algo], x0d.data_ptr(), float(mu_0), C.byref(o), x.data_ptr(),
                               C.byref(res), f_hist.ctypes.data, cap)
        _raise_for(rc, res)
        tt = time.perf_counter() - t0
        fh = f_hist[:res.iters]
        if logger.isEnabledFor(logging.DEBUG):
            s_hist = np.empty(max(int(res.iters), 1), dtype=np.float64)
            got = lib.glp_solve_sparsity_hist(self._h, s_hist.ctypes.data, int(res.iters))
            _log_history(fh, res, merged, float(mu_0), s_hist[:max(got, 0)])
        return x, int(res.iters), _format_out(res, fh, tt)

    def metrics(self, x, ref=None, mu=0.0, group=None):
        """Driver metrics of code/main.py:30-34,48-50 for an iterate that stays on the device.

        x, ref: (n_local, l) CUDA tensors (ref = the ground truth u, or another solver's solution).
        Returns {"objective", "sparsity", "err_exact", "err_ref", "norm_x", ...}.  On a handle attached
        to peers the shard sums are combined over `group` (sum / max of a dozen doubles; the count
        is taken against the global max) and the objective is left to the solve's out["fval"]."""
        import torch
        lib = _lib.load()
        xd = self._iterate(x, "x")
        rd = None if ref is None else self._iterate(ref, "ref")
        self._order_after_caller()
        met = _lib.GlpMetrics()
        call = lambda mx: _lib.check(lib.glp_metrics_eval(self._h, xd.data_ptr(), rd.data_ptr() if rd is not None else None,
                                                          float(mu), float(mx), C.byref(met)))
        with torch.cuda.device(self.device):
            call(-1.0)
            size = float(self.n * self.l)
            sums = [met.sumsq_x, met.sumsq_ref, met.sumsq_diff, met.reg, met.count, size]
            if getattr(self, "world", 1) > 1:
                import torch.distributed as dist
                on = self.device if dist.get_backend(group) == "nccl" else "cpu"
                mx = torch.tensor([met.max_abs], dtype=torch.float64, device=on)
                dist.all_reduce(mx, op=dist.ReduceOp.MAX, group=group)
                call(float(mx.item()))
                t = torch.tensor([met.sumsq_x, met.sumsq_ref, met.sumsq_diff, met.reg, met.count, size],
                                 dtype=torch.float64, device=on)
                dist.all_reduce(t, group=group)
                sums = [float(v) for v in t.cpu()]
                max_abs = float(mx.item())
            else:
                max_abs = met.max_abs
        return combine_metrics(sums, max_abs, None if getattr(self, "world", 1) > 1 else met.rr, mu, ref is not None)

    def solve_admm_dual(self, x0, mu, opts=None):
        """gl_Admm_dual (code/gl_ADMM_dual.py) on this problem; x0 a CUDA tensor, returns (x, k, out)."""
        import torch
        lib = _lib.load()
        merged = dict(ADMM_DUAL_DEFAULT_OPTS)
        merged.update(opts or {})
        o = _lib.GlpAdmmOpts(maxit=int(merged["maxit"]), converge_len=int(merged["converge_len"]),
                             thres=float(merged["thres"]), tau=float(merged["tau"]), rho=float(merged["rho"]))
        x0d = self._iterate(x0)
        x = torch.empty_like(x0d)
        cap = int(o.maxit) + 8
        f_hist = np.empty(cap, dtype=np.float64)
        res = _lib.GlpResult()
        t0 = time.perf_counter()
        with torch.cuda.device(self.device):
            self._order_after_caller()
            _lib.check(lib.glp_admm_dual(self._h, x0d.data_ptr(), float(mu), C.byref(o), x.data_ptr(), C.byref(res),
                                         f_hist.ctypes.data, cap))
        tt = time.perf_counter() - t0
        return x, int(res.iters), _format_out(res, f_hist[:res.iters], tt)


    def _fp64_twin(self):
        """gl_Admm_primal / gl_Alm_dual compute in fp64 (their m x m forms cancel ~1e5 of the leading digits): an
        fp32 problem is widened into a temporary fp64 one for the call."""
        import torch
        if self.dtype == torch.float64:
            return self, False
        if getattr(self, "world", 1) > 1:
            raise ValueError("gl_Admm_primal / gl_Alm_dual run on one GPU")
        return GroupLassoProblem(self.A.double(), self.b.double()), True

    def _solve_admm2(self, which, x0, mu, o, cap):
        import torch
        lib = _lib.load()
        prob, temp = self._fp64_twin()
        try:
            x0d = prob._iterate(x0.double() if isinstance(x0, torch.Tensor) else x0)
            x = torch.empty_like(x0d)
            f_hist = np.empty(cap, dtype=np.float64)
            res = _lib.GlpResult()
            t0 = time.perf_counter()
            with torch.cuda.device(prob.device):
                prob._order_after_caller()
                fn = lib.glp_admm_primal if which == "ADMM_primal" else lib.glp_alm_dual
                rc = fn(prob._h, x0d.data_ptr(), float(mu), C.byref(o), x.data_ptr(), C.byref(res), f_hist.ctypes.data, cap)
                _raise_for(rc, res)
            tt = time.perf_counter() - t0
        finally:
            if temp:
                prob.close()
        return x.to(self.dtype), int(res.iters), _format_out(res, f_hist[:res.iters], tt)

    def solve_admm_primal(self, x0, mu, opts=None):
        """gl_Admm_primal (code/gl_ADMM_primal.py) on this problem; x0 a CUDA tensor, returns (x, k, out)."""
        merged = dict(ADMM_PRIMAL_DEFAULT_OPTS)
        merged.update(opts or {})
        steps = {k: _lib.STEP_TYPES[k] for k in ("fixed", "diminishing", "diminishing2")}
        if merged["step_type"] not in steps:
            # the reference's set_step returns None for any other name and the next line raises TypeError (:66-80)
            raise TypeError("unsupported operand type(s) for *: 'NoneType' and 'float' (step_type %r)" % (merged["step_type"],))
        o = _lib.GlpAdmmPrimalOpts(maxit=int(merged["maxit"]), converge_len=int(merged["converge_len"]),
                                   step_type=steps[merged["step_type"]], pad=0, thres=float(merged["thres"]),
                                   tau=float(merged["tau"]), rho=float(merged["rho"]), eta_0=float(merged["eta_0"]))
        return self._solve_admm2("ADMM_primal", x0, mu, o, int(o.maxit) + 8)

    def solve_alm_dual(self, x0, mu, opts=None):
        """gl_Alm_dual (code/gl_ALM_dual.py) on this problem; x0 a CUDA tensor, returns (x, k, out)."""
        merged = dict(ALM_DUAL_DEFAULT_OPTS)
        merged.update(opts or {})
        o = _lib.GlpAdmmOpts(maxit=int(merged["maxit"]), converge_len=int(merged["converge_len"]),
                             thres=float(merged["thres"]), tau=float(merged["tau"]), rho=float(merged["rho"]))
        return self._solve_admm2("ALM_dual", x0, mu, o, int(o.maxit) + 8)


def combine_metrics(sums, max_abs, rr, mu, has_ref):
    """Finish the metrics from (sumsq_x, sumsq_ref, sumsq_diff, reg, count, size) totals."""
    sx, sr, sd, reg, cnt, size = sums
    nan = float("nan")
    nx, nr, nd = np.sqrt(sx), np.sqrt(sr), np.sqrt(sd)
    return {"objective": nan if rr is None else 0.5 * rr + mu * reg,
            "sparsity": cnt / size if size else nan,
            "err_exact": nd / (1.0 + nx) if has_ref else nan,   # main.py:49
            "err_ref": nd / (1.0 + nr) if has_ref else nan,     # main.py:48 errfun(ref, x)
            "norm_x": float(nx), "norm_ref": float(nr), "norm_diff": float(nd), "reg": reg,
            "max_abs": max_abs, "count": int(cnt)}


def solve_admm_dual(x0, A, b, mu, opts):
    """Drop-in body of gl_Admm_dual: NumPy arrays (uploaded for the call) or CUDA tensors."""
    merged = dict(ADMM_DUAL_DEFAULT_OPTS)
    merged.update(opts or {})
    import torch
    if isinstance(A, torch.Tensor):
        if not A.is_cuda:
            raise ValueError("torch tensors must live on a CUDA device (NumPy arrays take the host path)")
        prob = GroupLassoProblem(A, b)
        try:
            return prob.solve_admm_dual(x0, mu, merged)
        finally:
            prob.close()
    A = np.asarray(A)
    x0, b = np.asarray(x0), np.asarray(b)
    if A.ndim != 2 or x0.ndim != 2 or b.ndim != 2 or x0.shape[0] != A.shape[1] or b.shape[0] != A.shape[0] or x0.shape[1] != b.shape[1]:
        raise ValueError("shape mismatch: A %s, x0 %s, b %s" % (A.shape, x0.shape, b.shape))
    dt = _mode_dtype(A.dtype, merged)
    dev = torch.device("cuda", int(merged.get("device", 0)))
    t0 = time.perf_counter()
    to = lambda z: torch.from_numpy(np.ascontiguousarray(z, dtype=dt)).to(dev)
    prob = GroupLassoProblem(to(A), to(b))
    try:
        x, k, out = prob.solve_admm_dual(to(x0), mu, merged)
        xh = x.cpu().numpy()
    finally:
        prob.close()
    out["tt"] = time.perf_counter() - t0
    return xh, k, out


def _solve_fp64_dropin(method, defaults, x0, A, b, mu, opts):
    """Drop-in body shared by gl_Admm_primal and gl_Alm_dual: NumPy arrays (uploaded as fp64 for the call) or CUDA
    tensors; the result comes back in the dtype of A."""
    merged = dict(defaults)
    merged.update(opts or {})
    import torch
    if isinstance(A, torch.Tensor):
        if not A.is_cuda:
            raise ValueError("torch tensors must live on a CUDA device (NumPy arrays take the host path)")
        prob = GroupLassoProblem(A, b)
        try:
            return getattr(prob, method)(x0, mu, merged)
        finally:
            prob.close()
    A = np.asarray(A)
    x0, b = np.asarray(x0), np.asarray(b)
    if A.ndim != 2 or x0.ndim != 2 or b.ndim != 2 or x0.shape[0] != A.shape[1] or b.shape[0] != A.shape[0] or x0.shape[1] != b.shape[1]:
        raise ValueError("shape mismatch: A %s, x0 %s, b %s" % (A.shape, x0.shape, b.shape))
    dev = torch.device("cuda", int(merged.get("device", 0)))
    t0 = time.perf_counter()
    to = lambda z: torch.from_numpy(np.ascontiguousarray(z, dtype=np.float64)).to(dev)
    prob = GroupLassoProblem(to(A), to(b))
    try:
        x, k, out = getattr(prob, method)(to(x0), mu, merged)
        xh = x.cpu().numpy().astype(A.dtype if A.dtype in (np.float32, np.float64) else np.float64, copy=False)
    finally:
        prob.close()
    out["tt"] = time.perf_counter() - t0
    return xh, k, out


def solve_admm_primal(x0, A, b, mu, opts):
    """Drop-in body of gl_Admm_primal (code/gl_ADMM_primal.py:10)."""
    return _solve_fp64_dropin("solve_admm_primal", ADMM_PRIMAL_DEFAULT_OPTS, x0, A, b, mu, opts)


def solve_alm_dual(x0, A, b, mu, opts):
    """Drop-in body of gl_Alm_dual (code/gl_ALM_dual.py:66)."""
    return _solve_fp64_dropin("solve_alm_dual", ALM_DUAL_DEFAULT_OPTS, x0, A, b, mu, opts)


def shard_bounds(n, world, align=128):
    """Contiguous column blocks J_0..J_{world-1} of [0, n): sizes differ by at most `align` and all
    but the last are multiples of `align` (TMA tile width in elements), so that every x row group
    lives on exactly one rank and the prox stays rank-local."""
    if world < 1:
        raise ValueError("world must be >= 1")
    blocks = -(-n // align)
    cuts = [min(n, ((r * blocks) // world) * align) for r in range(world)] + [n]
    if any(cuts[r + 1] <= cuts[r] for r in range(world)):
        raise ValueError("n=%d is too small to shard over %d ranks in blocks of %d columns" % (n, world, align))
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def exchange_handles(mine, group=None):
    """All-gather one fixed-size byte string per rank over torch.distributed (any backend); returns
    the concatenation in rank order.  This is the only host-side communication of a sharded solve."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor(list(mine), dtype=torch.uint8, device=dev)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    return b"".join(bytes(o.cpu().numpy().tobytes()) for o in out)


def attach_distributed(prob, n_total, group=None, exchange="peer"):
    """One process per GPU: export this rank's mailbox, all-gather the IPC handles, map the peers.

    exchange="nccl" selects the baseline exchange instead (two ncclAllReduce calls per forward stream on a
    communicator created here from an id broadcast over `group`); it exists for A/B measurements."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if exchange == "nccl":
        return _attach_nccl(prob, n_total, rank, world, group)
    if exchange != "peer":
        raise ValueError("exchange must be 'peer' or 'nccl'")
    blob, _ = prob.export_mailbox(world)
    prob.attach_peers(rank, world, n_total, exchange_handles(blob, group), kind=0)
    dist.barrier(group)  # every mailbox is mapped everywhere before the first push
    return prob


def _attach_nccl(prob, n_total, rank, world, group):
    import torch
    import torch.distributed as dist
    lib = _lib.load()
    ident = (C.c_char * 128)()
    if rank == 0:
        _lib.check(lib.glp_nccl_unique_id(ident))
    on = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor(list(ident.raw), dtype=torch.uint8, device=on)
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    ident = (C.c_char * 128).from_buffer_copy(bytes(t.cpu().numpy().tobytes()))
    comm = C.c_void_p()
    with prob._device_ctx():
        _lib.check(lib.glp_nccl_comm_create(C.byref(comm), ident, int(rank), int(world), prob.device.index))
        _lib.check(lib.glp_comm_attach_nccl(prob.handle, comm, int(rank), int(world), int(n_total)))
    prob.rank, prob.world, prob.n_total = int(rank), int(world), int(n_total)
    prob._nccl_comm = comm
    dist.barrier(group)
    return prob


def solve_group(probs, algo, x0_shards, mu_0, opts=None):
    """One process driving G shards (G devices, or G logical ranks on one device + stream).
    Returns (list of x shards, k, out)."""
    import torch
    lib = _lib.load()
    G = len(probs)
    merged = merge_opts(algo, opts)
    o = pack_opts(algo, merged)
    if len(x0_shards) != G:
        raise ValueError("expected %d x0 shards, got %d" % (G, len(x0_shards)))
    x0d = [p._iterate(x0, "x0 shard %d" % i) for i, (p, x0) in enumerate(zip(probs, x0_shards))]
    for p in probs:
        p._order_after_caller()
    xs = [torch.empty_like(x) for x in x0d]
    cap = 3 * int(o.maxit) + 8
    f_hist = np.empty(cap, dtype=np.float64)
    res = (_lib.GlpResult * G)()
    hs = (C.c_void_p * G)(*[p.handle.value for p in probs])
    x0p = (C.c_void_p * G)(*[x.data_ptr() for x in x0d])
    xp = (C.c_void_p * G)(*[x.data_ptr() for x in xs])
    t0 = time.perf_counter()
    rc = lib.glp_group_solve(hs, G, ALGO_ID[algo], x0p, float(mu_0), C.byref(o), xp, res, f_hist.ctypes.data, cap)
    _raise_for(rc, res[0])
    tt = time.perf_counter() - t0
    fh = f_hist[:res[0].iters]
    out = _format_out(res[0], fh, tt)
    out["glprox"]["per_rank"] = [{"iters": int(r.iters), "fval": r.fval, "n_fwd": int(r.n_fwd)} for r in res]
    return xs, int(res[0].iters), out


def _solve_torch(algo, x0, A, b, mu_0, merged, o):
    import torch
    if not A.is_cuda:
        raise ValueError("torch tensors must live on a CUDA device (NumPy arrays take the host path)")
    want = _mode_dtype(np.float32 if A.dtype == torch.float32 else np.float64, merged)
    tdt = torch.float32 if want == np.float32 else torch.float64
    prob = GroupLassoProblem(A.to(tdt), b)
    try:
        x, k, out = prob.solve(algo, x0, mu_0, merged)
    finally:
        prob.close()
    return x, k, out
