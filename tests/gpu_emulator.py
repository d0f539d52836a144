"""CPU emulation of the GPU schedule for host-logic tests (no GPU, no CUDA).

The control flow is the *real* one: csrc/glp_decide.h compiled for the host (tests/host_shim).
The O(mn) products are NumPy, the row-group epilogue is oracle.prox_epilogue.  What this proves on
the CPU box: the device state machine + double-buffered commit protocol reproduce the reference.
"""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

from oracle import glprox_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "convex-optimization_b200", "csrc")


class Ctrl(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("done", "stage", "inner", "k", "stable", "trial", "cur", "retry",
                                          "tested", "first_commit", "ran_any", "fault")] + \
               [("stage_end", C.c_int32 * 3), ("gy_sel", C.c_int32)] + \
               [(n, C.c_double) for n in ("t", "t_k", "theta", "mu", "f_cur", "s_cur", "f_prev", "s_prev",
                                          "f_best", "g_base")] + \
               [("n_fwd", C.c_int64), ("n_tpass", C.c_int64)]


class SolveConst(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("algo", "step_type", "maxit", "stable_cap", "ls_cap", "hist_cap")] + \
               [(n, C.c_double) for n in ("thres", "alpha0", "ftol", "atten", "mu0", "delta")]


class Reduced(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("rr", "gd", "dd", "reg", "cnt", "gbase", "sx", "sy")]


class DecideOut(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("action", "reset_v", "first_commit", "new_cur")] + [("theta_old", C.c_double)]


ACT_NONE, ACT_RETRY, ACT_ITERATE, ACT_FINISH = 0, 1, 2, 3
STEP_ID = {"line_search": 0, "fixed": 1, "diminishing": 2, "diminishing2": 3, "bb": 4}
ALGO_ID = {"ProxGD": 0, "FProxGD": 1, "FGD": 2, "GD": 3, "SGD": 4}
_shim = None


def shim():
    global _shim
    if _shim is None:
        out = os.path.join(tempfile.mkdtemp(prefix="glp_shim_"), "decide_shim.so")
        subprocess.check_call(["g++", "-O1", "-ffp-contract=off", "-shared", "-fPIC", "-I", CSRC, "-o", out,
                               os.path.join(ROOT, "tests", "host_shim", "decide_shim.cpp")])
        lib = C.CDLL(out)
        assert lib.shim_sizeof_ctrl() == C.sizeof(Ctrl)
        for nm in ("shim_tile_lo", "shim_tile_owner", "shim_contrib_first", "shim_contrib_last"):
            getattr(lib, nm).restype = C.c_longlong
            getattr(lib, nm).argtypes = [C.c_longlong] * (3 if nm in ("shim_tile_lo", "shim_tile_owner") else 4)
        lib.shim_decide.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p]
        _shim = lib
    return _shim


def emulate(algo, x0, A, b, mu0, opts=None, matmul=None, rmatmul=None, n_shards=1, decide_log=None):
    """Run the GPU schedule on the CPU.  Returns (x, k, out, info).  decide_log, if a list, receives one
    (SolveConst, Ctrl in, Reduced, size_total, Ctrl out, DecideOut, (f, f_best) recorded or None) per decide step."""
    lib = shim()
    o = dict(orc.DEFAULTS[algo])
    o.update(opts or {})
    if algo in ("GD", "SGD") and o.get("continuous_subgradient_flag"):
        # host prologue of GroupLassoProblem.solve: alpha0 <- 1 / lambda_max(A^T A) (= lambda_max(A A^T))
        o["alpha0"] = 1.0 / float(np.max(np.linalg.eigvalsh(A.astype(np.float64) @ A.astype(np.float64).T)))
    fwd = matmul or (lambda X: A @ X)
    bwd = rmatmul or (lambda R: A.T @ R)
    n, l = x0.shape
    accel = algo in ("FProxGD", "FGD")
    delta = float(o.get("delta", 0.0))
    sc = SolveConst(algo=ALGO_ID[algo], step_type=STEP_ID[o["step_type"]], maxit=o["maxit"],
                    stable_cap=o["stable_len_threshold"], ls_cap=o.get("maxit_line_search_iter", 0),
                    hist_cap=3 * o["maxit"] + 8, thres=o["thres"], alpha0=o["alpha0"], ftol=o["ftol"],
                    atten=o.get("line_search_attenuation_coeffi", 1.0), mu0=mu0, delta=delta)
    f_hist = np.zeros(sc.hist_cap)
    f_best = np.zeros(sc.hist_cap)
    ctrl = [Ctrl(), Ctrl()]
    lib.shim_ctrl_init(C.byref(ctrl[0]), C.byref(sc))
    ctrl[1] = Ctrl.from_buffer_copy(ctrl[0])
    # double-buffered state {x, xt, v} and products {Ax, Axt}; Av and r single
    st = [dict(x=None, xt=None, v=None), dict(x=None, xt=None, v=None)]
    AX = [dict(Ax=None, Axt=None), dict(Ax=None, Axt=None)]
    Av = r = grad = None
    gy = [0.0, 0.0]
    # init: candidate of cur=0 is slot 1
    st[1]["x"] = np.array(x0, dtype=np.float64)
    st[1]["xt"] = orc.hard_threshold(st[1]["x"], o["thres"])
    scal = dict(gd=0.0, dd=0.0, reg=float(np.sum(orc.row_norms(st[1]["x"]))), mx=float(np.max(np.abs(st[1]["x"]))),
                sx=0.0, sy=0.0)

    def forward(cand):
        prod = fwd(np.concatenate([st[cand]["x"], st[cand]["xt"]], axis=1))
        AX[cand]["Ax"], AX[cand]["Axt"] = prod[:, :l].copy(), prod[:, l:].copy()
        rr = float(np.sum((AX[cand]["Ax"] - b) ** 2))
        cnt = float(np.sum(np.abs(st[cand]["x"]) > 1e-6 * scal["mx"]))
        return rr, cnt

    rr, cnt = forward(1)
    j = 0
    n_slots = 0
    while True:
        cin, cout = ctrl[j & 1], ctrl[(j + 1) & 1]
        red = Reduced(rr=rr, gd=scal["gd"], dd=scal["dd"], reg=scal["reg"], cnt=cnt, gbase=gy[cin.gy_sel],
                      sx=scal["sx"], sy=scal["sy"])
        act = DecideOut()
        k_before = cin.k
        cin_copy = Ctrl.from_buffer_copy(cin)
        lib.shim_decide(C.byref(sc), C.byref(cin), C.byref(red), float(n * l), C.byref(cout), C.byref(act),
                        f_hist.ctypes.data, f_best.ctypes.data)
        if decide_log is not None:
            rec = (float(f_hist[k_before]), float(f_best[k_before])) if cout.k > k_before else None
            decide_log.append((sc, cin_copy, Reduced.from_buffer_copy(red), float(n * l), Ctrl.from_buffer_copy(cout),
                               DecideOut.from_buffer_copy(act), rec))
        if act.action == ACT_ITERATE:
            cout.gy_sel = 1 - cin.gy_sel
        j += 1
        n_slots += 1
        if act.action in (ACT_FINISH, ACT_NONE):
            break
        cur = cout.cur
        if act.action == ACT_ITERATE:
            nc = act.new_cur
            Axn, Axnt, Axt_old = AX[nc]["Ax"], AX[nc]["Axt"], AX[1 - nc]["Axt"]
            if accel:
                if act.reset_v:
                    Av = Axn.copy()
                    st[nc]["v"] = st[nc]["x"].copy()
                else:
                    Av = Axt_old + (Axn - Axt_old) / act.theta_old
                base = (1 - cout.theta) * Axnt + cout.theta * Av
            else:
                base = Axnt
            r = base - b
            gy[cout.gy_sel] = float(np.sum(r ** 2))
            grad = bwd(r)
        # transposed epilogue (fresh or retry) writes the candidates of slot 1-cur
        xn, xnt, vn, gd, dd, reg, mx, sx, sy = orc.prox_epilogue(algo, grad, st[cur]["xt"], st[cur]["v"], cout.theta,
                                                                 cout.t, cout.mu, o["thres"], delta, full=True)
        st[1 - cur].update(x=xn, xt=xnt, v=vn)
        scal = dict(gd=float(gd), dd=float(dd), reg=float(reg), mx=float(mx), sx=float(sx), sy=float(sy))
        rr, cnt = forward(1 - cur)
    fin = ctrl[j & 1]
    x = st[fin.cur]["x"]
    out = {"fval": fin.f_cur, "f_hist": list(f_hist[:fin.k]), "f_hist_best": list(f_best[:fin.k])}
    info = dict(n_slots=n_slots, n_fwd=fin.n_fwd, n_tpass=fin.n_tpass, ran_any=fin.ran_any,
                stage_end=list(fin.stage_end), sparsity=fin.s_cur)
    return x, fin.k, out, info
