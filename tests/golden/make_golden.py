#!/usr/bin/env python
"""Generate golden vectors by running the UNMODIFIED reference solvers.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
Writes tests/golden/ref_*.npz.  The reference functions are imported from
/root/reference/code (gl_ProxGD_primal.py, gl_FProxGD_primal.py); the inputs come from
oracle.glprox_oracle.gen_data, which restates code/main.py:37-51 (main.py itself is not
importable: it needs cvxpy/matplotlib/mosek/gurobi).
"""
import os
import sys

import numpy as np

sys.dont_write_bytecode = True  # /root/reference is a read-only mount
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference/code")

from gl_ADMM_dual import gl_Admm_dual  # noqa: E402  (reference)
from gl_ADMM_primal import gl_Admm_primal  # noqa: E402  (reference)
from gl_ALM_dual import gl_Alm_dual  # noqa: E402  (reference)
from gl_FGD_primal import gl_FGD_primal  # noqa: E402  (reference)
from gl_FProxGD_primal import gl_FProxGD_primal  # noqa: E402  (reference)
from gl_GD_primal import gl_GD_primal  # noqa: E402  (reference)
from gl_ProxGD_primal import gl_ProxGD_primal  # noqa: E402  (reference)
from gl_SGD_primal import gl_SGD_primal  # noqa: E402  (reference)
from oracle.glprox_oracle import gen_data, sparsity  # noqa: E402

REF = {"ProxGD": gl_ProxGD_primal, "FProxGD": gl_FProxGD_primal}
REF_SMOOTH = {"FGD": gl_FGD_primal, "GD": gl_GD_primal, "SGD": gl_SGD_primal}


def mp_step(m, n):
    """Marchenko-Pastur bound step used for every synthetic shape."""
    return float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)


# name -> (seed, m, n, l, dtype, opts)
CASES = {
    "fixture": (97006855, 256, 512, 2, "f8", {}),
    "fixture_maxit3": (97006855, 256, 512, 2, "f8", {"maxit": 3}),
    "fixture_maxit40": (97006855, 256, 512, 2, "f8", {"maxit": 40}),
    "seed114514": (114514, 256, 512, 2, "f8", {}),
    "seed1": (1, 256, 512, 2, "f8", {}),
    "seed2": (2, 256, 512, 2, "f8", {}),
    "fixed_step": (97006855, 256, 512, 2, "f8", {"step_type": "fixed", "alpha0": 5e-4}),
    "diminishing": (97006855, 256, 512, 2, "f8", {"step_type": "diminishing", "alpha0": 5e-4,
                                                  "maxit": 1200}),
    "diminishing2": (97006855, 256, 512, 2, "f8", {"step_type": "diminishing2", "alpha0": 5e-4,
                                                   "maxit": 1200}),
    "thres1e-8": (97006855, 256, 512, 2, "f8", {"thres": 1e-8}),
    "ragged": (5, 193, 391, 2, "f8", {"alpha0": mp_step(193, 391), "maxit": 60}),
    "l1": (6, 128, 300, 1, "f8", {"alpha0": mp_step(128, 300), "maxit": 60}),
    "l3": (8, 96, 260, 3, "f8", {"alpha0": mp_step(96, 260), "maxit": 60}),
    "l5": (9, 64, 256, 5, "f8", {"alpha0": mp_step(64, 256), "maxit": 40}),
    "mid": (11, 1024, 4096, 2, "f8", {"alpha0": mp_step(1024, 4096), "maxit": 50}),
    "harsh_ls": (97006855, 256, 512, 2, "f8", {"alpha0": 1.2e-3, "maxit": 80}),
    # fp32 configs: the oracle is the fp64 reference on the fp32-rounded inputs
    "mid_f32in": (11, 1024, 4096, 2, "f4", {"alpha0": mp_step(1024, 4096), "maxit": 50}),
    "fixture_f32in": (97006855, 256, 512, 2, "f4", {}),
}


# cases for the smoothed / subgradient solvers; per-algorithm opts where the defaults differ too much
SMOOTH_CASES = {
    "smooth_fixture": (97006855, 256, 512, 2, "f8", {}),
    "smooth_maxit40": (97006855, 256, 512, 2, "f8", {"maxit": 40}),
    "smooth_seed2": (2, 256, 512, 2, "f8", {"maxit": 700}),
    "smooth_ragged": (5, 193, 391, 2, "f8", {"alpha0": mp_step(193, 391), "maxit": 60}),
    "smooth_l3": (8, 96, 260, 3, "f8", {"alpha0": mp_step(96, 260), "maxit": 60}),
    "smooth_fixed": (97006855, 256, 512, 2, "f8", {"step_type": "fixed", "alpha0": 5e-4, "maxit": 400}),
    "smooth_dim2": (97006855, 256, 512, 2, "f8", {"step_type": "diminishing2", "alpha0": 5e-4, "maxit": 1100}),
    "smooth_f32in": (97006855, 256, 512, 2, "f4", {"maxit": 200}),
    # gl_GD_primal.py:43-45 / gl_SGD_primal.py:35-37: alpha0 = 1 / lambda_max(A^T A)
    "smooth_csf": (97006855, 256, 512, 2, "f8", {"continuous_subgradient_flag": True, "maxit": 300}),
    "smooth_csf_ragged": (5, 193, 391, 2, "f8", {"continuous_subgradient_flag": True, "maxit": 80, "step_type": "fixed"}),
}


# gl_Admm_dual (no continuation; its own option set)
ADMM_CASES = {
    "admm_fixture": (97006855, 256, 512, 2, "f8", {}),
    "admm_seed1": (1, 256, 512, 2, "f8", {}),
    "admm_seed2": (2, 256, 512, 2, "f8", {}),
    "admm_maxit7": (97006855, 256, 512, 2, "f8", {"maxit": 7}),
    "admm_ragged": (5, 193, 391, 2, "f8", {"maxit": 60}),
    "admm_l1": (6, 128, 300, 1, "f8", {"maxit": 60}),
    "admm_rho10": (97006855, 256, 512, 2, "f8", {"rho": 10.0, "tau": 1.0, "maxit": 150, "converge_len": 5}),
    "admm_loose": (3, 96, 300, 2, "f8", {"thres": 1e-1, "converge_len": 3}),
    "admm_f32in": (97006855, 256, 512, 2, "f4", {}),
}
REF_ADMM = {"ADMM_dual": gl_Admm_dual}

# gl_Admm_primal (code/gl_ADMM_primal.py) and gl_Alm_dual (code/gl_ALM_dual.py): the two remaining solvers of the
# reference's registry (code/main.py:165-183)
ADMMP_CASES = {
    "admmp_fixture": (97006855, 256, 512, 2, "f8", {}),
    "admmp_seed1": (1, 256, 512, 2, "f8", {}),
    "admmp_maxit7": (97006855, 256, 512, 2, "f8", {"maxit": 7}),
    "admmp_ragged": (5, 193, 391, 2, "f8", {}),
    "admmp_l1": (6, 128, 300, 1, "f8", {}),
    "admmp_l3": (8, 96, 260, 3, "f8", {}),
    "admmp_dimin": (97006855, 256, 512, 2, "f8", {"step_type": "diminishing", "maxit": 60}),
    "admmp_dimin2": (2, 256, 512, 2, "f8", {"step_type": "diminishing2", "maxit": 40}),
    "admmp_rho1": (3, 96, 300, 2, "f8", {"rho": 1.0, "eta_0": 1.0, "tau": 1.0, "thres": 1e-2, "converge_len": 3, "maxit": 200}),
    "admmp_f32in": (97006855, 256, 512, 2, "f4", {}),
}
REF_ADMMP = {"ADMM_primal": gl_Admm_primal}
ALM_CASES = {
    "alm_fixture": (97006855, 256, 512, 2, "f8", {}),
    "alm_maxit3": (1, 256, 512, 2, "f8", {"maxit": 3}),
    "alm_ragged": (5, 193, 391, 2, "f8", {"maxit": 8}),
    "alm_l1": (6, 128, 300, 1, "f8", {"maxit": 6}),
    "alm_l3": (8, 96, 260, 3, "f8", {"maxit": 6}),
    "alm_loose": (3, 96, 300, 2, "f8", {"thres": 1e-1, "converge_len": 3, "rho": 10.0, "tau": 1.0}),
    "alm_f32in": (97006855, 256, 512, 2, "f4", {"maxit": 5}),
}
REF_ALM = {"ALM_dual": gl_Alm_dual}


def run_case(name):
    if name in ADMMP_CASES:
        seed, m, n, l, dt, opts = ADMMP_CASES[name]
        return _run(name, seed, m, n, l, dt, opts, REF_ADMMP)
    if name in ALM_CASES:
        seed, m, n, l, dt, opts = ALM_CASES[name]
        return _run(name, seed, m, n, l, dt, opts, REF_ALM)
    if name in ADMM_CASES:
        seed, m, n, l, dt, opts = ADMM_CASES[name]
        return _run(name, seed, m, n, l, dt, opts, REF_ADMM)
    if name in SMOOTH_CASES:
        seed, m, n, l, dt, opts = SMOOTH_CASES[name]
        algos = REF_SMOOTH
    else:
        seed, m, n, l, dt, opts = CASES[name]
        algos = REF
    return _run(name, seed, m, n, l, dt, opts, algos)


def _run(name, seed, m, n, l, dt, opts, algos):
    A, b, x0, u, mu = gen_data(seed, m, n, l)
    if dt == "f4":  # round the inputs to fp32, then run the reference in fp64 on them
        A, b, x0 = (z.astype(np.float32).astype(np.float64) for z in (A, b, x0))
    rec = {"seed": seed, "m": m, "n": n, "l": l, "in_dtype": dt, "mu": mu,
           "opts_keys": np.array(sorted(opts.keys())),
           "opts_vals": np.array([str(opts[k]) for k in sorted(opts.keys())])}
    for algo, fn in algos.items():
        A_keep, b_keep, x0_keep = A.copy(), b.copy(), x0.copy()
        x, k, out = fn(x0, A, b, mu, dict(opts))
        assert np.array_equal(A, A_keep) and np.array_equal(b, b_keep) and np.array_equal(x0, x0_keep)
        ax = np.abs(x)
        rec[algo + "_k"] = k
        rec[algo + "_fval"] = float(out["fval"])
        rec[algo + "_f_hist"] = np.asarray(out["f_hist"], dtype=np.float64)
        rec[algo + "_f_hist_best"] = np.asarray(out["f_hist_best"], dtype=np.float64)
        rec[algo + "_x"] = x
        rec[algo + "_support_nz"] = np.flatnonzero(np.sqrt((x * x).sum(1)) > 0)
        rec[algo + "_support_rule"] = np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1))
        rec[algo + "_sparsity"] = float(sparsity(x))
        print("%-16s %-8s k=%5d fval=%.16e sparsity=%.6f" % (name, algo, k, out["fval"], sparsity(x)))
    return rec


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES) + list(SMOOTH_CASES) + list(ADMM_CASES) + list(ADMMP_CASES) + list(ALM_CASES)
    for nm in names:
        np.savez_compressed(os.path.join(HERE, "ref_%s.npz" % nm), **run_case(nm))
