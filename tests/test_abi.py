"""CPU checks of the drop-in boundary: libglprox.so loads without a GPU, exports every symbol that
include/glprox.h declares, the POD mirrors have the C layout, and the host-side option semantics
follow the reference.  No compute call is made here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from glprox import _lib
    if not os.path.exists(_lib.LIB_PATH):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "convex-optimization_b200", "csrc")])
    return _lib.load()


def test_every_declared_symbol_is_exported_and_bound(lib):
    from glprox import _lib
    hdr = open(os.path.join(ROOT, "include", "glprox.h")).read()
    declared = set(re.findall(r"^\s*(?:int|const char\*)\s+(glp_\w+)\s*\(", hdr, flags=re.M))
    assert len(declared) >= 18
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name)
    assert lib.glp_version() == int(re.search(r"#define GLP_VERSION (\d+)", hdr).group(1))


def test_pod_layouts_match_the_header(lib):
    from glprox import _lib
    assert C.sizeof(_lib.GlpOpts) == 64 and C.sizeof(_lib.GlpStepParams) == 40
    assert C.sizeof(_lib.GlpResult) == 80
    for algo, maxit, alpha0, att in ((0, 2500, 2e-3, 0.9), (1, 1500, 1e-3, 0.98)):
        o = _lib.GlpOpts()
        assert lib.glp_default_opts(algo, C.byref(o)) == 0
        # gl_ProxGD_primal.py:10-19 / gl_FProxGD_primal.py:10-19
        assert (o.maxit, o.alpha0, o.line_search_attenuation_coeffi) == (maxit, alpha0, att)
        assert (o.thres, o.ftol, o.stable_len_threshold, o.maxit_line_search_iter, o.step_type) == (1e-3, 1e-6, 70, 5, 0)


def test_no_cpu_fallback(lib):
    """Without a device every computing entry point fails loudly (this test runs on the CPU box)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from gl_FProxGD_primal import gl_FProxGD_primal
    from glprox import _lib
    assert lib.glp_device_count() == 0
    with pytest.raises(_lib.GlproxError) as ei:
        gl_FProxGD_primal(np.zeros((8, 2)), np.zeros((4, 8)), np.zeros((4, 2)), 1e-2, {})
    assert ei.value.code == _lib.GLP_ERR_NO_DEVICE


def test_option_semantics_mirror_the_reference():
    from glprox import solver
    from oracle import glprox_oracle as orc
    for algo in ("ProxGD", "FProxGD"):
        assert solver.DEFAULT_OPTS[algo] == orc.DEFAULTS[algo]
        m = solver.merge_opts(algo, {"maxit": 7, "bogus": 1})
        assert m["maxit"] == 7 and m["thres"] == 1e-3 and m["bogus"] == 1
        o = solver.pack_opts(algo, m)
        assert o.maxit == 7 and o.step_type == 0
        with pytest.raises(ValueError):
            solver.pack_opts(algo, solver.merge_opts(algo, {"step_type": "no_such_rule"}))


def test_additive_keys_are_validated_before_any_launch():
    """tensor_cores / backend / dtype are additive keys (not in the reference): a bad value is a ValueError raised on the
    host, on every entry path, before the library is asked to compute anything."""
    from gl_FProxGD_primal import gl_FProxGD_primal
    from glprox import solver
    assert solver._tensor_mode({}) == 0 and solver._tensor_mode({"tensor_cores": "f16x2"}) == 4
    assert sorted(solver.TENSOR_MODES.values()) == [0, 1, 2, 3, 4, 5]   # the modes of glp_set_tensor_mode (include/glprox.h)
    x0, A, b = np.zeros((8, 2)), np.zeros((4, 8)), np.zeros((4, 2))
    for bad in ({"tensor_cores": "fp8"}, {"backend": "cpu"}, {"dtype": "bf16"}):
        with pytest.raises(ValueError):
            gl_FProxGD_primal(x0, A, b, 1e-2, bad)


def test_product_path_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "convex-optimization_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inl")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "/root/reference" not in src, f
