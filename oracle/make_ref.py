"""Recipe that stages the UNMODIFIED reference solvers of the hot path under oracle/_ref/ (TEST INFRASTRUCTURE).

    python oracle/make_ref.py            # run in the build container, where /root/reference exists

oracle/_ref/ is git-ignored (reference sources never enter the history) but not gpurun-ignored, so the staged
files travel to the GPU box next to the built .so files.  There `bench.py --impl reference` and the
`cpu_baseline` leg time THESE functions (kind "reference") on the box's host cores; when the directory is
absent they fall back to the oracle port (kind "port").  Nothing under convex-optimization_b200/ reads it.

Staged, byte for byte: the two solvers of the path and what they import
(code/gl_ProxGD_primal.py, code/gl_FProxGD_primal.py, code/utils/__init__.py, code/utils/stopwatch.py).
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("GLPROX_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ["gl_ProxGD_primal.py", "gl_FProxGD_primal.py", os.path.join("utils", "__init__.py"),
         os.path.join("utils", "stopwatch.py")]


def stage(verbose=True):
    """Copy the files; returns the manifest {relative path: sha256} or None when the reference is absent."""
    src_dir = os.path.join(REF_ROOT, "code")
    if not os.path.isdir(src_dir):
        return None
    manifest = {}
    for rel in FILES:
        dst = os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(src_dir, rel), dst)
        manifest[rel] = hashlib.sha256(open(dst, "rb").read()).hexdigest()
    json.dump({"source": src_dir, "files": manifest}, open(os.path.join(DST, "MANIFEST.json"), "w"), indent=1)
    if verbose:
        print("staged %d reference files under %s" % (len(manifest), DST))
    return manifest


def load():
    """(gl_ProxGD_primal, gl_FProxGD_primal) of the staged reference, or None when oracle/_ref is absent."""
    if not os.path.exists(os.path.join(DST, "gl_FProxGD_primal.py")):
        return None
    import importlib.util
    if DST not in sys.path:
        sys.path.insert(0, DST)      # the solvers do `from utils.stopwatch import Stopwatch`
    mods = []
    for name in ("gl_ProxGD_primal", "gl_FProxGD_primal"):
        spec = importlib.util.spec_from_file_location("glprox_ref_" + name, os.path.join(DST, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mods.append(getattr(mod, name))
    return tuple(mods)


if __name__ == "__main__":
    if stage() is None:
        print("reference not found under %s: nothing staged" % REF_ROOT)
