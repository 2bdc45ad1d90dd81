"""Loader of the reference's own extension (baseline/_ref/diff_gauss, built by baseline/build_ref.sh
from the unmodified sources) under the module name `diff_gauss_ref`, so that it can live in one
process with this repo's `diff_gauss`.

    ref = load()                      # the reference's diff_gauss/__init__.py, its _C beside it
    ref.GaussianRasterizer, ref.GaussianRasterizationSettings, ref._C.rasterize_gaussians, ...

The reference's Python (diff_gauss/__init__.py:21-250), its torch binding (rasterize_points.cu:35-239,
ext.cpp:15-19) and its kernels all run as they are; nothing of this repo is on that path."""
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(HERE, "_ref", "diff_gauss")
_mod = None


def available():
    if not os.path.isfile(os.path.join(PKG, "__init__.py")):
        return False
    return any(f.startswith("_C.") and f.endswith(".so") for f in os.listdir(PKG))


def load():
    """The reference package as module `diff_gauss_ref`; raises if baseline/_ref has not been built."""
    global _mod
    if _mod is None:
        if not available():
            raise RuntimeError("baseline/_ref/diff_gauss is missing: run baseline/build_ref.sh where /root/reference exists")
        import torch  # noqa: F401  (the extension links against libtorch, which `import torch` loads)
        spec = importlib.util.spec_from_file_location("diff_gauss_ref", os.path.join(PKG, "__init__.py"),
                                                      submodule_search_locations=[PKG])
        mod = importlib.util.module_from_spec(spec)
        sys.modules["diff_gauss_ref"] = mod
        spec.loader.exec_module(mod)
        _mod = mod
    return _mod
