"""pip-installable build of the drop-in (the reference installs its rasterizer with
`pip install submodules/diff-gaussian-rasterization`, DGR setup.py:12-35):

    pip install --no-build-isolation .        # nvcc for sm_100a + g++; needs torch in the environment

installs the packages `diff_gauss` (the import name gaussian_renderer/__init__.py:15 uses) and
`instag_b200` with libigs_raster.so and _igs_torch.so built in place by instag_b200/build.py.
"""
import os
import sys

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildWithLibrary(build_py):
    def run(self):
        sys.path.insert(0, ROOT)
        from instag_b200 import build as B
        B.build()           # libigs_raster.so (nvcc -gencode arch=compute_100a,code=sm_100a)
        B.build_binding()   # _igs_torch.so (g++, links the library by $ORIGIN)
        super().run()


setup(
    name="diff_gauss_b200",
    version="0.2.0",
    description="B200-native drop-in for InsTaG's diff_gauss Gaussian-splatting rasterizer (libigs_raster)",
    packages=["diff_gauss", "instag_b200"],
    package_data={"instag_b200": ["libigs_raster.so", "_igs_torch.so", "csrc/*", "../include/igs_raster.h"]},
    include_package_data=True,
    cmdclass={"build_py": BuildWithLibrary},
    python_requires=">=3.9",
)
