"""instag_b200: B200-native (sm_100a) differentiable Gaussian-splatting rasterizer
for InsTaG -- the hot path only.

    csrc/            CUDA kernels + the C ABI (include/igs_raster.h)
    _native.py       ctypes binding of libigs_raster.so
    rasterizer.py    the reference's Python API (GaussianRasterizer, ...)
    synthetic.py     seeded clouds / cameras of the BASELINE.json shapes
    build.py         in-tree nvcc build
"""
