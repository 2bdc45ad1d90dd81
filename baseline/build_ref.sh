#!/bin/bash
# Build the reference's own torch extension (diff_gauss: rasterize_points.cu + ext.cpp +
# cuda_rasterizer/*.cu + diff_gauss/__init__.py), UNMODIFIED, for sm_100a, into baseline/_ref/.
#   * /root/reference is read-only and setup.py builds in-tree: the sources are copied to /tmp first;
#   * the only change to the build is on the command line: cuda_rasterizer/rasterizer_impl.h uses
#     uint32_t / std::uintptr_t without <cstdint> (fails under gcc 13), so `-include cstdint` is
#     force-included for nvcc (NVCC_APPEND_FLAGS) and the host compiler (CXXFLAGS); the arch list of
#     torch.utils.cpp_extension is set to 10.0a (the reference's CMake file lists sm_70/75/86 only).
# ~6 minutes on 8 cores.  baseline/_ref is git-ignored and travels to the GPU box with the snapshot.
set -euo pipefail
SRC=${1:-/root/reference/submodules/diff-gaussian-rasterization}
HERE="$(cd "$(dirname "$0")" && pwd)"
WORK=$(mktemp -d /tmp/dgr_ref.XXXXXX)
cp -r "$SRC"/. "$WORK"/
mkdir -p "$HERE/_ref"
cd "$WORK"
TORCH_CUDA_ARCH_LIST="10.0a" NVCC_APPEND_FLAGS="-include cstdint" CXXFLAGS="-include cstdint" CFLAGS="-include cstdint" \
  MAX_JOBS=${MAX_JOBS:-8} python -m pip install --no-index --no-build-isolation --no-deps \
  --find-links /opt/wheelhouse --upgrade --target "$HERE/_ref" "$WORK"
rm -rf "$WORK"
# the caller of the path, as it is: gaussian_renderer/__init__.py (render, render_motion, ...), for the test that
# runs it unmodified on top of either rasterizer (tests/test_gpu_reference_module.py stubs its other imports)
REPO_ROOT="$(cd "$SRC/../.." && pwd)"
if [ -f "$REPO_ROOT/gaussian_renderer/__init__.py" ]; then
  mkdir -p "$HERE/_ref/instag_caller/gaussian_renderer"
  cp "$REPO_ROOT/gaussian_renderer/__init__.py" "$HERE/_ref/instag_caller/gaussian_renderer/__init__.py"
fi
ls -la "$HERE/_ref/diff_gauss"
