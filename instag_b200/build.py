"""In-tree build of libigs_raster.so (nvcc, sm_100a only).

    python -m instag_b200.build [--force]

The shared library is written next to this file so that it travels to the GPU
box with the repo snapshot; it is git-ignored.  No JIT cache is involved.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libigs_raster.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
SOURCES = ["igs_api.cu", "igs_preprocess.cu", "igs_binning.cu", "igs_render.cu",
           "igs_render_backward.cu", "igs_preprocess_backward.cu", "igs_adjacent.cu", "igs_loss.cu", "igs_geom_loss.cu", "igs_collective.cu"]
# IGS_NVCC_DEFS="-DX=1 -DY=2": extra macro definitions for kernel experiments (profiles/ notes say which)
FLAGS = os.environ.get("IGS_NVCC_DEFS", "").split() + ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _deps():
    out = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    out.append(os.path.join(HERE, "..", "include", "igs_raster.h"))
    return out


STAMP = os.path.join(HERE, "libigs_raster.stamp")


def _digest():
    """Content hash of every source (mtimes do not survive the copy to the GPU box)."""
    import hashlib
    h = hashlib.sha256(" ".join(FLAGS).encode())
    for p in sorted(_deps()):
        if p.endswith((".cu", ".cuh", ".h")):
            h.update(open(p, "rb").read())
    return h.hexdigest()


def needs_build():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    return open(STAMP).read().strip() != _digest()


def _compile(src):
    obj = os.path.join(CSRC, src[:-3] + ".o")
    cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return src, obj, r


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a and link libigs_raster.so."""
    if not force and not needs_build():
        return LIB
    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        results = list(ex.map(_compile, SOURCES))
    log = []
    for src, obj, r in results:
        log.append(f"== {src}\n{r.stderr}")
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{r.stdout}\n{r.stderr}")
    objs = [obj for _, obj, _ in results]
    r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs],
                       capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    with open(os.path.join(CSRC, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    with open(STAMP, "w") as f:
        f.write(_digest())
    if verbose:
        print("\n".join(log))
    return LIB


# ---- the compiled host side (csrc/igs_torch.cpp): a torch extension module linked against the library ----
BINDING = os.path.join(HERE, "_igs_torch.so")
BINDING_SRC = os.path.join(CSRC, "igs_torch.cpp")
BINDING_STAMP = os.path.join(HERE, "_igs_torch.stamp")
CXX = os.environ.get("CXX", "g++")


def _binding_digest():
    import hashlib
    import torch
    h = hashlib.sha256(torch.__version__.encode())
    for p in (BINDING_SRC, os.path.join(HERE, "..", "include", "igs_raster.h")):
        h.update(open(p, "rb").read())
    return h.hexdigest()


def build_binding(force=False):
    """g++ build of _igs_torch.so (no nvcc: the file launches no kernel).  Links libigs_raster.so by
    rpath $ORIGIN and torch's libraries, which `import torch` has already loaded."""
    if not force and os.path.exists(BINDING) and os.path.exists(BINDING_STAMP) \
            and open(BINDING_STAMP).read().strip() == _binding_digest():
        return BINDING
    import sysconfig
    import torch
    tdir = os.path.dirname(torch.__file__)
    inc = [os.path.join(tdir, "include"), os.path.join(tdir, "include", "torch", "csrc", "api", "include"),
           os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include"),
           sysconfig.get_paths()["include"], os.path.join(HERE, "..", "include")]
    cmd = [CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-fvisibility=hidden", "-Wno-attributes",
           f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}",
           "-DTORCH_EXTENSION_NAME=_igs_torch", "-DTORCH_API_INCLUDE_EXTENSION_H",
           *[f"-I{i}" for i in inc], BINDING_SRC, "-o", BINDING,
           f"-L{os.path.join(tdir, 'lib')}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
           "-ltorch_python", f"-L{HERE}", "-l:libigs_raster.so", "-Wl,-rpath,$ORIGIN"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"g++ failed on igs_torch.cpp:\n{r.stdout}\n{r.stderr}")
    with open(BINDING_STAMP, "w") as f:
        f.write(_binding_digest())
    return BINDING


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_binding(force="--force" in sys.argv))
