"""Generate the golden vectors of tests/golden/ FROM THE REFERENCE ITSELF.

Run on a B200 box (needs oracle/_ref/libdgr_ref.so = the unmodified reference
rasterizer compiled by oracle/Makefile):

    python tests/golden/make_golden.py [outdir]

For every case in CASES the seeded synthetic scene is rasterized forward and
backward by the reference; inputs are NOT stored (they are regenerated from the
seed by instag_b200.synthetic, which is numpy-only and deterministic), outputs
and all intermediates are.  tests/test_oracle_golden.py checks the CPU oracle
against these files; tests/test_gpu_parity.py checks the product against them.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import util as U  # noqa: E402
from instag_b200 import synthetic as S  # noqa: E402

# name -> make_scene kwargs (small: <= 2k Gaussians, <= 80x64 pixels)
CASES = {
    "precomp_64": dict(P=1500, W=64, H=64, sh_degree=None, ED=1, scale_mult=2.0, seed=1),
    "sh1_m4_64": dict(P=1500, W=64, H=64, sh_degree=1, M=4, ED=1, scale_mult=2.0, seed=2),
    "sh3_m16_ragged": dict(P=1200, W=75, H=50, sh_degree=3, M=16, ED=1, scale_mult=2.5, seed=3),
    "sh0_m9_noextra": dict(P=1000, W=48, H=64, sh_degree=0, M=9, ED=0, scale_mult=2.0, seed=4),
    "big_splats_ed3": dict(P=300, W=64, H=64, sh_degree=2, M=9, ED=3, scale_mult=12.0, seed=5),
}

INT_KEYS = ("radii", "tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib")
FLOAT_KEYS = ("color", "depth", "norm", "alpha", "extra", "means2D", "depths", "conic_opacity", "rgb", "norm3D",
              "cov3D", "clamped")


def main(outdir):
    os.makedirs(outdir, exist_ok=True)
    for name, kw in CASES.items():
        sc = U.make_scene(**kw)
        g = S.upstream_grads(sc["H"], sc["W"], extra_dims=kw.get("ED", 1), seed=11)
        r = U.run_ref(sc, grads=g)
        r2 = U.run_ref(sc, grads=g)  # second run: the reference's own atomic-order spread
        spread = {k: U.rel_err(r[k], r2[k]) for k in U.GRAD_NAMES}
        vis = r["radii"] > 0
        save = {"num_rendered": np.int64(r["num_rendered"])}
        for k in INT_KEYS + FLOAT_KEYS + U.GRAD_NAMES:
            v = np.asarray(r[k])
            # per-Gaussian forward state of culled Gaussians is uninitialised in the reference: zero it
            if k in ("means2D", "depths", "conic_opacity", "rgb", "norm3D", "cov3D", "clamped"):
                v = v.copy()
                v[~vis] = 0
                if k in ("rgb", "clamped") and sc["shs"] is None:
                    v[:] = 0  # not written by the reference with precomputed colours
            save[k] = v
        save["spread"] = np.array([spread[k] for k in U.GRAD_NAMES], np.float64)
        np.savez_compressed(os.path.join(outdir, name + ".npz"), **save)
        print(name, "num_rendered", r["num_rendered"], "visible", int(vis.sum()), "max spread",
              max(spread.values()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
