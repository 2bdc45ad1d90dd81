"""CPU: the on-disk formats of SURVEY 8f row f4 (instag_b200/assets.py) against the reference's
format description (scene/gaussian_model.py:115-132, 429-527)."""
import os

import numpy as np
import pytest

import util as U  # noqa: F401
from instag_b200 import assets as AS


def _model(P=37, degree=2, seed=0):
    rng = np.random.default_rng(seed)
    K = (degree + 1) ** 2 - 1
    return dict(xyz=rng.standard_normal((P, 3)).astype(np.float32),
                features_dc=rng.standard_normal((P, 1, 3)).astype(np.float32),
                features_rest=rng.standard_normal((P, K, 3)).astype(np.float32),
                opacity=rng.standard_normal((P, 1)).astype(np.float32),
                scaling=rng.standard_normal((P, 3)).astype(np.float32),
                rotation=rng.standard_normal((P, 4)).astype(np.float32))


def test_attribute_order_is_the_references():
    names = AS.attribute_names(3, 24)
    assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert names[9] == "f_rest_0" and names[32] == "f_rest_23" and names[33] == "opacity"
    assert names[34:] == ["scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]


@pytest.mark.parametrize("degree", [0, 1, 3])
def test_ply_round_trip_and_layout(tmp_path, degree):
    m = _model(degree=degree)
    path = os.path.join(tmp_path, "pc.ply")
    AS.write_gaussian_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 37"]
    assert [l.split()[2] for l in lines[3:]] == AS.attribute_names(3, 3 * ((degree + 1) ** 2 - 1))
    assert all(l.split()[:2] == ["property", "float"] for l in lines[3:])
    row = np.frombuffer(body, "<f4").reshape(37, -1)
    assert np.array_equal(row[:, 0:3], m["xyz"]) and not row[:, 3:6].any()
    # f_dc / f_rest are the [P,K,3] tensors transposed to [P,3,K] and flattened (gaussian_model.py:448-449)
    assert np.array_equal(row[:, 6:9], m["features_dc"].transpose(0, 2, 1).reshape(37, -1))
    K3 = 3 * ((degree + 1) ** 2 - 1)
    assert np.array_equal(row[:, 9:9 + K3], m["features_rest"].transpose(0, 2, 1).reshape(37, -1))
    back = AS.read_gaussian_ply(path, max_sh_degree=degree)
    for k in m:
        assert np.array_equal(back[k], m[k]), k
    assert back["sh_degree"] == degree
    if degree != 1:
        with pytest.raises(AssertionError):
            AS.read_gaussian_ply(path, max_sh_degree=1)


def test_ascii_and_big_endian_ply_are_read(tmp_path):
    m = _model(P=5, degree=0)
    names = AS.attribute_names(3, 0)
    table = np.concatenate([m["xyz"], np.zeros((5, 3), np.float32), m["features_dc"].reshape(5, 3), m["opacity"],
                            m["scaling"], m["rotation"]], axis=1)
    hdr = "ply\nformat {}\ncomment made by a test\nelement vertex 5\n" + "".join(f"property float {n}\n" for n in names) + \
          "element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    pa = os.path.join(tmp_path, "a.ply")
    with open(pa, "w") as f:
        f.write(hdr.format("ascii 1.0"))
        for r in table:
            f.write(" ".join(repr(float(x)) for x in r) + "\n")
    pb = os.path.join(tmp_path, "b.ply")
    with open(pb, "wb") as f:
        f.write(hdr.format("binary_big_endian 1.0").encode())
        f.write(table.astype(">f4").tobytes())
    for p in (pa, pb):
        back = AS.read_gaussian_ply(p)
        assert np.allclose(back["xyz"], m["xyz"]) and np.allclose(back["rotation"], m["rotation"])
        assert back["features_rest"].shape == (5, 0, 3) and back["sh_degree"] == 0


def test_capture_tuple_fields_and_fuse_checkpoint(tmp_path):
    import torch
    m = _model(P=4, degree=1)
    tup = (1, torch.from_numpy(m["xyz"]), torch.from_numpy(m["features_dc"]), torch.from_numpy(m["features_rest"]),
           torch.zeros(4, 1), torch.from_numpy(m["scaling"]), torch.from_numpy(m["rotation"]),
           torch.from_numpy(m["opacity"]), torch.zeros(4), torch.zeros(4, 1), torch.zeros(4, 1), {}, 1.0, {}, {})
    d = AS.unpack_capture(tup)
    assert d["active_sh_degree"] == 1 and d["rotation"].shape == (4, 4) and d["denom"].shape == (4, 1)
    with pytest.raises(ValueError):
        AS.unpack_capture(tup[:-1])
    path = os.path.join(tmp_path, "chkpnt_fuse_latest.pth")
    torch.save((tup, {"w": torch.zeros(1)}, tup, {}), path)
    face, mouth = AS.load_fuse_checkpoint(path)
    assert torch.equal(face["xyz"], tup[1]) and mouth["active_sh_degree"] == 1
