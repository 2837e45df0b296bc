"""Model file format: gsr_ply.save_ply / load_ply against the layout GaussianModel.save_ply /
load_ply define (scene/gaussian_model.py:263-364).  `plyfile` is not in this image, so the checks restate what the
reference hands to plyfile (attribute list, channel-major SH blocks, all 'f4') and what plyfile's binary writer emits
for it (header lines, little-endian interleaved records).  CPU only."""
import os
import struct

import numpy as np
import pytest
import torch

import gsr_ply


def _model(P, deg, seed=0):
    g = torch.Generator().manual_seed(seed)
    K = (deg + 1) ** 2
    return dict(xyz=torch.randn(P, 3, generator=g), features_dc=torch.randn(P, 1, 3, generator=g),
                features_rest=torch.randn(P, K - 1, 3, generator=g), opacity=torch.randn(P, 1, generator=g),
                scaling=torch.randn(P, 3, generator=g), rotation=torch.randn(P, 4, generator=g))


def _reference_attributes(m):
    """The (P, n_attributes) matrix the reference concatenates before handing it to plyfile (:279-293)."""
    xyz = m["xyz"].numpy()
    f_dc = m["features_dc"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    f_rest = m["features_rest"].transpose(1, 2).flatten(start_dim=1).contiguous().numpy()
    return np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, m["opacity"].numpy(), m["scaling"].numpy(),
                           m["rotation"].numpy()), axis=1)


@pytest.mark.parametrize("deg", [0, 1, 3])
def test_file_bytes_are_what_the_reference_writes(tmp_path, deg):
    P = 37
    m = _model(P, deg)
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")     # directories are created (mkdir_p, :278)
    gsr_ply.save_ply(path, **m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode("ascii").splitlines()
    K = (deg + 1) ** 2
    names = (["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + ["f_rest_%d" % i for i in range(3 * (K - 1))]
             + ["opacity"] + ["scale_%d" % i for i in range(3)] + ["rot_%d" % i for i in range(4)])
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex %d" % P]
    assert lines[3:] == ["property float %s" % n for n in names]
    attrs = _reference_attributes(m)
    assert len(body) == 4 * attrs.size
    assert body == attrs.astype("<f4").tobytes()
    # spot check of the channel-major rule: f_rest_j = rest[:, j % (K-1), j // (K-1)]
    if K > 1:
        j = min(2 * (K - 1) + 1, 3 * (K - 1) - 1)
        col = 9 + j
        first = struct.unpack_from("<f", body, 4 * col)[0]
        assert first == float(m["features_rest"][0, j % (K - 1), j // (K - 1)])


@pytest.mark.parametrize("deg", [0, 2, 3])
def test_round_trip_is_bit_exact_and_shaped_like_the_model(tmp_path, deg):
    m = _model(101, deg, seed=3)
    path = str(tmp_path / "pc.ply")
    gsr_ply.save_ply(path, **m)
    back = gsr_ply.load_ply(path, max_sh_degree=deg)
    K = (deg + 1) ** 2
    assert back["features_dc"].shape == (101, 1, 3) and back["features_rest"].shape == (101, K - 1, 3)
    for k, v in m.items():
        assert back[k].dtype == np.float32 and back[k].flags["C_CONTIGUOUS"]
        assert np.array_equal(back[k], v.numpy()), k
    t = gsr_ply.gaussians_from_ply(path, max_sh_degree=deg, device="cpu")
    assert all(torch.equal(t[k], m[k]) for k in m)


def test_wrong_sh_degree_asserts_like_the_reference(tmp_path):
    path = str(tmp_path / "pc.ply")
    gsr_ply.save_ply(path, **_model(5, 1))
    with pytest.raises(AssertionError):           # scene/gaussian_model.py:330
        gsr_ply.load_ply(path, max_sh_degree=3)


def test_empty_model_and_bad_inputs(tmp_path):
    path = str(tmp_path / "empty.ply")
    gsr_ply.save_ply(path, **_model(0, 3))
    back = gsr_ply.load_ply(path, 3)
    assert back["xyz"].shape == (0, 3) and back["features_rest"].shape == (0, 15, 3)
    with pytest.raises(ValueError):
        gsr_ply.save_ply(path, torch.zeros(4, 2), torch.zeros(4, 1, 3), torch.zeros(4, 0, 3), torch.zeros(4, 1), torch.zeros(4, 3),
                         torch.zeros(4, 4))
    bad = tmp_path / "bad.ply"
    bad.write_bytes(b"plx\n")
    with pytest.raises(ValueError):
        gsr_ply.read_ply_vertices(str(bad))
    m = _model(9, 0)
    gsr_ply.save_ply(path, **m)
    raw = open(path, "rb").read()
    open(path, "wb").write(raw[:-5])
    with pytest.raises(ValueError, match="truncated"):
        gsr_ply.load_ply(path, 0)


def test_reader_accepts_ascii_big_endian_doubles_and_shuffled_property_order(tmp_path):
    """plyfile reads any scalar PLY; files converted by other tools (ASCII, doubles, reordered columns) must load too."""
    m = _model(6, 1, seed=5)
    attrs = _reference_attributes(m).astype(np.float64)
    names = gsr_ply.construct_list_of_attributes(3, 9)
    perm = np.random.RandomState(0).permutation(len(names))
    # ASCII, shuffled columns, a comment line, a trailing face element that must be ignored
    p1 = tmp_path / "ascii.ply"
    with open(p1, "w") as f:
        f.write("ply\nformat ascii 1.0\ncomment made by a test\nelement vertex 6\n")
        f.write("".join("property double %s\n" % names[j] for j in perm))
        f.write("element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for row in attrs:
            f.write(" ".join(repr(float(row[j])) for j in perm) + "\n")
    back = gsr_ply.load_ply(str(p1), 1)
    for k, v in m.items():
        assert np.array_equal(back[k], v.numpy()), k
    # big-endian binary
    p2 = tmp_path / "be.ply"
    with open(p2, "wb") as f:
        f.write(("ply\nformat binary_big_endian 1.0\nelement vertex 6\n" + "".join("property float32 %s\n" % n for n in names)
                 + "end_header\n").encode())
        f.write(attrs.astype(">f4").tobytes())
    back = gsr_ply.load_ply(str(p2), 1)
    for k, v in m.items():
        assert np.array_equal(back[k], v.numpy()), k
    p3 = tmp_path / "list.ply"
    p3.write_bytes(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty list uchar int idx\nend_header\n1 0\n")
    with pytest.raises(ValueError, match="list"):
        gsr_ply.read_ply_vertices(str(p3))
