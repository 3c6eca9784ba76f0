"""PLY I/O of Gaussian scenes (gsr_ply) against the reference's on-disk layout (scene/gaussian_model.py:377-467)."""
import numpy as np
import pytest

import gsr_ply


def _scene(P=37, M=16, K=16, seed=3):
    r = np.random.default_rng(seed)
    f = lambda *s: r.standard_normal(s).astype(np.float32)
    return dict(xyz=f(P, 3), features_dc=f(P, 1, 3), features_rest=f(P, M - 1, 3), opacity=f(P, 1), scaling=f(P, 3),
                rotation=f(P, 4), objects_dc=f(P, 1, K))


def test_round_trip_is_bit_exact(tmp_path):
    s = _scene()
    p = tmp_path / "point_cloud.ply"
    gsr_ply.save_ply(str(p), s)
    t = gsr_ply.load_ply(str(p))
    for k in s:
        assert t[k].dtype == np.float32 and t[k].shape == s[k].shape, k
        assert np.array_equal(t[k], s[k]), k


def test_header_and_channel_major_layout(tmp_path):
    s = _scene(P=5)
    p = tmp_path / "a.ply"
    gsr_ply.save_ply(str(p), s)
    raw = open(p, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().split("\n")
    assert lines[0] == "ply" and lines[1] == "format binary_little_endian 1.0" and lines[2] == "element vertex 5"
    names = [l.split()[2] for l in lines if l.startswith("property")]
    assert names == gsr_ply.property_names(45, 16)
    assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] and names[54] == "opacity"
    assert all(l.split()[1] == "float" for l in lines if l.startswith("property"))
    tab = np.frombuffer(body, "<f4").reshape(5, len(names))
    assert np.array_equal(tab[:, 3:6], np.zeros((5, 3), np.float32))            # normals are written as zeros
    # f_rest_{c*15+k} = coefficient k+1 of channel c (transpose(1,2).flatten)
    for c in range(3):
        for k in (0, 7, 14):
            assert np.array_equal(tab[:, names.index(f"f_rest_{c * 15 + k}")], s["features_rest"][:, k, c])
    assert np.array_equal(tab[:, names.index("obj_dc_3")], s["objects_dc"][:, 0, 3])
    assert np.array_equal(tab[:, names.index("rot_2")], s["rotation"][:, 2])


def test_missing_object_features_and_shuffled_property_order(tmp_path):
    s = _scene(P=4)
    names = gsr_ply.property_names(45, 0)
    order = list(reversed(names))  # the reader looks properties up by name
    cols = {"x": s["xyz"][:, 0], "y": s["xyz"][:, 1], "z": s["xyz"][:, 2], "opacity": s["opacity"][:, 0]}
    for n in ("nx", "ny", "nz"):
        cols[n] = np.zeros(4, np.float32)
    for c in range(3):
        cols[f"f_dc_{c}"] = s["features_dc"][:, 0, c]
        cols[f"scale_{c}"] = s["scaling"][:, c]
        for k in range(15):
            cols[f"f_rest_{c * 15 + k}"] = s["features_rest"][:, k, c]
    for c in range(4):
        cols[f"rot_{c}"] = s["rotation"][:, c]
    p = tmp_path / "b.ply"
    with open(p, "wb") as f:
        f.write(("ply\nformat binary_little_endian 1.0\ncomment made by a test\nelement vertex 4\n" +
                 "".join(f"property float {n}\n" for n in order) + "end_header\n").encode())
        f.write(np.stack([cols[n] for n in order], axis=1).astype("<f4").tobytes())
    t = gsr_ply.load_ply(str(p))
    for k in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation"):
        assert np.array_equal(t[k], s[k]), k
    assert t["objects_dc"].shape == (4, 1, 16) and not t["objects_dc"].any()


def test_wrong_sh_count_asserts_and_activations(tmp_path):
    s = _scene(P=3, M=4)
    p = tmp_path / "c.ply"
    gsr_ply.save_ply(str(p), s)
    with pytest.raises(AssertionError):
        gsr_ply.load_ply(str(p), max_sh_degree=3)
    t = gsr_ply.load_ply(str(p), max_sh_degree=1)
    a = gsr_ply.activated(t)
    assert a["shs"].shape == (3, 4, 3) and np.array_equal(a["shs"][:, :1], s["features_dc"])
    assert np.allclose(np.linalg.norm(a["rotations"], axis=1), 1.0, atol=1e-6)
    assert np.allclose(a["opacities"], 1 / (1 + np.exp(-s["opacity"])), atol=1e-6)
    assert np.allclose(a["scales"], np.exp(s["scaling"]), rtol=1e-6)


def test_combine_splats_concatenates_pads_and_masks(tmp_path):
    """combine_splats (scene/gaussian_model.py:469-556): order kept, f_rest padded, objects zero, masks per file."""
    a, b = _scene(P=11, seed=1), _scene(P=7, M=4, seed=2)  # b is a degree-1 splat (9 f_rest columns)
    pa, pb = str(tmp_path / "a.ply"), str(tmp_path / "b.ply")
    gsr_ply.save_ply(pa, a)
    gsr_ply.save_ply(pb, b)
    s, masks = gsr_ply.combine_splats([pa, pb])
    assert s["xyz"].shape == (18, 3) and s["features_rest"].shape == (18, 15, 3) and s["objects_dc"].shape == (18, 1, 16)
    for k in ("xyz", "features_dc", "opacity", "scaling", "rotation"):
        assert np.array_equal(s[k][:11], a[k]) and np.array_equal(s[k][11:], b[k]), k
    assert np.array_equal(s["features_rest"][:11], a["features_rest"])
    # the 9 stored columns are channel-major with 3 per channel; padded to 45 and re-read as (3, 15): the
    # reference's own reshape, so the first 9 flat values are b's, the rest zero
    flat = s["features_rest"][11:].transpose(0, 2, 1).reshape(7, 45)
    assert np.array_equal(flat[:, :9], b["features_rest"].transpose(0, 2, 1).reshape(7, 9)) and not flat[:, 9:].any()
    assert not s["objects_dc"].any()
    assert [m.shape for m in masks] == [(18,), (18,)]
    assert masks[0][:11].all() and not masks[0][11:].any() and masks[1][11:].all() and not masks[1][:11].any()
    _, ref = gsr_ply.combine_splats([pa, pb], reference_masks=True)
    assert ref[0].shape == (11,) and ref[0].all() and ref[1].shape == (7,) and not ref[1].any()
    with pytest.raises(ValueError):
        gsr_ply.combine_splats([])
