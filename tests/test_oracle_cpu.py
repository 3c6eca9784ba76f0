"""Pins the CPU oracle (oracle/cpu_ref.cpp) against outputs of the reference CUDA rasterizer itself
(tests/golden/*.npz, produced on a B200 by tools/make_golden.py from the unmodified
DGR/cuda_rasterizer sources) and against independent restatements of the reference's Python
helpers (utils/sh_utils.py eval_sh, utils/general_utils.py build_scaling_rotation).

Tolerances: integer/index outputs (radii, tiles_touched, point_offsets, sort keys, point_list, tile
ranges, n_contrib) bit-exact; image <= 1e-5 abs (north_star); gradients |d| <= 1e-3*|ref| + 1e-4*max|ref|.
"""
import numpy as np
import pytest

import helpers as h

CASES = ["tiny_sh3", "tiny_sh1_bg", "tiny_precomp"]
INT_KEYS = ("radii", "tiles_touched", "point_offsets", "keys_sorted", "point_list", "ranges", "n_contrib")


@pytest.fixture(scope="module", params=CASES)
def case(request):
    scene, kw, gold = h.golden_case(request.param)
    o = h.run_oracle(scene, **kw)
    return scene, kw, gold, o


def test_oracle_integer_outputs_bit_exact_vs_reference_cuda(case):
    scene, kw, gold, o = case
    assert o["num_rendered"] == gold["num_rendered"]
    for k in INT_KEYS:
        a, b = np.asarray(o[k]).reshape(-1), np.asarray(gold[k]).reshape(-1)
        assert a.shape == b.shape, k
        assert np.array_equal(a, b), f"{k}: {(a != b).sum()} mismatches"


def test_oracle_geometry_close_to_reference_cuda(case):
    scene, kw, gold, o = case
    P = scene["P"]
    vis = gold["radii"] > 0
    assert np.array_equal(o["depths"][vis].view(np.uint32), gold["depths"][vis].view(np.uint32))  # depth bits feed the keys
    for k, w, tol in (("means2D", 2, 0.0), ("conic_opacity", 4, 2e-6), ("rgb", 3, 2e-7)):
        if k == "rgb" and "colors_precomp" in kw:
            continue  # the reference reads colors_precomp directly; its geometry-state rgb is not written
        a, b = o[k].reshape(P, w)[vis], gold[k].reshape(P, w)[vis]
        err = np.abs(a.astype(np.float64) - b)
        assert (err <= tol * np.maximum(1.0, np.abs(b))).all(), (k, err.max())
    if "cov3D_precomp" not in kw:
        a, b = o["cov3D"].reshape(P, 6)[vis], gold["cov3D"].reshape(P, 6)[vis]
        assert np.abs(a - b).max() <= 1e-6 * np.abs(b).max()
        # clamped flags: bool[3P] (forward.cu:66-69)
        gc = gold["clamped"].reshape(P, 3).astype(bool)
        oc = o["clamped"].reshape(P, 3).astype(bool)
        if "colors_precomp" not in kw:
            assert np.array_equal(oc[vis], gc[vis])


def test_oracle_image_within_1e5_of_reference_cuda(case):
    scene, kw, gold, o = case
    assert np.abs(o["color"] - gold["color"]).max() <= 1e-5
    assert np.abs(o["final_T"] - gold["final_T"]).max() <= 1e-5


def test_oracle_gradients_within_tolerance_of_reference_cuda(case):
    scene, kw, gold, o = case
    seen = 0
    for k, ref in gold["grads"].items():
        if k not in o["grads"]:
            continue
        if k == "sh" and "colors_precomp" in kw:
            continue
        if k in ("scales", "rotations") and "cov3D_precomp" in kw:
            continue
        a = o["grads"][k].reshape(-1)
        b = ref.reshape(-1)[: a.size] if k != "conic" else ref.reshape(-1)
        ok, worst = h.grad_close(a, b)
        assert ok, f"grad {k}: tolerance ratio {worst}"
        seen += 1
    assert seen >= 5


# ------------------------------------------------------------ independent cross-checks
def _eval_sh_deg3(sh, dirs):
    """Restatement of eval_sh (utils/sh_utils.py:57-112) in float64; sh (P,16,3), dirs (P,3) unit."""
    C0 = 0.28209479177387814
    C1 = 0.4886025119029199
    C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
    C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435]
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    r = C0 * sh[:, 0]
    r = r - C1 * y * sh[:, 1] + C1 * z * sh[:, 2] - C1 * x * sh[:, 3]
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    r = (r + C2[0] * xy * sh[:, 4] + C2[1] * yz * sh[:, 5] + C2[2] * (2 * zz - xx - yy) * sh[:, 6]
         + C2[3] * xz * sh[:, 7] + C2[4] * (xx - yy) * sh[:, 8])
    r = (r + C3[0] * y * (3 * xx - yy) * sh[:, 9] + C3[1] * xy * z * sh[:, 10] + C3[2] * y * (4 * zz - xx - yy) * sh[:, 11]
         + C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12] + C3[4] * x * (4 * zz - xx - yy) * sh[:, 13]
         + C3[5] * z * (xx - yy) * sh[:, 14] + C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return r


def test_oracle_sh_colour_matches_python_eval_sh():
    """gaussian_renderer/__init__.py:73-78: colour = clamp_min(eval_sh(...) + 0.5, 0)."""
    s = h.make_scene("tiny")
    o = h.run_oracle(s, grad=False)
    vis = o["radii"] > 0
    d = s["means3D"].astype(np.float64) - s["cam"]["campos"].astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    want = np.maximum(_eval_sh_deg3(s["shs"].astype(np.float64), d) + 0.5, 0.0)
    got = o["rgb"].reshape(-1, 3)
    assert np.abs(got[vis] - want[vis]).max() < 5e-6


def test_oracle_cov3d_matches_python_build_scaling_rotation():
    """utils/general_utils.py:78-110 + scene/gaussian_model.py:25-29 (unit quaternions)."""
    s = h.make_scene("tiny")
    o = h.run_oracle(s, grad=False)
    vis = o["radii"] > 0
    q = s["rotations"].astype(np.float64)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(-1, 3, 3)
    Lm = R * s["scales"].astype(np.float64)[:, None, :]
    S = Lm @ Lm.transpose(0, 2, 1)
    want = np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1)
    got = o["cov3D"].reshape(-1, 6)
    assert np.abs(got[vis] - want[vis]).max() <= 1e-6 * np.abs(want).max()


def test_oracle_sort_and_ranges_invariants():
    s = h.make_scene("small")
    o = h.run_oracle(s, grad=False)
    keys = o["keys_sorted"]
    assert (np.diff(keys.astype(np.uint64)) >= 0).all() if keys.size > 1 else True
    tile = (keys >> np.uint64(32)).astype(np.int64)
    ranges = o["ranges"].reshape(-1, 2)
    counts = np.bincount(tile, minlength=ranges.shape[0])
    assert np.array_equal(ranges[:, 1].astype(np.int64) - ranges[:, 0], counts)
    assert int(o["tiles_touched"].sum()) == o["num_rendered"] == keys.size
    assert np.array_equal(np.cumsum(o["tiles_touched"], dtype=np.uint32), o["point_offsets"])
    # stable: equal keys keep ascending Gaussian index (emission order)
    pl = o["point_list"].astype(np.int64)
    same = keys[1:] == keys[:-1]
    assert (pl[1:][same] > pl[:-1][same]).all()


def test_oracle_gradient_matches_finite_differences():
    """Directional finite difference of L = <image, dL_dpix> w.r.t. opacity and SH DC (smooth directions)."""
    s = h.make_scene("tiny")
    base = h.run_oracle(s, grad=True)
    rng = np.random.default_rng(0)

    def loss(sc):
        return float((h.run_oracle(sc, grad=False)["color"].astype(np.float64) * s["dL_dpix"]).sum())

    # SH enters linearly (up to the clamp at 0): tight.  Opacity moves pairs across the alpha < 1/255 and
    # T < 1e-4 cut-offs, where L jumps and the analytic gradient (like the reference's) ignores the jump: loose.
    for name, gkey, eps, rtol in (("opacities", "opacity", 1e-3, 0.25), ("shs", "sh", 1e-2, 0.02)):
        d = rng.standard_normal(s[name].shape).astype(np.float32)
        if name == "opacities":
            d *= np.minimum(s[name], 1 - s[name])  # stay inside (0,1)
        sp, sm = dict(s), dict(s)
        sp[name] = (s[name] + eps * d).astype(np.float32)
        sm[name] = (s[name] - eps * d).astype(np.float32)
        fd = (loss(sp) - loss(sm)) / (2 * eps)
        an = float((base["grads"][gkey].reshape(d.shape).astype(np.float64) * d).sum())
        assert abs(fd - an) <= rtol * max(abs(an), abs(fd)) + 1e-9, (name, fd, an)


def test_oracle_empty_and_all_culled():
    s = h.make_scene("tiny")
    behind = dict(s)
    behind["means3D"] = (s["means3D"] + np.array([0, 0, -10.0], np.float32)).astype(np.float32)  # behind the camera
    o = h.run_oracle(behind, grad=True)
    assert o["num_rendered"] == 0 and (o["radii"] == 0).all()
    assert np.abs(o["color"]).max() == 0.0  # bg is zero
    for g in o["grads"].values():
        assert np.abs(g).max() == 0.0
    from oracle import cpu_ref
    vis = cpu_ref.mark_visible(behind["means3D"], s["cam"]["viewmatrix"], s["cam"]["projmatrix"])
    assert not vis.any()
    vis = cpu_ref.mark_visible(s["means3D"], s["cam"]["viewmatrix"], s["cam"]["projmatrix"])
    assert vis.all()


def test_oracle_matches_reference_cuda_digest_at_C1():
    """BASELINE configs[0] size (100k Gaussians SH3, 800x800, forward + backward): the CPU oracle against a digest of
    the reference CUDA's own run on a B200 (tests/golden/c1_digest.npz, made by `tools/make_golden.py --digest`):
    SHA-256 of every integer / index output (bit-exact: radii, tiles_touched, point_offsets, the 593k sorted 64-bit
    keys, point_list, tile ranges, n_contrib), depth bits, image and final_T on a 4x4 pixel grid (<= 1e-5), every
    64th row of each gradient and each gradient's float64 sums (the gradient tolerance of the other tests)."""
    import hashlib
    import os
    z = np.load(os.path.join(h.GOLDEN_DIR, "c1_digest.npz"))
    scene = h.make_scene("C1")
    P, W, H, D = (int(x) for x in z["meta"])
    assert (scene["P"], scene["cam"]["image_width"], scene["cam"]["image_height"], scene["sh_degree"]) == (P, W, H, D)
    o = h.run_oracle(scene)
    assert o["num_rendered"] == int(z["num_rendered"])
    for k in INT_KEYS:
        a = np.ascontiguousarray(o[k])
        assert a.size == int(z["len_" + k]), k
        assert hashlib.sha256(a.tobytes()).digest() == z["sha_" + k].tobytes(), k
    assert np.array_equal(o["radii"][::64], z["radii_sub"])
    vis = z["radii_sub"] > 0
    assert np.array_equal(o["depths"].view(np.uint32)[::64][vis], z["depth_bits_sub"][vis])
    assert np.abs(o["color"].reshape(3, H, W)[:, ::4, ::4] - z["color_sub"]).max() <= 1e-5
    assert np.abs(o["final_T"].reshape(H, W)[::4, ::4] - z["final_T_sub"]).max() <= 1e-5
    for k in ("means3D", "means2D", "opacity", "sh", "scales", "rotations"):
        g = np.asarray(o["grads"][k]).reshape(P, -1)
        ref_sub, gmax = z["grad_sub_" + k], float(z["grad_max_" + k])
        sub = g[::64][:, : ref_sub.shape[1]]
        tol = 1e-3 * np.abs(ref_sub) + 1e-4 * gmax
        assert (np.abs(sub - ref_sub) <= tol).all(), (k, float((np.abs(sub - ref_sub) / tol).max()))
        s_abs = float(z["grad_abs_" + k])
        assert abs(float(g[:, : ref_sub.shape[1]].astype(np.float64).sum()) - float(z["grad_sum_" + k])) <= 1e-3 * s_abs, k
