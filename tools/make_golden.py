"""Generates tests/golden/*.npz on the GPU box by running the UNMODIFIED reference CUDA rasterizer
(oracle/_ref/libdgr_ref.so = DGR/cuda_rasterizer/*.cu compiled for sm_100a behind oracle/ref_shim.cu)
on small deterministic scenes.  The fixtures pin the CPU oracle (tests/test_oracle_cpu.py) and are a
second check of the product (tests/test_gpu_parity.py).

    gpurun -- 'python tools/make_golden.py gpurun_out/golden'   then   cp gpurun_out/golden/*.npz tests/golden/
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as h  # noqa: E402

KEEP = ("color", "radii", "depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped", "tiles_touched",
        "point_offsets", "final_T", "n_contrib", "ranges", "keys_sorted", "point_list")


def variants():
    """name -> (scene, kwargs for run_ref / run_oracle / run_gsr)."""
    out = {}
    out["tiny_sh3"] = (h.make_scene(**h.GOLDEN_SCENES["tiny_sh3"]), {})
    out["tiny_sh1_bg"] = (h.make_scene(**h.GOLDEN_SCENES["tiny_sh1_bg"]), dict(bg=[0.3, 0.5, 0.7]))
    s = h.make_scene(**h.GOLDEN_SCENES["tiny_precomp"])
    rng = np.random.default_rng(5)
    base = h.run_ref(s, grad=False)
    cov = base["cov3D"].reshape(s["P"], 6).copy()
    invis = base["radii"] <= 0  # culled Gaussians have no cov3D in the geometry state: give them a valid one
    cov[invis] = np.array([1e-4, 0, 0, 1e-4, 0, 1e-4], np.float32)
    out["tiny_precomp"] = (s, dict(colors_precomp=rng.uniform(0, 1, (s["P"], 3)).astype(np.float32), cov3D_precomp=cov,
                                   bg=[0.1, 0.2, 0.9]))
    return out


INT_KEYS = ("radii", "tiles_touched", "point_offsets", "keys_sorted", "point_list", "ranges", "n_contrib")
DIGEST_PIX_STRIDE, DIGEST_ROW_STRIDE = 4, 64


def digest(scene, r):
    """A compact record of a reference-CUDA run that is too large to commit whole (BASELINE configs[0]: 100k Gaussians,
    800x800): SHA-256 of every integer / index array, the image and final_T on every 4th pixel of every 4th row, every
    64th row of each gradient, and float64 sums of each gradient.  tests/test_oracle_cpu.py holds the CPU oracle to it."""
    import hashlib
    d = {"num_rendered": np.int64(r["num_rendered"]),
         "meta": np.array([scene["P"], scene["cam"]["image_width"], scene["cam"]["image_height"], scene["sh_degree"]], np.int64)}
    for k in INT_KEYS:
        a = np.ascontiguousarray(r[k])
        d["sha_" + k] = np.frombuffer(hashlib.sha256(a.tobytes()).digest(), np.uint8).copy()
        d["len_" + k] = np.int64(a.size)
    H, W = scene["cam"]["image_height"], scene["cam"]["image_width"]
    d["color_sub"] = r["color"].reshape(3, H, W)[:, ::DIGEST_PIX_STRIDE, ::DIGEST_PIX_STRIDE].copy()
    d["final_T_sub"] = r["final_T"].reshape(H, W)[::DIGEST_PIX_STRIDE, ::DIGEST_PIX_STRIDE].copy()
    d["depth_bits_sub"] = r["depths"].view(np.uint32)[::DIGEST_ROW_STRIDE].copy()
    d["radii_sub"] = r["radii"][::DIGEST_ROW_STRIDE].copy()
    for k, v in r["grads"].items():
        v = np.asarray(v)
        rows = v.reshape(scene["P"], -1)
        d["grad_sub_" + k] = rows[::DIGEST_ROW_STRIDE].copy()
        d["grad_sum_" + k] = np.float64(rows.astype(np.float64).sum())
        d["grad_abs_" + k] = np.float64(np.abs(rows.astype(np.float64)).sum())
        d["grad_max_" + k] = np.float64(np.abs(rows).max())
    return d


def main():
    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(dst, exist_ok=True)
    if "--digest" in sys.argv:  # python tools/make_golden.py <dst> --digest : the C1-size digest only
        scene = h.make_scene("C1")
        r = h.run_ref(scene)
        path = os.path.join(dst, "c1_digest.npz")
        np.savez_compressed(path, **digest(scene, r))
        print("c1_digest L =", r["num_rendered"], os.path.getsize(path), "bytes")
        return
    for name, (scene, kw) in variants().items():
        r = h.run_ref(scene, **kw)
        arrs = {k: r[k] for k in KEEP}
        arrs["num_rendered"] = np.int64(r["num_rendered"])
        for k, v in r["grads"].items():
            arrs["grad_" + k] = v
        for k, v in kw.items():
            arrs["in_" + k] = np.asarray(v, np.float32)
        arrs["meta"] = np.array([scene["P"], scene["cam"]["image_width"], scene["cam"]["image_height"],
                                 scene["sh_degree"]], np.int64)
        path = os.path.join(dst, name + ".npz")
        np.savez_compressed(path, **arrs)
        print(name, "L =", r["num_rendered"], os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
