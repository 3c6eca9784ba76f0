"""Diagnostic run on the GPU box: product vs reference CUDA (oracle/_ref) vs CPU oracle.
Prints mismatch statistics for every intermediate and output. Not a test; see tests/ for assertions."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as h  # noqa: E402


def cmp_int(name, a, b):
    a, b = np.asarray(a).reshape(-1), np.asarray(b).reshape(-1)
    if a.shape != b.shape:
        print(f"   {name:14s} SHAPE {a.shape} vs {b.shape}")
        return
    bad = int((a != b).sum())
    print(f"   {name:14s} mismatches {bad} / {a.size}")


def cmp_f(name, a, b, mask=None):
    a, b = np.asarray(a, np.float64).reshape(-1), np.asarray(b, np.float64).reshape(-1)
    if a.shape != b.shape:
        print(f"   {name:14s} SHAPE {a.shape} vs {b.shape}")
        return
    bits = int((np.asarray(a, np.float32).view(np.uint32) != np.asarray(b, np.float32).view(np.uint32)).sum())
    d = np.abs(a - b)
    print(f"   {name:14s} bit-mismatch {bits} / {a.size}  max|d| {d.max() if d.size else 0:.3e}  max|ref| {np.abs(b).max() if b.size else 0:.3e}")


def compare(tag, a, b, vis):
    print(f"  -- {tag}: L {a['num_rendered']} vs {b['num_rendered']}")
    cmp_int("radii", a["radii"], b["radii"])
    for k in ("tiles_touched", "point_offsets"):
        cmp_int(k, a[k], b[k])
    P = a["radii"].size
    for k, w in (("depths", 1), ("means2D", 2), ("conic_opacity", 4), ("cov3D", 6), ("rgb", 3)):
        cmp_f(k, a[k].reshape(P, w)[vis], b[k].reshape(P, w)[vis])
    for k in ("keys_sorted", "point_list", "ranges", "n_contrib"):
        cmp_int(k, a[k], b[k])
    cmp_f("final_T", a["final_T"], b["final_T"])
    cmp_f("color", a["color"], b["color"])
    if "grads" in a and "grads" in b:
        for k in a["grads"]:
            if k in b["grads"]:
                ok, worst = h.grad_close(a["grads"][k], b["grads"][k])
                x, y = a["grads"][k].reshape(-1).astype(np.float64), b["grads"][k].reshape(-1).astype(np.float64)
                rel = np.abs(x - y).max() / (np.abs(y).max() + 1e-30)
                print(f"   grad {k:10s} max|d|/max|ref| {rel:.3e}   tol-ratio {worst:.3f} {'ok' if ok else 'FAIL'}")


def main():
    names = sys.argv[1:] or ["tiny", "small", "C1"]
    for name in names:
        s = h.make_scene(name)
        print(f"== scene {name}: P={s['P']} {s['cam']['image_width']}x{s['cam']['image_height']}")
        t0 = time.time(); g = h.run_gsr(s); torch.cuda.synchronize(); print(f"  gsr  {time.time() - t0:.3f}s")
        vis = g["radii"] > 0
        if h.ref_available():
            t0 = time.time(); r = h.run_ref(s); print(f"  ref  {time.time() - t0:.3f}s")
            compare("gsr vs refCUDA", g, r, vis)
        if s["P"] <= 200_000:
            t0 = time.time(); o = h.run_oracle(s); print(f"  oracle {time.time() - t0:.3f}s n_pairs={o['n_pairs']}")
            compare("gsr vs oracle", g, o, vis)
            if h.ref_available():
                compare("oracle vs refCUDA", o, r, vis)


if __name__ == "__main__":
    main()
