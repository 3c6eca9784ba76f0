"""GPU box: worst gradient tolerance ratio (|d| / (1e-3|ref| + 1e-4 max|ref|)) of the product vs the reference CUDA
on a scene, for the library named by GSR_B200_LIB.  Also ref-vs-ref (two runs) to show the reference's own noise."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import helpers as h
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
s = h.make_scene(name)
g = h.run_gsr(s, fetch=False)
r = h.run_ref(s, fetch=False)
r2 = h.run_ref(s, fetch=False)
out = {}
for k in ("means3D", "means2D", "opacity", "sh", "scales", "rotations"):
    a, b, c = g["grads"][k].reshape(-1), r["grads"][k].reshape(-1), r2["grads"][k].reshape(-1)
    b, c = b[: a.size], c[: a.size]
    out[k] = (round(h.grad_close(a, b)[1], 3), round(h.grad_close(c, b)[1], 3))
print(os.environ.get("GSR_B200_LIB", "default").split("/")[-1], name, "product-vs-ref / ref-vs-ref:", out)
