"""GPU box, library built with EXTRA=-DGSR_SORT_PROFILE (tools/build_variant.sh prof "-DGSR_SORT_PROFILE"; GSR_B200_LIB=...):
average cycles per phase of the sort's scatter kernel over the tiles of real forward passes (C2 by default)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as h  # noqa: E402
import diff_gaussian_rasterization as dgr  # noqa: E402
from diff_gaussian_rasterization import _abi  # noqa: E402

lib = _abi.lib()
s = h.make_scene(sys.argv[1] if len(sys.argv) > 1 else "C2")
rs = h.settings_for(s)
t = h.scene_tensors(s, requires_grad=False)
m2 = torch.zeros((s["P"], 3), device="cuda")
out = (C.c_uint64 * 8)()
for rep in range(4):
    if rep == 1:
        lib.gsr_debug_sort_profile(out, 1)
    dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"],
                               scales=t["scales"], rotations=t["rotations"])
    torch.cuda.synchronize()
lib.gsr_debug_sort_profile(out, 0)
v = np.array(list(out), dtype=np.float64)
names = ["tiles", "load", "rank", "count+scan", "look-back", "stage", "write-out"]
print("tiles (3 views, all passes):", int(v[0]))
for n, x in zip(names[1:], v[1:7]):
    print(f"  {n:12s} {x / max(v[0], 1):9.0f} cycles/tile")
print(f"  (of look-back: in-group walk of thread 0 {v[7] / max(v[0], 1):9.0f} cycles/tile)")
print(f"  total        {v[1:8].sum() / max(v[0], 1):9.0f} cycles/tile")
