"""Prints the blend-kernel counters of a `make STATS=1` build for one fwd+bwd view of a scene."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers as h  # noqa: E402
from diff_gaussian_rasterization import _abi  # noqa: E402

for name in sys.argv[1:] or ["C2"]:
    s = h.make_scene(name)
    lib = _abi.lib()
    buf = (C.c_uint64 * 16)()
    lib.gsr_debug_counters(buf, 1)
    g = h.run_gsr(s, fetch=True)
    torch.cuda.synchronize()
    lib.gsr_debug_counters(buf, 1)
    v = list(buf)
    L = g["num_rendered"]
    print(f"{name}: P={s['P']} L={L} sum_n_contrib={int(g['n_contrib'].astype(np.int64).sum())}")
    for tag, o in (("fwd", 0), ("bwd", 8)):
        rounds, tested, kept, whit, pairs, act = v[o:o + 6]
        print(f"  {tag}: rounds={rounds} tested={tested} (x{tested / max(L, 1):.2f} of 8L={8 * L}) kept={kept} ({kept / max(tested, 1):.3f})"
              f" warp_evals_with_hit={whit} ({whit / max(kept, 1):.3f} of kept) pairs={pairs} ({pairs / max(whit, 1):.2f}/warp-hit)"
              f" active_lanes/kept={act / max(kept, 1):.1f}")
