"""GPU box: times the (u32 key, u32 value) sort alone on the two shapes the forward pass uses -- depth bits of P = 1M
Gaussians (32 key bits, 4 passes) and tile ids of L = 5.1M pairs in depth order (13 bits at 1080p, 2 passes) -- plus a
large case.  Prints microseconds per sort (CUDA events, median of 20) and GB/s on 16 B/pair/pass + 4 B/pair histogram."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"))
from diff_gaussian_rasterization import _abi  # noqa: E402

lib = _abi.lib()
p = lambda t: C.c_void_p(t.data_ptr())


def run(name, keys, end_bit, reps=20):
    n = keys.numel()
    vals = torch.arange(n, dtype=torch.int32, device="cuda")
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    tb = lib.gsr_sort_u32_temp_bytes(n)
    temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for r in range(reps + 3):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = lib.gsr_sort_pairs_u32(p(keys), p(ko), p(vals), p(vo), n, None, end_bit, p(temp), tb, None)
        b.record()
        assert rc == 0, _abi.last_error()
        torch.cuda.synchronize()
        if r >= 3:
            ts.append(a.elapsed_time(b) * 1e3)
    passes = (end_bit + 7) // 8
    us = float(np.median(ts))
    copy = 0 if passes % 2 == 0 else 8 * n
    print(f"{name:28s} n={n:>11d} bits={end_bit:2d} passes={passes}  {us:9.1f} us  "
          f"{(n * (4 + 16 * passes) + 2 * copy) / us / 1e3:7.1f} GB/s (own bytes)")
    chk = torch.sort(keys.long() & ((1 << end_bit) - 1), stable=True)
    assert torch.equal(vo.long(), chk.indices), name


g = torch.Generator(device="cuda").manual_seed(1)
depth = (torch.rand(1_000_000, device="cuda", generator=g) * 3.0 + 1.5).view(torch.int32)
run("depth 1M", depth, 32)
run("depth 3M", (torch.rand(3_000_000, device="cuda", generator=g) * 3.0 + 1.5).view(torch.int32), 32)
run("tiles 5.1M (13 bit)", torch.randint(0, 8160, (5_100_000,), device="cuda", generator=g, dtype=torch.int32), 13)
run("tiles 15M (13 bit)", torch.randint(0, 8160, (15_000_000,), device="cuda", generator=g, dtype=torch.int32), 13)
run("tiles 200M (13 bit)", torch.randint(0, 8160, (200_000_000,), device="cuda", generator=g, dtype=torch.int32), 13)
