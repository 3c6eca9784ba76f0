"""gsr_sort_pairs_u32_static (include/gsr_b200.h, csrc/static_scratch.cu): the (u32 key, u32 value) radix sort with
its temp block in a file-scope __device__ array of the library instead of a caller buffer -- for callers that cannot
allocate.  No counterpart in the reference.  Checked through the C ABI
against gsr_sort_pairs_u32 (the existing entry point, caller-supplied temp) on the same inputs: integer work, so the
outputs must be identical bit for bit; and against numpy's stable argsort.  Every case runs twice in a row (the
array keeps the previous call's contents), and the cases follow each other from large to small and back.

Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_PAIRS = 1 << 20  # GSR_STATIC_SORT_MAX_PAIRS
GSR_ERR_CAPACITY = -3


def _p(t):
    return C.c_void_p(t.data_ptr())


def _inputs(n, seed):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    if n > 1000:
        keys[: n // 2] = keys[rng.integers(0, 37, n // 2)]  # many duplicates -> stability matters
    vals = rng.permutation(n).astype(np.uint32)
    return keys, vals


def _existing(lib, k_in, v_in, n, end_bit):
    from diff_gaussian_rasterization import _abi
    k_out, v_out = torch.full_like(k_in, -1), torch.full_like(v_in, -1)
    tb = lib.gsr_sort_u32_temp_bytes(n)
    temp = torch.empty(max(tb, 1), dtype=torch.uint8, device="cuda")
    r = lib.gsr_sort_pairs_u32(_p(k_in), _p(k_out), _p(v_in), _p(v_out), n, None, end_bit, _p(temp), tb, None)
    assert r == 0, _abi.last_error()
    torch.cuda.synchronize()
    return k_out, v_out


def _static(lib, k_in, v_in, n, end_bit):
    from diff_gaussian_rasterization import _abi
    k_out, v_out = torch.full_like(k_in, -1), torch.full_like(v_in, -1)
    torch.cuda.synchronize()
    r = lib.gsr_sort_pairs_u32_static(_p(k_in), _p(k_out), _p(v_in), _p(v_out), n, end_bit, None)
    assert r == 0, _abi.last_error()
    torch.cuda.synchronize()
    return k_out, v_out


def test_static_scratch_sort_matches_the_caller_buffer_sort_twice_in_a_row():
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()
    cases = [(MAX_PAIRS, 32), (4097, 13), (300_000, 17), (1, 32), (MAX_PAIRS, 13), (70_001, 9), (MAX_PAIRS - 3, 24)]
    for i, (n, end_bit) in enumerate(cases):
        for rep in range(2):  # two calls in a row, fresh inputs each time
            keys, vals = _inputs(n, 1000 * i + rep)
            k_in = torch.from_numpy(keys.view(np.int32)).cuda()
            v_in = torch.from_numpy(vals.view(np.int32)).cuda()
            k_want, v_want = _existing(lib, k_in, v_in, n, end_bit)
            k_got, v_got = _static(lib, k_in, v_in, n, end_bit)
            assert torch.equal(k_got, k_want) and torch.equal(v_got, v_want), (n, end_bit, rep)
            assert torch.equal(k_in, torch.from_numpy(keys.view(np.int32)).cuda())  # inputs untouched
            mask = np.uint32((1 << end_bit) - 1) if end_bit < 32 else np.uint32(0xFFFFFFFF)
            order = np.argsort(keys & mask, kind="stable")
            assert np.array_equal(k_got.cpu().numpy().view(np.uint32), keys[order]), (n, end_bit, rep)
            assert np.array_equal(v_got.cpu().numpy().view(np.uint32), vals[order]), (n, end_bit, rep)


def test_static_scratch_sort_back_to_back_on_one_stream_without_a_host_sync():
    """Calls on one stream are ordered by the stream: the second may be queued while the first still runs."""
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()
    sets = []
    for j, (n, end_bit) in enumerate([(900_000, 32), (123_457, 13), (900_000, 20)]):
        keys, vals = _inputs(n, 77 + j)
        k_in = torch.from_numpy(keys.view(np.int32)).cuda()
        v_in = torch.from_numpy(vals.view(np.int32)).cuda()
        sets.append((n, end_bit, keys, vals, k_in, v_in, torch.full_like(k_in, -1), torch.full_like(v_in, -1)))
    torch.cuda.synchronize()
    for n, end_bit, _, _, k_in, v_in, k_out, v_out in sets:
        r = lib.gsr_sort_pairs_u32_static(_p(k_in), _p(k_out), _p(v_in), _p(v_out), n, end_bit, None)
        assert r == 0, _abi.last_error()
    torch.cuda.synchronize()
    for n, end_bit, keys, vals, k_in, v_in, k_out, v_out in sets:
        k_want, v_want = _existing(lib, k_in, v_in, n, end_bit)
        assert torch.equal(k_out, k_want) and torch.equal(v_out, v_want), (n, end_bit)


def test_static_scratch_sort_rejects_sizes_beyond_its_maximum():
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()
    n = MAX_PAIRS + 1
    k = torch.zeros(n, dtype=torch.int32, device="cuda")
    v = torch.zeros(n, dtype=torch.int32, device="cuda")
    k_out, v_out = torch.full_like(k, -1), torch.full_like(v, -1)
    r = lib.gsr_sort_pairs_u32_static(_p(k), _p(k_out), _p(v), _p(v_out), n, 32, None)
    assert r == GSR_ERR_CAPACITY
    assert "exceed the built-in scratch" in _abi.last_error() and str(MAX_PAIRS) in _abi.last_error()
    torch.cuda.synchronize()
    assert bool((k_out == -1).all()) and bool((v_out == -1).all())  # nothing was launched
    assert lib.gsr_sort_pairs_u32_static(None, None, None, None, 0, 32, None) == 0  # empty input: nothing to do
    assert lib.gsr_sort_pairs_u32_static(_p(k), _p(k_out), _p(v), _p(v_out), 10, 0, None) != 0  # bad end_bit
