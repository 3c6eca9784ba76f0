"""world_size-2 gloo tests (CPU) of the view-sharded multi-GPU path (diff_gaussian_rasterization.dist):
sharding covers every view exactly once, the flat bucket receives autograd's accumulation, and the
all-reduced gradient / loss equal the single-process result.  The per-view "render" here is a small
differentiable stand-in (the rasterizer itself needs a GPU; its kernels are covered by the gpu tests)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_params(seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = {"xyz": (50, 3), "shs": (50, 16, 3), "scales": (50, 3), "rotations": (50, 4), "opacity": (50, 1)}
    return {n: torch.randn(s, generator=g).requires_grad_(True) for n, s in shapes.items()}


def _views(n):
    g = torch.Generator().manual_seed(123)
    return [torch.randn(5, generator=g) for _ in range(n)]


def _fake_render_backward(params):
    def fn(view):
        loss = sum(((p * view[i]).sin() * (i + 1)).sum() for i, p in enumerate(params.values()))
        loss.backward()
        return loss
    return fn


def _worker(rank, world, port, n_views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from diff_gaussian_rasterization.dist import FlatGradBucket, sharded_step
        params = _make_params()
        bucket = FlatGradBucket(params)
        total = sharded_step(_views(n_views), _fake_render_backward(params), bucket)
        # a second step must start from zero (no accumulation across steps)
        total2 = sharded_step(_views(n_views), _fake_render_backward(params), bucket)
        out[rank] = (bucket.flat.clone(), float(total), float(total2), float(bucket.global_norm()))
    finally:
        dist.destroy_process_group()


def _single(n_views):
    from diff_gaussian_rasterization.dist import FlatGradBucket, sharded_step
    params = _make_params()
    bucket = FlatGradBucket(params)
    total = sharded_step(_views(n_views), _fake_render_backward(params), bucket, rank=0, world=1)
    return bucket.flat.clone(), float(total)


@pytest.mark.parametrize("n_views", [8, 5])
def test_two_rank_sharded_step_matches_single_process(n_views):
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "3d-gaussian-splat-attack_b200"))
    mgr = mp.Manager()
    out = mgr.dict()
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n_views, out), nprocs=2, join=True)
    ref_flat, ref_total = _single(n_views)
    f0, t0, t0b, n0 = out[0]
    f1, t1, t1b, n1 = out[1]
    assert torch.equal(f0, f1), "ranks disagree after all-reduce"
    assert torch.allclose(f0, ref_flat, rtol=1e-5, atol=1e-6)
    assert abs(t0 - ref_total) <= 1e-4 * abs(ref_total) and t0 == t1
    assert abs(t0b - t0) <= 1e-5 * abs(t0)
    assert abs(n0 - float(ref_flat.norm())) <= 1e-4 * n0


def test_shard_views_partition():
    from diff_gaussian_rasterization.dist import shard_views
    for n in (0, 1, 7, 64):
        for world in (1, 2, 3, 8):
            seen = sorted(k for r in range(world) for k in shard_views(n, r, world))
            assert seen == list(range(n))
            sizes = [len(shard_views(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    assert shard_views(64, 3, 8) == list(range(3, 64, 8))
    with pytest.raises(ValueError):
        shard_views(4, 2, 2)


def test_bucket_binds_grads_and_rebinds_after_set_to_none():
    from diff_gaussian_rasterization.dist import FlatGradBucket
    params = _make_params()
    b = FlatGradBucket(params)
    # every slice starts on a 256-byte boundary (P = 50 is not a multiple of 4: 16-byte vector accesses on the
    # shs / rotations slices need it), so the bucket is the parameters plus padding
    assert b.numel >= sum(p.numel() for p in params.values()) == 50 * (3 + 48 + 3 + 4 + 1)
    assert all(a % 64 == 0 and e - a == params[n].numel() for n, (a, e) in b.slices.items())
    assert all(b.grad(n).data_ptr() % 256 == b.flat.data_ptr() % 256 for n in params)
    (params["xyz"] * 2).sum().backward()
    (params["xyz"] * 3).sum().backward()
    assert torch.equal(b.grad("xyz"), torch.full((50, 3), 5.0))
    assert b.grad("xyz").data_ptr() == b.flat.data_ptr()
    params["xyz"].grad = None  # optimizer.zero_grad(set_to_none=True), attack.py:603
    b.zero()
    (params["xyz"] * 7).sum().backward()
    assert torch.equal(b.flat[:150], torch.full((150,), 7.0))
    assert b.all_reduce() is None  # no process group: no-op
    with pytest.raises(ValueError):
        FlatGradBucket({"a": torch.zeros(3)})


# ---------------------------------------------------------------- accumulate-API step (views.forward_view / backward_view)
def _fake_view_api():
    """Stand-in for views.forward_view / backward_view on CPU: image = a fixed linear map of sin(inputs), whose
    backward ADDS into the given buffers (like gsr_backward_accumulate)."""
    def fwd(rs, means3D, opacities, shs=None, scales=None, rotations=None):
        ins = dict(means3D=means3D, opacities=opacities, shs=shs, scales=scales, rotations=rotations)
        img = sum(((t * rs[i]).sin() * (i + 1)).sum() for i, t in enumerate(ins.values())) * torch.ones(3, 2, 2)
        return img, None, (rs, ins)

    def bwd(state, dL_dimage, grads):
        rs, ins = state
        s = dL_dimage.sum()
        for i, (n, t) in enumerate(ins.items()):
            grads[n] += s * (i + 1) * rs[i] * (t * rs[i]).cos()
    return fwd, bwd


def _acc_inputs():
    raw = _make_params(seed=4)
    raw = {"means3D": raw["xyz"], "opacities": raw["opacity"], "shs": raw["shs"], "scales": raw["scales"],
           "rotations": raw["rotations"]}
    # activations as in scene/gaussian_model.py:97-121
    act = {"means3D": raw["means3D"], "opacities": torch.sigmoid(raw["opacities"]), "shs": raw["shs"],
           "scales": torch.exp(raw["scales"]), "rotations": torch.nn.functional.normalize(raw["rotations"])}
    return raw, act


def _acc_worker(rank, world, port, n_views, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from diff_gaussian_rasterization.dist import FlatGradBucket, sharded_accumulate_step, backprop_inputs
        raw, act = _acc_inputs()
        bucket = FlatGradBucket(act, bind=False)
        fwd, bwd = _fake_view_api()
        total = sharded_accumulate_step(_views(n_views), act, lambda img, k: (img * (k + 1)).sum(), bucket,
                                        forward_view=fwd, backward_view=bwd)
        backprop_inputs(act, bucket)
        out[rank] = (bucket.flat.clone(), float(total), {n: p.grad.clone() for n, p in raw.items()})
    finally:
        dist.destroy_process_group()


def test_two_rank_accumulate_step_matches_autograd_reference():
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "3d-gaussian-splat-attack_b200"))
    n_views = 5
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_acc_worker, args=(2, _free_port(), n_views, out), nprocs=2, join=True)
    # reference: plain autograd through the same fake renderer, one process
    raw, act = _acc_inputs()
    fwd, _ = _fake_view_api()
    loss = sum((fwd(v, **act)[0] * (k + 1)).sum() for k, v in enumerate(_views(n_views)))
    loss.backward()
    f0, t0, g0 = out[0]
    f1, t1, g1 = out[1]
    assert torch.equal(f0, f1) and t0 == t1
    assert abs(t0 - float(loss)) <= 1e-4 * abs(float(loss))
    for n, p in raw.items():
        assert torch.allclose(g0[n], p.grad, rtol=1e-4, atol=1e-5), n
        assert torch.equal(g0[n], g1[n])


def test_host_io_shard_layout_tiles_the_flat_vector():
    """pipeline.shard_layout: the per-rank host slices of the sharded upload / read-back tile [0, n_param) exactly, the
    padded device length splits into equal slices (all_gather_into_tensor / reduce_scatter_tensor), for any world size."""
    from diff_gaussian_rasterization.pipeline import shard_layout
    from diff_gaussian_rasterization.dist import flat_layout
    for P in (1, 7, 2001, 1_000_000):
        _, n_param = flat_layout({"means3D": 3 * P, "shs": 48 * P, "scales": 3 * P, "rotations": 4 * P, "opacities": P})
        for world in (1, 2, 3, 4, 8):
            seen = 0
            for rank in range(world):
                n_pad, shard, (lo, hi) = shard_layout(n_param, world, rank)
                assert n_pad == shard * world and n_param <= n_pad < n_param + world
                assert lo == seen and hi - lo <= shard
                seen = hi
            assert seen == n_param
    with pytest.raises(ValueError):
        shard_layout(10, 2, 2)
