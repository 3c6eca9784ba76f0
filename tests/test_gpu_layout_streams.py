"""GPU tests of buffer layout and stream ordering: Gaussian counts that are not a multiple of 4 through every flat-buffer
path (FlatGradBucket, HostStepPipeline), misaligned views, camera tensors as the reference's Camera produces them
(transposed / on the CPU) through the three-stream ViewPipeline, more than 65536 tiles, and two streams through the
drop-in module at once."""
import numpy as np
import pytest
import torch

import helpers as h

pytestmark = pytest.mark.gpu


def _cams_settings(s, n, W, H):
    from gsr_synth import ring_cameras
    cams = ring_cameras(n, width=W, height=H)
    out = []
    for c in cams:
        sc = dict(s); sc["cam"] = c
        out.append(h.settings_for(sc))
    return cams, out


def _autograd_grads(s, settings, dLs):
    import diff_gaussian_rasterization as dgr
    t = h.scene_tensors(s)
    for rs, dL in zip(settings, dLs):
        m2 = torch.zeros((s["P"], 3), device="cuda", requires_grad=True)
        img, _ = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"],
                                            scales=t["scales"], rotations=t["rotations"])
        (img * dL).sum().backward()
    return {k: v.grad for k, v in t.items()}


@pytest.mark.parametrize("P", [1999, 2001, 2002])
def test_odd_gaussian_counts_through_bucket_and_host_pipeline(P):
    """Slices of the flat buffers start on 256-byte boundaries for any P (dist.flat_layout): the 16-byte vector loads
    of shs / rotations and the vector reductions into their gradients stay aligned."""
    from diff_gaussian_rasterization.dist import FlatGradBucket, sharded_accumulate_step, flat_layout
    from diff_gaussian_rasterization.pipeline import HostStepPipeline
    s = h.make_scene("tiny", P=P)
    W, H = 160, 120
    cams, settings = _cams_settings(s, 3, W, H)
    rng = np.random.default_rng(P)
    dL_np = (rng.standard_normal((3, 3, H, W)) / (3 * H * W)).astype(np.float32)
    dLs = [torch.from_numpy(d).cuda() for d in dL_np]
    want = _autograd_grads(s, settings, dLs)

    t = h.scene_tensors(s, requires_grad=False)
    bucket = FlatGradBucket(t, bind=False)
    assert all(bucket.grad(n).data_ptr() % 256 == 0 for n in t)
    sharded_accumulate_step(settings, t, lambda img, k: (img * dLs[k]).sum(), bucket, rank=0, world=1)
    for k in t:
        ok, worst = h.grad_close(bucket.grad(k).cpu().numpy(), want[k].cpu().numpy())
        assert ok, ("bucket", k, worst)

    widths = dict(means3D=3, shs=48, scales=3, rotations=4, opacities=1)
    pipe = HostStepPipeline(P, 16, W, H, 3, cams[0]["tanfovx"], cams[0]["tanfovy"], 3, "cuda", depth=2, group=False)
    slices, n_flat = flat_layout({n: P * w for n, w in widths.items()})
    assert slices == pipe.slices and n_flat == pipe.n_param
    host_param = torch.zeros(n_flat).pin_memory()
    for n in widths:
        host_param[slices[n][0]:slices[n][1]] = torch.from_numpy(s[n].reshape(-1))
    host_cam = torch.from_numpy(np.stack([np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1),
                                                          c["campos"], np.zeros(1, np.float32)]) for c in cams])).pin_memory()
    host_dL = torch.from_numpy(dL_np).pin_memory()
    outs = [torch.empty(n_flat).pin_memory() for _ in range(3)]
    done = []
    for o in outs:
        done += pipe.submit(host_param, host_cam, host_dL, o)
    done += pipe.flush()
    assert len(done) == 3
    for o in done:
        for n in widths:
            ok, worst = h.grad_close(o[slices[n][0]:slices[n][1]].numpy(), want[n].cpu().numpy().reshape(-1))
            assert ok, ("host pipeline", n, worst)


def test_misaligned_views_are_realigned_or_rejected():
    """Inputs that are views at a 4-byte offset into a packed buffer are re-aligned by the Python layer; a misaligned
    gradient buffer of the accumulate API is an error (the C ABI returns GSR_ERR_INVALID rather than faulting)."""
    import diff_gaussian_rasterization as dgr
    from diff_gaussian_rasterization import views, attack_ops as ao
    s = h.make_scene("tiny", P=501)
    P = s["P"]
    rs = h.settings_for(s)
    t = h.scene_tensors(s, requires_grad=False)
    packed = torch.zeros(1 + P * 52, device="cuda")
    packed[1:1 + P * 48] = t["shs"].reshape(-1)
    packed[1 + P * 48:] = t["rotations"].reshape(-1)
    shs_v, rot_v = packed[1:1 + P * 48].view(P, 16, 3), packed[1 + P * 48:].view(P, 4)
    assert shs_v.data_ptr() % 16 != 0
    m2 = torch.zeros((P, 3), device="cuda")
    a, _ = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=shs_v,
                                      scales=t["scales"], rotations=rot_v)
    b, _ = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"],
                                      scales=t["scales"], rotations=t["rotations"])
    assert torch.equal(a, b)
    img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
                                        rotations=t["rotations"])
    grads = {k: torch.zeros_like(v) for k, v in t.items()}
    grads["shs"] = torch.zeros(1 + P * 48, device="cuda")[1:].view(P, 16, 3)
    with pytest.raises(RuntimeError, match="16-byte aligned"):
        views.backward_view(st, torch.ones_like(img), grads)
    # the fused PGD update takes slices that start anywhere (scalar path when a pointer is not 16-byte aligned)
    p0 = torch.randn(4099, device="cuda")
    g = torch.randn(4099, device="cuda")
    p = p0.clone() + 0.01 * torch.randn(4099, device="cuda")
    want = ((p - 0.02 * torch.sign(g)) - p0).clamp(-0.05, 0.05) + p0
    pv = torch.zeros(4100, device="cuda")
    pv[1:] = p
    ao.pgd_step_({"x": pv[1:]}, {"x": g}, {"x": p0}, 0.02, 0.05, norm="linf")
    assert torch.equal(pv[1:], want)
    torch.cuda.synchronize()


@pytest.mark.parametrize("where", ["transposed", "cpu"])
def test_view_pipeline_takes_reference_style_camera_tensors(where):
    """Camera.world_view_transform is `.transpose(0, 1).cuda()` (scene/cameras.py:54): a non-contiguous CUDA tensor whose
    .contiguous() copy runs on the caller's stream; the geometry stream must not read the matrix before it is written.
    A large fill queued on the caller's stream first makes an unordered read observable."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("small")
    W, H = 320, 240
    _, settings = _cams_settings(s, 6, W, H)
    t = h.scene_tensors(s, requires_grad=False)
    rng = np.random.default_rng(3)
    dLs = [torch.from_numpy(rng.standard_normal((3, H, W)).astype(np.float32)).cuda() for _ in settings]
    want_imgs, want = [], {k: torch.zeros_like(v) for k, v in t.items()}
    for rs, dL in zip(settings, dLs):
        img, _, st = views.forward_view(rs, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
                                        rotations=t["rotations"])
        views.backward_view(st, dL, want)
        want_imgs.append(img)

    def restyle(rs):
        if where == "cpu":
            return rs._replace(viewmatrix=rs.viewmatrix.cpu(), projmatrix=rs.projmatrix.cpu(), campos=rs.campos.cpu(),
                               bg=rs.bg.cpu())
        tr = lambda m: m.t().contiguous().t()  # same values, column-major storage: not contiguous
        v, p = tr(rs.viewmatrix), tr(rs.projmatrix)
        assert not v.is_contiguous() and torch.equal(v, rs.viewmatrix)
        return rs._replace(viewmatrix=v, projmatrix=p)

    vp = views.ViewPipeline(overlap=True)
    ballast = torch.empty(64 << 20, device="cuda")
    for rep in range(2):
        styled = [restyle(rs) for rs in settings]
        got, imgs = {k: torch.zeros_like(v) for k, v in t.items()}, []
        for _ in range(4):
            ballast.normal_()  # ~1 ms of work on the caller's stream ahead of the camera copies
        vp.step(styled, lambda img, k: (imgs.append(img.clone()), dLs[k])[1], got, t["means3D"], t["opacities"],
                shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
        torch.cuda.synchronize()
        for a, b in zip(imgs, want_imgs):
            assert torch.equal(a, b)
        for k in want:
            ok, worst = h.grad_close(got[k].cpu().numpy(), want[k].cpu().numpy())
            assert ok, (k, worst)


def test_more_than_65536_tiles():
    """4096 x 4112 pixels = 65792 tiles: tile ids need 17 bits, i.e. three sort passes and no fused histogram rows."""
    s = h.make_scene("tiny", width=4096, height=4112)
    g = h.run_gsr(s, grad=False, fetch=False)
    assert h.ref_available()
    r = h.run_ref(s, grad=False, fetch=False)
    assert np.array_equal(g["color"], r["color"]) and np.array_equal(g["radii"], r["radii"])


def test_two_streams_through_the_drop_in_module():
    """Scratch, accumulators and the count word are per (device, stream): forward + backward issued on two streams
    without host synchronisation in between give the single-stream results."""
    import diff_gaussian_rasterization as dgr
    s = h.make_scene("small")
    _, settings = _cams_settings(s, 2, 320, 240)
    dL = torch.from_numpy(np.random.default_rng(0).standard_normal((3, 240, 320)).astype(np.float32)).cuda()

    def run(rs):
        t = h.scene_tensors(s)
        m2 = torch.zeros((s["P"], 3), device="cuda", requires_grad=True)
        img, _ = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"],
                                            scales=t["scales"], rotations=t["rotations"])
        (img * dL).sum().backward()
        return img.detach(), {k: v.grad for k, v in t.items()}

    want = [run(rs) for rs in settings]
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    got = [None, None]
    for rep in range(3):
        for i, (rs, st) in enumerate(zip(settings, streams)):
            st.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(st):
                got[i] = run(rs)
        torch.cuda.synchronize()
        for (ia, ga), (ib, gb) in zip(got, want):
            assert torch.equal(ia, ib)
            for k in ga:
                ok, worst = h.grad_close(ga[k].cpu().numpy(), gb[k].cpu().numpy())
                assert ok, (k, worst)


def test_speculative_forward_overflow_is_redone_exactly():
    """Steady-state forwards size point_list / scratch from the previous view of the same shape and launch binning +
    blending before the host has seen the duplicate count (gsr_forward_render_speculative).  A view whose count
    exceeds 1.25 x the previous one overflows that capacity and must be redone with the exact size: same image,
    same saved state, same gradients as with speculation off; a count far BELOW the capacity must work as well."""
    import diff_gaussian_rasterization as dgr
    from gsr_synth import look_at_camera
    s = h.make_scene("small")
    W, H = 320, 240
    far, near = look_at_camera((0, 0, -9.0), width=W, height=H), look_at_camera((0, 0, -2.2), width=W, height=H)
    dL = torch.from_numpy(np.random.default_rng(1).standard_normal((3, H, W)).astype(np.float32)).cuda()

    def run(cam):
        sc = dict(s); sc["cam"] = cam
        t = h.scene_tensors(s)
        m2 = torch.zeros((s["P"], 3), device="cuda", requires_grad=True)
        img, radii = dgr.GaussianRasterizer(h.settings_for(sc))(
            means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"], scales=t["scales"],
            rotations=t["rotations"])
        L = img.grad_fn.num_rendered
        (img * dL).sum().backward()
        return img.detach().clone(), L, {k: v.grad.clone() for k, v in t.items()}

    old = dgr.SPECULATE
    try:
        dgr.SPECULATE = False
        want = [run(c) for c in (far, near, far)]
        assert want[1][1] > 1.25 * want[0][1] + 4096, "the near view must overflow the far view's capacity"
        dgr.SPECULATE = True
        dgr._states.clear()  # forget the shapes seen so far
        got = [run(c) for c in (far, near, far)]  # not speculated / overflow + redo / capacity far above the count
        again = [run(c) for c in (far, near, far)]  # speculated with a fitting capacity
    finally:
        dgr.SPECULATE = old
    for res in (got, again):
        for (ia, La, ga), (ib, Lb, gb) in zip(res, want):
            assert La == Lb and torch.equal(ia, ib)
            for k in ga:
                ok, worst = h.grad_close(ga[k].cpu().numpy(), gb[k].cpu().numpy())
                assert ok, (k, worst)
