"""GPU tests of the multi-view backward (views.backward_views -> gsr_backward_blend_view per view +
gsr_backward_preprocess_views once per group): it must equal the sum of per-view backward_view calls -- the arithmetic
of Rasterizer::backward (rasterizer_impl.cu:340-433) per view -- up to summation order."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers as h

pytestmark = pytest.mark.gpu

TOL = dict(rtol=1e-3, atol_frac=1e-4)  # the gradient tolerance of every other test (helpers.grad_close)


def _views_of(s, n, W=160, H=120):
    from gsr_synth import ring_cameras
    cams = ring_cameras(n, width=W, height=H)
    out = []
    for c in cams:
        sc = dict(s); sc["cam"] = c
        out.append(h.settings_for(sc))
    return out


def _forward_all(settings, t, **kw):
    from diff_gaussian_rasterization import views
    states = []
    for rs in settings:
        img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], **kw)
        states.append(st)
    return states


def _per_view_reference(states, dLs, grads_like, with_m2=False):
    from diff_gaussian_rasterization import views
    ref = {k: torch.zeros_like(v) for k, v in grads_like.items()}
    m2s = []
    for st, dL in zip(states, dLs):
        if with_m2:
            ref["means2D"] = torch.zeros((st.P, 3), device="cuda")
        views.backward_view(st, dL, ref)
        if with_m2:
            m2s.append(ref.pop("means2D"))
    return ref, m2s


@pytest.mark.parametrize("n_views", [1, 3, 8, 11, 16, 19])
def test_backward_views_equals_sum_of_per_view_backwards(n_views):
    """SH degree 3, scale / rotation inputs; 19 views = one full group of 16 plus a group of 3."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("small")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, n_views)
    rng = np.random.default_rng(11 + n_views)
    dLs = [torch.from_numpy(rng.standard_normal((3, 120, 160)).astype(np.float32)).cuda() for _ in settings]
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    states = _forward_all(settings, t, **kw)
    ref, ref_m2 = _per_view_reference(states, dLs, t, with_m2=True)
    # (a) overwrite mode into buffers full of garbage: every row must be written
    got = {k: torch.full_like(v, 123.0) for k, v in t.items()}
    m2 = [torch.full((s["P"], 3), 7.0, device="cuda") for _ in settings]
    views.backward_views(states, dLs, got, accumulate=False, means2D_grads=m2)
    for k in ref:
        ok, worst = h.grad_close(got[k].cpu().numpy(), ref[k].cpu().numpy(), **TOL)
        assert ok, ("overwrite", k, worst)
    for a, b in zip(m2, ref_m2):
        ok, worst = h.grad_close(a.cpu().numpy(), b.cpu().numpy(), **TOL)
        assert ok, ("means2D", worst)
    # (b) accumulate mode adds to what is there
    base = {k: torch.from_numpy(rng.standard_normal(tuple(v.shape)).astype(np.float32)).cuda() for k, v in t.items()}
    got = {k: v.clone() for k, v in base.items()}
    views.backward_views(states, dLs, got, accumulate=True)
    for k in ref:
        ok, worst = h.grad_close((got[k] - base[k]).cpu().numpy(), ref[k].cpu().numpy(), rtol=1e-3, atol_frac=2e-4)
        assert ok, ("accumulate", k, worst)


def test_backward_views_rows_no_view_rendered():
    """Gaussians behind every camera: overwrite mode zeroes their rows, accumulate mode leaves them alone."""
    from diff_gaussian_rasterization import views
    from gsr_synth import look_at_camera
    s = h.make_scene("tiny")
    t = h.scene_tensors(s, requires_grad=False)
    sc = dict(s); sc["cam"] = look_at_camera((0.0, 0.0, -0.3), width=160, height=120)  # inside the cloud: part of it
    settings = [h.settings_for(sc)] * 3                                                # is behind the camera
    dLs = [torch.full((3, 120, 160), float(i + 1), device="cuda") for i in range(3)]
    states = _forward_all(settings, t, shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    seen = torch.zeros(s["P"], dtype=torch.bool, device="cuda")
    for st in states:
        seen |= st.radii > 0
    assert bool((~seen).any()) and bool(seen.any())
    got = {k: torch.full_like(v, 5.0) for k, v in t.items()}
    views.backward_views(states, dLs, got, accumulate=False)
    for k, v in got.items():
        assert float(v[~seen].abs().max()) == 0.0, k
    got = {k: torch.full_like(v, 5.0) for k, v in t.items()}
    views.backward_views(states, dLs, got, accumulate=True)
    for k, v in got.items():
        assert bool((v[~seen] == 5.0).all()), k


def test_backward_views_precomputed_colour_and_covariance_and_low_sh_degree():
    from diff_gaussian_rasterization import views
    # colours + 3D covariances given: dL_dcolor and dL_dcov3D are outputs, no SH / scale / rotation path
    sc, kw, gold = h.golden_case("tiny_precomp")
    t = h.scene_tensors(sc, requires_grad=False)
    col = torch.from_numpy(kw["colors_precomp"]).cuda()
    cov = torch.from_numpy(kw["cov3D_precomp"]).cuda()
    settings = _views_of(sc, 4, W=sc["cam"]["image_width"], H=sc["cam"]["image_height"])
    rng = np.random.default_rng(3)
    dLs = [torch.from_numpy(rng.standard_normal((3, settings[0].image_height, settings[0].image_width)).astype(np.float32)).cuda()
           for _ in settings]
    states = _forward_all(settings, t, colors_precomp=col, cov3D_precomp=cov)
    like = dict(means3D=t["means3D"], opacities=t["opacities"], colors_precomp=col, cov3D_precomp=cov)
    ref, _ = _per_view_reference(states, dLs, like)
    got = {k: torch.full_like(v, -3.0) for k, v in like.items()}
    views.backward_views(states, dLs, got, accumulate=False)
    for k in ref:
        ok, worst = h.grad_close(got[k].cpu().numpy(), ref[k].cpu().numpy(), **TOL)
        assert ok, (k, worst)
    # SH degree 1 with M = 16 rows, and an M = 4 layout (scalar row path)
    sc, kw, gold = h.golden_case("tiny_sh1_bg")
    t = h.scene_tensors(sc, requires_grad=False)
    settings = _views_of(sc, 3, W=sc["cam"]["image_width"], H=sc["cam"]["image_height"])
    dLs = [torch.from_numpy(rng.standard_normal((3, settings[0].image_height, settings[0].image_width)).astype(np.float32)).cuda()
           for _ in settings]
    for shs in (t["shs"], t["shs"][:, :4, :].contiguous()):
        tt = dict(t); tt["shs"] = shs
        states = _forward_all(settings, tt, shs=shs, scales=t["scales"], rotations=t["rotations"])
        ref, _ = _per_view_reference(states, dLs, tt)
        got = {k: torch.full_like(v, 9.0) for k, v in tt.items()}
        views.backward_views(states, dLs, got, accumulate=False)
        for k in ref:
            ok, worst = h.grad_close(got[k].cpu().numpy(), ref[k].cpu().numpy(), **TOL)
            assert ok, (tuple(shs.shape), k, worst)


def test_backward_views_argument_errors():
    from diff_gaussian_rasterization import views, _abi
    s = h.make_scene("tiny")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, 2)
    states = _forward_all(settings, t, shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    dLs = [torch.ones((3, 120, 160), device="cuda") for _ in settings]
    grads = {k: torch.zeros_like(v) for k, v in t.items()}
    with pytest.raises(Exception):
        views.backward_views(states, dLs[:1], grads)
    with pytest.raises(Exception):
        views.backward_views(states, dLs, {k: v for k, v in grads.items() if k != "shs"})
    other = h.scene_tensors(s, requires_grad=False)  # same values, different tensors: not "the same inputs"
    st2 = _forward_all(settings[:1], other, shs=other["shs"], scales=other["scales"], rotations=other["rotations"])
    with pytest.raises(Exception):
        views.backward_views([states[0], st2[0]], dLs, grads)
    # the C entry point rejects more than GSR_MAX_FUSED_VIEWS views per call
    lib = _abi.lib()
    refs = (_abi.GsrViewRef * 17)()
    inp = _abi.GsrInputs()
    inp.P = 1
    assert lib.gsr_backward_preprocess_views(C.byref(inp), 17, refs, 1, *([None] * 7), None) == _abi.GSR_ERR_INVALID
    assert "views per call" in _abi.last_error()


@pytest.mark.parametrize("fuse", [1, 2, 8])
def test_view_pipeline_fused_and_unfused_agree(fuse):
    """ViewPipeline with the per-Gaussian backward deferred over groups of `fuse` cameras against fuse = 1."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("small")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, 7, W=320, H=240)
    rng = np.random.default_rng(21)
    dLs = [torch.from_numpy(rng.standard_normal((3, 240, 320)).astype(np.float32)).cuda() for _ in settings]
    states = _forward_all(settings, t, shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    ref, _ = _per_view_reference(states, dLs, t)
    vp = views.ViewPipeline(fuse_backward=fuse)
    assert len(vp.slots) >= max(3, fuse + 2 if fuse > 1 else 3)
    for rep in range(3):  # slot reuse across steps
        over = rep == 2  # overwrite mode: the buffers start full of garbage and are SET to the step's sum
        got = {k: (torch.full_like(v, 3.0) if over else torch.zeros_like(v)) for k, v in t.items()}
        skip = {2} if rep == 1 else set()  # a camera without a backward inside a group

        def dL_fn(img, k):
            return None if k in skip else dLs[k]
        vp.step(settings, dL_fn, got, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
                rotations=t["rotations"], overwrite=over)
        torch.cuda.synchronize()
        want = ref
        if skip:
            keep = [i for i in range(len(states)) if i not in skip]
            want, _ = _per_view_reference([states[i] for i in keep], [dLs[i] for i in keep], t)
        for k in want:
            ok, worst = h.grad_close(got[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
            assert ok, (fuse, rep, k, worst)
    # overwrite with no backward at all: the sum over no cameras is zero
    got = {k: torch.full_like(v, 3.0) for k, v in t.items()}
    vp.step(settings[:2], lambda img, k: None, got, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
            rotations=t["rotations"], overwrite=True)
    assert all(float(v.abs().max()) == 0.0 for v in got.values())
    # a summed screen-space gradient buffer (grads["means2D"]) keeps its meaning with the deferred pass
    g1 = {k: torch.zeros_like(v) for k, v in t.items()}; g1["means2D"] = torch.zeros((s["P"], 3), device="cuda")
    vp.step(settings, lambda img, k: dLs[k], g1, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
            rotations=t["rotations"], overwrite=True)
    g0 = {k: torch.zeros_like(v) for k, v in t.items()}; g0["means2D"] = torch.zeros((s["P"], 3), device="cuda")
    for st, dL in zip(states, dLs):
        views.backward_view(st, dL, g0)
    for k in g0:
        ok, worst = h.grad_close(g1[k].cpu().numpy(), g0[k].cpu().numpy(), **TOL)
        assert ok, (fuse, "means2D run", k, worst)


def test_forward_views_is_bit_identical_to_forward_view():
    """gsr_forward_preprocess_views + gsr_forward_order against gsr_forward_geometry: the whole geometry state, radii,
    duplicate count, point list and image of every view, byte for byte; 19 views = a group of 16 and a group of 3."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("small")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, 19, W=320, H=240)
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    got = views.forward_views(settings, t["means3D"], t["opacities"], **kw)
    assert len(got) == len(settings)
    for rs, (img, radii, st) in zip(settings, got):
        ref_img, ref_radii, ref_st = views.forward_view(rs, t["means3D"], t["opacities"], **kw)
        assert st.L == ref_st.L
        assert torch.equal(radii, ref_radii)
        assert torch.equal(img, ref_img)
        assert torch.equal(st.point_list[: st.L], ref_st.point_list[: st.L])
        P, L, W, H = st.P, st.L, 320, 240
        for name, dt in (("depths", np.float32), ("means2D", np.float32), ("conic_opacity", np.float32), ("rgb", np.float32),
                         ("clamped", np.uint8), ("tiles_touched", np.uint32), ("rects", np.uint16), ("perm", np.uint32),
                         ("offsets_sorted", np.uint32), ("cov3D", np.float32), ("n_contrib", np.uint32),
                         ("final_T", np.float32), ("ranges", np.uint32)):
            a = h.gsr_fetch(name, dt, P, W, H, L, st.geom, st.image, None, st.point_list)
            b = h.gsr_fetch(name, dt, P, W, H, L, ref_st.geom, ref_st.image, None, ref_st.point_list)
            if name in ("depths", "means2D", "conic_opacity", "rgb", "clamped", "cov3D"):
                vis = (ref_radii > 0).cpu().numpy()  # rows of culled Gaussians are never written (either path)
                k = a.size // P
                a, b = a.reshape(P, k)[vis], b.reshape(P, k)[vis]
            assert np.array_equal(a.view(np.uint8), b.view(np.uint8)), name
    # gradients through the states of the grouped forward
    rng = np.random.default_rng(4)
    dLs = [torch.from_numpy(rng.standard_normal((3, 240, 320)).astype(np.float32)).cuda() for _ in settings]
    g1 = {k: torch.zeros_like(v) for k, v in t.items()}
    views.backward_views([st for _, _, st in got], dLs, g1, accumulate=False)
    states = _forward_all(settings, t, **kw)
    g0, _ = _per_view_reference(states, dLs, t)
    for k in g0:
        ok, worst = h.grad_close(g1[k].cpu().numpy(), g0[k].cpu().numpy(), **TOL)
        assert ok, (k, worst)


def test_forward_views_precomputed_inputs_and_low_sh_degree():
    from diff_gaussian_rasterization import views
    sc, kw, gold = h.golden_case("tiny_precomp")
    t = h.scene_tensors(sc, requires_grad=False)
    col = torch.from_numpy(kw["colors_precomp"]).cuda()
    cov = torch.from_numpy(kw["cov3D_precomp"]).cuda()
    settings = _views_of(sc, 3, W=sc["cam"]["image_width"], H=sc["cam"]["image_height"])
    for (img, radii, st), rs in zip(views.forward_views(settings, t["means3D"], t["opacities"], colors_precomp=col,
                                                        cov3D_precomp=cov), settings):
        ref_img, ref_radii, _ = views.forward_view(rs, t["means3D"], t["opacities"], colors_precomp=col, cov3D_precomp=cov)
        assert torch.equal(img, ref_img) and torch.equal(radii, ref_radii)
    sc, kw, gold = h.golden_case("tiny_sh1_bg")
    t = h.scene_tensors(sc, requires_grad=False)
    settings = _views_of(sc, 3, W=sc["cam"]["image_width"], H=sc["cam"]["image_height"])
    for shs in (t["shs"], t["shs"][:, :4, :].contiguous()):
        kw2 = dict(shs=shs, scales=t["scales"], rotations=t["rotations"])
        for (img, radii, st), rs in zip(views.forward_views(settings, t["means3D"], t["opacities"], **kw2), settings):
            ref_img, ref_radii, _ = views.forward_view(rs, t["means3D"], t["opacities"], **kw2)
            assert torch.equal(img, ref_img) and torch.equal(radii, ref_radii), tuple(shs.shape)


@pytest.mark.parametrize("fuse_forward", [True, False])
def test_view_pipeline_grouped_preprocess_images_bit_identical(fuse_forward):
    """ViewPipeline with the grouped preprocess: images bit-identical to forward_view, gradients equal to per-view
    sums; 37 cameras (more than the 34 slots of the ring), three steps, a camera without a backward."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("small")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, 37, W=320, H=240)
    rng = np.random.default_rng(23)
    dLs = [torch.from_numpy(rng.standard_normal((3, 240, 320)).astype(np.float32)).cuda() for _ in settings]
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    ref_imgs, states = [], []
    for rs in settings:
        img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], **kw)
        ref_imgs.append(img); states.append(st)
    vp = views.ViewPipeline(fuse_forward=fuse_forward)
    for rep in range(3):
        skip = {5, 6} if rep == 1 else set()
        keep = [i for i in range(len(settings)) if i not in skip]
        want, _ = _per_view_reference([states[i] for i in keep], [dLs[i] for i in keep], t)
        got = {k: torch.full_like(v, 2.0) for k, v in t.items()}
        imgs = []

        def dL_fn(img, k):
            imgs.append(img.clone())
            return None if k in skip else dLs[k]
        counts = vp.step(settings, dL_fn, got, t["means3D"], t["opacities"], overwrite=True, **kw)
        torch.cuda.synchronize()
        assert counts == [st.L for st in states]
        for a, b in zip(imgs, ref_imgs):
            assert torch.equal(a, b)
        for k in want:
            ok, worst = h.grad_close(got[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
            assert ok, (fuse_forward, rep, k, worst)


@pytest.mark.parametrize("P,W,H", [(1, 16, 16), (3, 17, 33), (50, 1, 40), (300, 161, 7)])
def test_grouped_paths_on_extreme_and_ragged_sizes(P, W, H):
    """One Gaussian, images smaller than a tile, widths / heights that are not multiples of 16, and views of
    DIFFERENT sizes inside one group: forward_views / backward_views / ViewPipeline against the per-view calls."""
    from diff_gaussian_rasterization import views
    from gsr_synth import ring_cameras
    s = h.make_scene("tiny", P=P, width=W, height=H, seed=21)
    s["means3D"] *= 0.3  # keep them in view
    t = h.scene_tensors(s, requires_grad=False)
    sizes = [(W, H), (W + 16, H), (W, H + 5)]
    settings = []
    for i, (w, hh) in enumerate(sizes):
        sc = dict(s); sc["cam"] = ring_cameras(3, width=w, height=hh)[i]
        settings.append(h.settings_for(sc))
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    rng = np.random.default_rng(P + W)
    dLs = [torch.from_numpy(rng.standard_normal((3, rs.image_height, rs.image_width)).astype(np.float32)).cuda()
           for rs in settings]
    ref_imgs, states = [], []
    for rs in settings:
        img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], **kw)
        ref_imgs.append(img); states.append(st)
    want, _ = _per_view_reference(states, dLs, t)
    got_fwd = views.forward_views(settings, t["means3D"], t["opacities"], **kw)
    for (img, radii, st), ref_img, ref_st in zip(got_fwd, ref_imgs, states):
        assert torch.equal(img, ref_img) and torch.equal(radii, ref_st.radii) and st.L == ref_st.L
    g = {k: torch.full_like(v, 4.0) for k, v in t.items()}
    views.backward_views([st for _, _, st in got_fwd], dLs, g, accumulate=False)
    for k in want:
        ok, worst = h.grad_close(g[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
        assert ok, ("backward_views", k, worst)
    vp = views.ViewPipeline()
    g = {k: torch.full_like(v, 4.0) for k, v in t.items()}
    imgs = []
    vp.step(settings, lambda img, k: (imgs.append(img.clone()), dLs[k])[1], g, t["means3D"], t["opacities"],
            overwrite=True, **kw)
    torch.cuda.synchronize()
    for a, b in zip(imgs, ref_imgs):
        assert torch.equal(a, b)
    for k in want:
        ok, worst = h.grad_close(g[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
        assert ok, ("pipeline", k, worst)


def test_view_pipeline_recovers_after_a_step_that_raised():
    """dL_fn raising in the middle of a group leaves deferred cameras behind; the next step must not apply them."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("tiny")
    t = h.scene_tensors(s, requires_grad=False)
    settings = _views_of(s, 5)
    rng = np.random.default_rng(31)
    dLs = [torch.from_numpy(rng.standard_normal((3, 120, 160)).astype(np.float32)).cuda() for _ in settings]
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    states = _forward_all(settings, t, **kw)
    want, _ = _per_view_reference(states, dLs, t)
    vp = views.ViewPipeline()
    junk = {k: torch.zeros_like(v) for k, v in t.items()}

    def bad(img, k):
        if k == 3:
            raise ValueError("loss failed")
        return dLs[k] * 1000.0
    with pytest.raises(ValueError):
        vp.step(settings, bad, junk, t["means3D"], t["opacities"], **kw)
    torch.cuda.synchronize()
    got = {k: torch.full_like(v, 6.0) for k, v in t.items()}
    vp.step(settings, lambda img, k: dLs[k], got, t["means3D"], t["opacities"], overwrite=True, **kw)
    torch.cuda.synchronize()
    for k in want:
        ok, worst = h.grad_close(got[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
        assert ok, (k, worst)


def test_views_that_differ_in_sh_degree_or_scale_modifier_are_not_grouped():
    """sh_degree and scale_modifier are per-view settings in the reference API but per-group inputs of the grouped
    kernels: mixed values must fall back to the per-view kernels (pipeline, forward_views) or be refused (backward_views)."""
    from diff_gaussian_rasterization import views
    s = h.make_scene("tiny")
    t = h.scene_tensors(s, requires_grad=False)
    base = _views_of(s, 3)
    settings = [base[0], base[1]._replace(sh_degree=1), base[2]._replace(scale_modifier=0.7)]
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    rng = np.random.default_rng(41)
    dLs = [torch.from_numpy(rng.standard_normal((3, 120, 160)).astype(np.float32)).cuda() for _ in settings]
    ref_imgs, states = [], []
    for rs in settings:
        img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], **kw)
        ref_imgs.append(img); states.append(st)
    assert not torch.equal(ref_imgs[1], views.forward_view(base[1], t["means3D"], t["opacities"], **kw)[0])
    want, _ = _per_view_reference(states, dLs, t)
    for (img, radii, st), ref in zip(views.forward_views(settings, t["means3D"], t["opacities"], **kw), ref_imgs):
        assert torch.equal(img, ref)
    with pytest.raises(Exception):
        views.backward_views(states, dLs, {k: torch.zeros_like(v) for k, v in t.items()})
    got = {k: torch.zeros_like(v) for k, v in t.items()}
    imgs = []
    views.ViewPipeline().step(settings, lambda img, k: (imgs.append(img.clone()), dLs[k])[1], got, t["means3D"],
                              t["opacities"], **kw)
    torch.cuda.synchronize()
    for a, b in zip(imgs, ref_imgs):
        assert torch.equal(a, b)
    for k in want:
        ok, worst = h.grad_close(got[k].cpu().numpy(), want[k].cpu().numpy(), **TOL)
        assert ok, (k, worst)
