"""Conformance: the reference's UNMODIFIED gaussian_renderer.render() (gaussian_renderer/__init__.py:18-103),
handed a reference GaussianModel and reference Camera objects, runs on this repo's diff_gaussian_rasterization
exactly the way attack.py:481-494 drives it -- render() per camera (it always passes sh_objs = pc.get_objects),
a loss on render_pkg["render"] only, one backward -- and its image / gradients are compared with the reference
CUDA rasterizer (oracle/_ref/libdgr_ref.so) fed the same activated tensors and camera matrices.

The reference modules are the byte-compiled files of oracle/_ref/refpy (tests/ref_harness.py); nothing under
/root/reference is read at run time.
"""
import numpy as np
import pytest
import torch

import helpers as h
import ref_harness as rh

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rh.available(), reason="oracle/_ref/refpy not built (make -C oracle refpy)")]


def _activated_scene(scene, gm, cam, bg):
    """The numpy `scene` dict h.run_ref / h.run_oracle take, built from what render() will hand the rasterizer."""
    n = lambda t: np.ascontiguousarray(t.detach().float().cpu().numpy())
    s = dict(scene)
    s.update(means3D=n(gm.get_xyz), scales=n(gm.get_scaling), rotations=n(gm.get_rotation), opacities=n(gm.get_opacity),
             shs=n(gm.get_features), bg=n(bg)[:3], cam=rh.camera_arrays(cam))
    return s


def _raw_grads_via_reference(gm, ref_grads):
    """Reference-CUDA gradients w.r.t. the activated attributes, pulled back through the reference model's own
    activations (scene/gaussian_model.py:97-121) by torch autograd -> gradients of the raw parameters."""
    params = [gm._xyz, gm._features_dc, gm._features_rest, gm._scaling, gm._rotation, gm._opacity]
    for p in params:
        p.grad = None
    outs = [gm.get_xyz, gm.get_features, gm.get_scaling, gm.get_rotation, gm.get_opacity]
    gts = [torch.from_numpy(ref_grads[k]).cuda().reshape(o.shape)
           for k, o in zip(("means3D", "sh", "scales", "rotations", "opacity"), outs)]
    torch.autograd.backward(outs, gts)
    res = {k: p.grad.detach().cpu().numpy() for k, p in
           zip(("xyz", "f_dc", "f_rest", "scaling", "rotation", "opacity"), params)}
    for p in params:
        p.grad = None
    return res


def _model_grads(gm):
    return {k: getattr(gm, a).grad.detach().cpu().numpy() for k, a in
            (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("scaling", "_scaling"),
             ("rotation", "_rotation"), ("opacity", "_opacity"))}


@pytest.mark.parametrize("size", ["tiny", "small"])
def test_unmodified_render_matches_reference_cuda(size):
    ns = rh.load()
    scene = h.make_scene(size)
    gm = rh.inverse_activated_model(ns, scene)
    cam = rh.camera_from(ns, scene["cam"])
    assert not cam.world_view_transform.is_contiguous()  # scene/cameras.py:54: a transposed view
    bg = torch.tensor([0.0, 0.0, 0.0, 0.0], device="cuda")  # attack.py:396
    dL = torch.from_numpy(scene["dL_dpix"]).cuda()

    pkg = ns.render(cam, gm, rh.pipe(), bg)
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii", "render_object"}
    H, W = scene["cam"]["image_height"], scene["cam"]["image_width"]
    assert pkg["render"].shape == (3, H, W) and pkg["render_object"].shape == (16, H, W)
    (pkg["render"] * dL).sum().backward()  # the loss reads "render" only (attack.py:485-494)
    got = _model_grads(gm)
    assert gm._objects_dc.grad is None  # nothing flowed into render_object: the K = 0 backward ran
    g_m2d = pkg["viewspace_points"].grad.detach().cpu().numpy()

    s_act = _activated_scene(scene, gm, cam, bg)
    assert h.ref_available(), "oracle/_ref/libdgr_ref.so missing"
    ref = h.run_ref(s_act, dL=scene["dL_dpix"], fetch=False)
    assert np.array_equal(pkg["render"].detach().cpu().numpy(), ref["color"])  # bit-identical image
    assert np.array_equal(pkg["radii"].cpu().numpy(), ref["radii"])
    assert np.array_equal(pkg["visibility_filter"].cpu().numpy(), ref["radii"] > 0)
    ok, worst = h.grad_close(g_m2d, ref["grads"]["means2D"])
    assert ok, ("viewspace_points", worst)
    want = _raw_grads_via_reference(gm, ref["grads"])
    for k in want:
        ok, worst = h.grad_close(got[k], want[k])
        assert ok, (k, worst)
    # render_object: Gaussian-Grouping's channel (parity unpinned) against the oracle's restatement
    o = h.run_oracle(s_act, grad=False, obj=gm.get_objects.detach().cpu().numpy().reshape(scene["P"], 16))
    assert np.abs(pkg["render_object"].detach().cpu().numpy() - o["objects"]).max() <= 1e-5


def test_attack_style_batch_of_cameras_then_one_backward():
    """attack.py:478-486: render every camera of the batch, stack, one loss, one backward."""
    ns = rh.load()
    scene = h.make_scene("tiny")
    gm = rh.inverse_activated_model(ns, scene)
    eyes = [(0.0, 0.0, -3.0), (2.0, 0.5, -2.2), (-1.5, -0.4, -2.6)]
    cams = [rh.camera_from(ns, scene["cam"], eye=e, uid=i) for i, e in enumerate(eyes)]
    bg = torch.zeros(4, device="cuda")
    dL = torch.from_numpy(scene["dL_dpix"]).cuda()
    renders = torch.stack([ns.render(c, gm, rh.pipe(), bg)["render"] for c in cams])
    (renders * dL).sum().backward()
    got = _model_grads(gm)
    tot = None
    for c in cams:
        ref = h.run_ref(_activated_scene(scene, gm, c, bg), dL=scene["dL_dpix"], fetch=False)
        w = _raw_grads_via_reference(gm, ref["grads"])
        tot = w if tot is None else {k: tot[k] + w[k] for k in w}
    for k in tot:
        ok, worst = h.grad_close(got[k], tot[k])
        assert ok, (k, worst)


def test_render_with_python_covariance_and_object_loss():
    """pipe.compute_cov3D_python (gaussian_renderer/__init__.py:62-63) feeds cov3D_precomp; a loss that DOES read
    render_object sends a gradient into _objects_dc (the K = 16 backward)."""
    ns = rh.load()
    scene = h.make_scene("tiny")
    gm = rh.inverse_activated_model(ns, scene)
    cam = rh.camera_from(ns, scene["cam"])
    bg = torch.zeros(3, device="cuda")
    dL = torch.from_numpy(scene["dL_dpix"]).cuda()
    pkg = ns.render(cam, gm, rh.pipe(compute_cov3D_python=True), bg)
    s_act = _activated_scene(scene, gm, cam, bg)
    cov = gm.get_covariance(1.0).detach().cpu().numpy()
    ref = h.run_ref(s_act, dL=scene["dL_dpix"], cov3D_precomp=cov, fetch=False)
    assert np.array_equal(pkg["render"].detach().cpu().numpy(), ref["color"])
    rng = np.random.default_rng(9)
    dLo = (rng.standard_normal(tuple(pkg["render_object"].shape)) / pkg["render_object"].numel()).astype(np.float32)
    ((pkg["render"] * dL).sum() + (pkg["render_object"] * torch.from_numpy(dLo).cuda()).sum()).backward()
    assert gm._objects_dc.grad is not None and float(gm._objects_dc.grad.abs().max()) > 0
    s_o = dict(s_act); s_o["dL_dobj"] = dLo
    o = h.run_oracle(s_o, cov3D_precomp=cov, obj=gm.get_objects.detach().cpu().numpy().reshape(scene["P"], 16))
    ok, worst = h.grad_close(gm._objects_dc.grad.cpu().numpy().reshape(-1, 16), o["grads"]["obj"])
    assert ok, worst
    ok, worst = h.grad_close(gm._features_dc.grad.cpu().numpy(), o["grads"]["sh"][:, :1])
    assert ok, worst


def test_reference_render_python_sh_branch_is_unreachable():
    """With convert_SHs_python (or override_color) the reference's own render() never binds `sh_objs`
    (gaussian_renderer/__init__.py:71-83) and raises before reaching the rasterizer; the colors_precomp path is
    therefore covered through GaussianRasterizer directly, with the reference's eval_sh producing the colours."""
    import diff_gaussian_rasterization as dgr
    ns = rh.load()
    scene = h.make_scene("tiny")
    gm = rh.inverse_activated_model(ns, scene)
    cam = rh.camera_from(ns, scene["cam"])
    bg = torch.zeros(3, device="cuda")
    with pytest.raises(UnboundLocalError):
        ns.render(cam, gm, rh.pipe(convert_SHs_python=True), bg)
    with torch.no_grad():  # gaussian_renderer/__init__.py:73-77
        shs_view = gm.get_features.transpose(1, 2).view(-1, 3, (gm.max_sh_degree + 1) ** 2)
        dir_pp = gm.get_xyz - cam.camera_center.repeat(gm.get_features.shape[0], 1)
        sh2rgb = ns.eval_sh(gm.active_sh_degree, shs_view, dir_pp / dir_pp.norm(dim=1, keepdim=True))
        colors = torch.clamp_min(sh2rgb + 0.5, 0.0)
    s_act = _activated_scene(scene, gm, cam, bg)
    rs = h.settings_for(s_act)
    rs = rs._replace(viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform, campos=cam.camera_center)
    img, radii = dgr.GaussianRasterizer(rs)(means3D=gm.get_xyz, means2D=torch.zeros_like(gm.get_xyz),
                                            opacities=gm.get_opacity, colors_precomp=colors,
                                            scales=gm.get_scaling, rotations=gm.get_rotation)
    ref = h.run_ref(s_act, grad=False, colors_precomp=colors.cpu().numpy(), fetch=False)
    assert np.array_equal(img.detach().cpu().numpy(), ref["color"]) and np.array_equal(radii.cpu().numpy(), ref["radii"])
