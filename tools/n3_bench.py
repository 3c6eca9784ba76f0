"""GPU box: per-view fwd+bwd time of (a) render()'s way -- torch activations + torch.cat every view + autograd through
them, (b) the accumulate API on pre-activated tensors, (c) the fused raw-parameter / split-SH path (N3), on C2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import helpers as h
import diff_gaussian_rasterization as dgr
from diff_gaussian_rasterization import views
s = h.make_scene(sys.argv[1] if len(sys.argv) > 1 else "C2")
t = h.scene_tensors(s, requires_grad=False)
rs = h.settings_for(s)
dL = torch.from_numpy(s["dL_dpix"]).cuda()
raw = dict(xyz=t["means3D"].clone(), f_dc=t["shs"][:, :1].contiguous(), f_rest=t["shs"][:, 1:].contiguous(),
           opacity=torch.logit(t["opacities"].clamp(1e-4, 1 - 1e-4)), scaling=torch.log(t["scales"]), rotation=t["rotations"] * 2.5)
leaf = {k: v.clone().requires_grad_(True) for k, v in raw.items()}
m2 = torch.zeros((s["P"], 3), device="cuda", requires_grad=True)


def a():
    img, _ = dgr.GaussianRasterizer(rs)(means3D=leaf["xyz"], means2D=m2, opacities=torch.sigmoid(leaf["opacity"]),
                                        shs=torch.cat((leaf["f_dc"], leaf["f_rest"]), dim=1), scales=torch.exp(leaf["scaling"]),
                                        rotations=torch.nn.functional.normalize(leaf["rotation"]))
    torch.autograd.backward([img], [dL])


gb = {k: torch.zeros_like(v) for k, v in t.items()}


def b():
    img, _, st = views.forward_view(rs, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    views.backward_view(st, dL, gb)


gc = dict(means3D=torch.zeros_like(raw["xyz"]), opacities=torch.zeros_like(raw["opacity"]), shs=torch.zeros_like(raw["f_dc"]),
          shs_rest=torch.zeros_like(raw["f_rest"]), scales=torch.zeros_like(raw["scaling"]), rotations=torch.zeros_like(raw["rotation"]))


def c():
    img, _, st = views.forward_view(rs, raw["xyz"], raw["opacity"], shs=raw["f_dc"], shs_rest=raw["f_rest"], scales=raw["scaling"],
                                    rotations=raw["rotation"], raw_params=True)
    views.backward_view(st, dL, gc)


for name, fn in (("render()-style: torch activations + cat + autograd, per view", a), ("accumulate API, pre-activated inputs", b),
                 ("fused raw parameters + split SH (N3)", c)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 10:.3f} ms / view")
