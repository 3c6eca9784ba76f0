import sys, os
ROOT="/root/repo"
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")): sys.path.insert(0, p)
import numpy as np, torch, helpers as h
for name in sys.argv[1:]:
    s = h.make_scene(name)
    import diff_gaussian_rasterization as dgr
    rs = h.settings_for(s); t = h.scene_tensors(s)
    m2 = torch.zeros((s["P"],3), device="cuda", requires_grad=True)
    out = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=m2, opacities=t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"])
    fn = out[0].grad_fn
    geom, pl, img = fn.saved_tensors[9], fn.saved_tensors[10], fn.saved_tensors[11]
    W,H = s["cam"]["image_width"], s["cam"]["image_height"]
    tt = h.gsr_fetch("tiles_touched", np.uint32, s["P"], W, H, fn.num_rendered, geom, img, dgr._state(torch.device("cuda",0)).scratch, pl)
    tt = tt[tt>0]
    bins=[1,2,3,5,7,10,17,33,65,129,1025,10**9]
    hist,_=np.histogram(tt,bins=bins)
    pairs=[int(tt[(tt>=bins[i])&(tt<bins[i+1])].sum()) for i in range(len(bins)-1)]
    print(name, "visible", tt.size, "L", int(tt.sum()))
    for i in range(len(bins)-1): print(f"  area [{bins[i]},{bins[i+1]}): gaussians {hist[i]:9d} pairs {pairs[i]:11d} ({100*pairs[i]/tt.sum():.1f}%)")
