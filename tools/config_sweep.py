"""GPU box: BASELINE configs C3 / C4 / C5 at full size -- parity against the live reference CUDA (oracle/_ref) and
per-view timing of both.  python tools/config_sweep.py [C3 C4 C5] -> one JSON line per config (also appended to
gpurun_out/config_sweep.jsonl)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3d-gaussian-splat-attack_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch
import helpers as h
from diff_gaussian_rasterization import views, _abi

FWD_ONLY = {"C4"}


def time_product(s, bwd, iters=3):
    t = h.scene_tensors(s, requires_grad=False)
    rs = h.settings_for(s)
    dL = torch.from_numpy(s["dL_dpix"]).cuda()
    grads = {k: torch.zeros_like(v) for k, v in t.items()}
    lib = _abi.lib()

    def one():
        img, radii, st = views.forward_view(rs, t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
                                            rotations=t["rotations"])
        if bwd:
            views.backward_view(st, dL, grads)
        return st

    one()
    torch.cuda.synchronize()
    import ctypes as C
    lib.gsr_profile_enable(1)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        st = one()
    b.record()
    torch.cuda.synchronize()
    ms = (C.c_float * len(_abi.STAGES))()
    calls = (C.c_int32 * len(_abi.STAGES))()
    lib.gsr_profile_read(ms, calls)
    lib.gsr_profile_enable(0)
    stages = {n: round(float(ms[i]) / iters, 4) for i, n in enumerate(_abi.STAGES)}
    return a.elapsed_time(b) / iters, stages, st.L, int((st.radii > 0).sum().item())


def time_pipeline(s, bwd, n_views=8, iters=3):
    """The same scene through views.ViewPipeline: `n_views` ring cameras per step, per-Gaussian work grouped over the
    cameras, three streams (forward-only when bwd is False).  Returns ms per view."""
    from gsr_synth import ring_cameras
    t = h.scene_tensors(s, requires_grad=False)
    W, H = s["cam"]["image_width"], s["cam"]["image_height"]
    settings = []
    for c in ring_cameras(n_views, width=W, height=H):
        sc = dict(s); sc["cam"] = c
        settings.append(h.settings_for(sc))
    dL = torch.from_numpy(s["dL_dpix"]).cuda()
    grads = {k: torch.zeros_like(v) for k, v in t.items()}
    vp = views.ViewPipeline()
    fn = (lambda img, k: dL) if bwd else (lambda img, k: None)
    kw = dict(shs=t["shs"], scales=t["scales"], rotations=t["rotations"], overwrite=bwd)
    vp.step(settings, fn, grads, t["means3D"], t["opacities"], **kw)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        vp.step(settings, fn, grads, t["means3D"], t["opacities"], **kw)
    b.record()
    torch.cuda.synchronize()
    del vp
    return a.elapsed_time(b) / iters / n_views


def time_ref(s, bwd, iters=2):
    h.run_ref(s, grad=bwd, fetch=False)
    t0 = time.perf_counter()
    for _ in range(iters):
        h.run_ref(s, grad=bwd, fetch=False)
    return None  # run_ref includes host copies; the bench's ref_cuda_baseline is the timed comparison


def main():
    names = sys.argv[1:] or ["C3", "C4", "C5"]
    for name in names:
        s = h.make_scene(name)
        bwd = name not in FWD_ONLY
        out = {"config": name, "P": s["P"], "W": s["cam"]["image_width"], "H": s["cam"]["image_height"], "backward": bwd}
        ms, stages, L, V = time_product(s, bwd)
        out.update(ms_per_view=round(ms, 3), views_per_s=round(1e3 / ms, 1), stages_ms=stages, L=int(L), V=V)
        torch.cuda.empty_cache()
        if name != "C5":  # (C5: 1.5e9 pairs per view -- one view's lists at a time)
            mp = time_pipeline(s, bwd)
            out.update(pipeline_ms_per_view=round(mp, 3), pipeline_views_per_s=round(1e3 / mp, 1),
                       pipeline="views.ViewPipeline, 8 ring cameras per step, grouped per-Gaussian kernels")
            torch.cuda.empty_cache()
        g = h.run_gsr(s, grad=bwd, fetch=False)
        assert np.isfinite(g["color"]).all()
        if h.ref_available():
            r = h.run_ref(s, grad=bwd, fetch=False)
            out["num_rendered_equal"] = bool(r["num_rendered"] == L)
            out["radii_equal"] = bool(np.array_equal(g["radii"], r["radii"]))
            out["image_bit_identical"] = bool(np.array_equal(g["color"].view(np.uint32), r["color"].view(np.uint32)))
            out["image_max_abs_diff"] = float(np.abs(g["color"] - r["color"]).max())
            if bwd:
                ratios = {}
                for k in ("means3D", "means2D", "opacity", "sh", "scales", "rotations"):
                    a, b = g["grads"][k].reshape(-1), r["grads"][k].reshape(-1)
                    ratios[k] = round(h.grad_close(a, b[: a.size])[1], 4)
                out["grad_tolerance_ratio"] = ratios
            del r
        del g
        torch.cuda.empty_cache()
        line = json.dumps(out)
        print(line, flush=True)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "config_sweep.jsonl"), "a") as f:
            f.write(line + "\n")


if __name__ == "__main__":
    main()
