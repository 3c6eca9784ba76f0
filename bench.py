#!/usr/bin/env python
"""bench.py -- fwd+bwd views/sec of the B200-native Gaussian-splat rasterizer (BASELINE.json metric).

A "step" is one DAGGER-style PGD gradient step on one rank: `--views` camera views (default 8,
config C3's 64 views / 8 GPUs) are rendered forward and backward against a replicated Gaussian
set (config C2: 1M Gaussians, SH degree 3, 1920x1080), attribute gradients accumulate over the
views and -- for N > 1 -- are summed with one NCCL all-reduce.  views/sec = N * views * K / t.

  value     inputs resident in HBM, public Python API: views.ViewPipeline.step (each view's gradients are added
            straight into the flat all-reduce bucket; preprocess / sorts / binning of the next camera run on a
            high-priority stream underneath the blending kernels of the current one).  `--no-overlap` keeps one
            stream (views.forward_view + backward_view); `--api autograd` times the drop-in GaussianRasterizer +
            autograd backward instead.  The `stages` table is always measured on one stream.
  e2e       same step driven from HOST buffers through pipeline.HostStepPipeline: Gaussian attributes,
            camera matrices and dL/dpixel copied from pinned host memory every step, gradients read back
            to the host every step; copies of neighbouring steps overlap compute (depth-2 pipeline, three
            streams); `e2e.serial` is the same step with nothing overlapped
  roofline  dominant kernel, timed live with CUDA events on the launching stream (gsr_profile_*)
  cpu_baseline / --impl reference: the CPU oracle (oracle/cpu_ref.cpp, OpenMP, all host cores) on
            a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "3d-gaussian-splat-attack_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "fwd+bwd views/sec at 1080p, 1M Gaussians"  # the default workload (C2); other --workload values are named below


def metric_name(P, W, H):
    if (W, H) == (1920, 1080) and P == 1_000_000:
        return METRIC
    return f"fwd+bwd views/sec at {W}x{H}, {P / 1e6:g}M Gaussians"
UNIT = "views/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gsr", choices=["gsr", "reference"])
    ap.add_argument("--workload", default="C2", help="scene config of gsr_synth (C2 = BASELINE configs[1])")
    ap.add_argument("--views", type=int, default=8, help="camera views per rank per step")
    ap.add_argument("--P", type=int, default=None, help="override Gaussian count (debug)")
    ap.add_argument("--api", default="views", choices=["views", "autograd"],
                    help="views: forward_view/backward_view accumulate API; autograd: drop-in GaussianRasterizer")
    ap.add_argument("--sh-objs", action="store_true",
                    help="with --api autograd: pass sh_objs (P,1,16) and build the screen-space tensor per call, exactly as "
                         "the reference's gaussian_renderer.render() does (gaussian_renderer/__init__.py:26,81-95)")
    ap.add_argument("--no-overlap", action="store_true",
                    help="views API without the two-stream ViewPipeline (everything on one stream)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-drop-in", action="store_true", help="skip the drop-in autograd API measurements")
    ap.add_argument("--no-c3", action="store_true",
                    help="skip the extra C3 leg (BASELINE configs[2]: 3M Gaussians, 64 views per step over all ranks)")
    ap.add_argument("--cpu-sample-views", type=int, default=1)
    ap.add_argument("--stages-out", default=None, help="write the per-stage table as JSON to this file")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback", 1965.0


def cameras_for(scene, n_total):
    from gsr_synth import ring_cameras
    c = scene["cam"]
    return ring_cameras(n_total, radius=3.0, width=c["image_width"], height=c["image_height"])


# ------------------------------------------------------------------ CPU arm (oracle = checker, timed as baseline)
def cpu_views_per_sec(scene, cams, steps, warmup, views):
    """Times the CPU oracle (the reference's rasterizer math restated for the host, oracle/cpu_ref.cpp;
    the reference has no CPU implementation of its own) on `views` views per step."""
    from oracle.cpu_ref import Oracle, num_threads, set_num_threads
    # torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arm is meant to use every host core
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    set_num_threads(avail)
    cores = num_threads()
    dL = scene["dL_dpix"]

    def one(cam):
        o = Oracle()
        o.forward(scene["bg"], scene["means3D"], scene["opacities"], cam["viewmatrix"], cam["projmatrix"],
                  cam["campos"], cam["image_width"], cam["image_height"], cam["tanfovx"], cam["tanfovy"],
                  sh_degree=scene["sh_degree"], shs=scene["shs"], scales=scene["scales"], rotations=scene["rotations"])
        o.backward(dL)

    for i in range(warmup):
        one(cams[i % len(cams)])
    t0 = time.perf_counter()
    n = 0
    for s in range(steps):
        for v in range(views):
            one(cams[(s * views + v) % len(cams)])
            n += 1
    dt = time.perf_counter() - t0
    return n / dt, dt, cores, n


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gsr_synth import make_scene
    scene = make_scene(args.workload, P=args.P)
    cams = cameras_for(scene, max(args.views * args.gpus, 1))
    steps, warmup = max(1, min(args.steps, 3)), min(args.warmup, 1)
    vps, dt, cores, n = cpu_views_per_sec(scene, cams, steps, warmup, args.cpu_sample_views)
    c = scene["cam"]
    sample = f"{n} full view(s) fwd+bwd of {args.workload} (P={scene['P']}, {c['image_width']}x{c['image_height']}), {warmup} warm-up"
    line = {
        "impl": "reference", "metric": metric_name(scene["P"], c["image_width"], c["image_height"]), "value": vps, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": 1e3 * dt / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {scene['P']} Gaussians SH3, {c['image_width']}x{c['image_height']}, fwd+bwd",
                   "views_per_step": args.cpu_sample_views, "note": "reference has no CPU rasterizer; this is the CPU "
                   "restatement of its CUDA math (oracle/cpu_ref.cpp) on all host cores.  A full C2 view takes ~2.3 s "
                   "of CPU time, so this arm is bounded to min(--steps, 3) steps of --cpu-sample-views view(s) and at "
                   "most 1 warm-up (steps / warmup in this line are the ones actually run)"},
        "cpu_baseline": {"value": vps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": vps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ GPU arm
ATTRS = (("means3D", 3), ("shs", 48), ("scales", 3), ("rotations", 4), ("opacities", 1))  # floats per Gaussian
SHAPES = {"means3D": (-1, 3), "shs": (-1, 16, 3), "scales": (-1, 3), "rotations": (-1, 4), "opacities": (-1, 1)}


class Workload:
    """Replicated Gaussian set in one flat HBM buffer (+ flat gradient buffer = the all-reduce bucket),
    this rank's cameras and dL/dpixel images; pinned host mirrors for the e2e path."""

    def __init__(self, args, torch, rank, world, dev):
        from gsr_synth import make_scene
        import diff_gaussian_rasterization as dgr
        self.torch, self.dgr, self.dev, self.rank, self.world = torch, dgr, dev, rank, world
        sc = self.scene = make_scene(args.workload, P=args.P)
        self.P, self.W, self.H = sc["P"], sc["cam"]["image_width"], sc["cam"]["image_height"]
        self.cams = cameras_for(sc, args.views * world)[rank::world]
        from diff_gaussian_rasterization.dist import FlatGradBucket, flat_layout
        slices, n_flat = flat_layout({n: self.P * k for n, k in ATTRS})  # every slice 256-byte aligned for any P
        host = np.zeros(n_flat, np.float32)
        for n, _ in ATTRS:
            host[slices[n][0]:slices[n][1]] = sc[n].reshape(-1)
        self.host_param = torch.from_numpy(host).pin_memory()
        self.host_grad = torch.empty_like(self.host_param).pin_memory()
        self.flat_param = self.host_param.to(dev)
        self.params = {}
        for n, k in ATTRS:
            a, b = slices[n]
            self.params[n] = self.flat_param[a:b].view(SHAPES[n]).detach().requires_grad_(True)
        self.bucket = FlatGradBucket(self.params)  # every .grad is a view into bucket.flat (the NCCL bucket)
        self.flat_grad = self.bucket.flat
        self.means2D = torch.zeros((self.P, 3), device=dev)  # dummy screen-space tensor (no grad wanted)
        rng = np.random.default_rng(99 + rank)
        nv = len(self.cams)
        g = (rng.standard_normal((nv, 3, self.H, self.W), dtype=np.float32) / (3.0 * self.H * self.W))
        self.host_dL = torch.from_numpy(g).pin_memory()
        self.dev_dL = self.host_dL.to(dev)
        cam_host = np.stack([np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1), c["campos"],
                                             np.zeros(1, np.float32)]) for c in self.cams]).astype(np.float32)
        self.host_cam = torch.from_numpy(cam_host).pin_memory()       # (nv, 36)
        self.dev_cam = self.host_cam.to(dev)
        self.bg = torch.zeros(3, device=dev)
        self.api = args.api
        self.sh_objs = bool(getattr(args, "sh_objs", False))
        self.obj = None
        self.grads = {n: self.bucket.grad(n) for n, _ in ATTRS}
        self.overlap = not args.no_overlap
        self.vp = None
        self.settings_list = None
        self.host_grad2 = torch.empty_like(self.host_param).pin_memory()
        self.pipe = None
        self.h2d_bytes = self.host_param.numel() * 4 + self.host_dL.numel() * 4 + self.host_cam.numel() * 4
        self.d2h_bytes = self.host_grad.numel() * 4
        self.last_fn = None

    def settings(self, v):
        c, cm = self.cams[v], self.dev_cam[v]
        return self.dgr.GaussianRasterizationSettings(
            image_height=self.H, image_width=self.W, tanfovx=c["tanfovx"], tanfovy=c["tanfovy"], bg=self.bg,
            scale_modifier=1.0, viewmatrix=cm[0:16].view(4, 4), projmatrix=cm[16:32].view(4, 4),
            sh_degree=self.scene["sh_degree"], campos=cm[32:35], prefiltered=False, debug=False)

    def view_fwd_bwd(self, v):
        p = self.params
        if self.api == "views":
            from diff_gaussian_rasterization import views
            color, radii, st = views.forward_view(self.settings(v), p["means3D"], p["opacities"], shs=p["shs"],
                                                  scales=p["scales"], rotations=p["rotations"])
            self.last_state, self.last_radii = st, radii
            views.backward_view(st, self.dev_dL[v], self.grads)
            return
        if self.sh_objs:
            self.view_render_like(v)
            return
        color, radii = self.dgr.GaussianRasterizer(self.settings(v))(
            means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"],
            scales=p["scales"], rotations=p["rotations"])
        self.last_fn, self.last_radii = color.grad_fn, radii
        self.torch.autograd.backward([color], [self.dev_dL[v]])

    def view_render_like(self, v):
        """What the reference's render() does around the rasterizer call (gaussian_renderer/__init__.py:26-103) on
        pre-activated attributes: a fresh screen-space tensor with retain_grad, sh_objs always passed, the
        visibility filter; the loss reads the colour image only (attack.py:485-494)."""
        t, p = self.torch, self.params
        if self.obj is None:
            g = t.Generator(device=self.dev).manual_seed(7)
            self.obj = t.rand((self.P, 1, 16), device=self.dev, generator=g).requires_grad_(True)
        ssp = t.zeros_like(p["means3D"], requires_grad=True) + 0
        ssp.retain_grad()
        color, radii, objects = self.dgr.GaussianRasterizer(self.settings(v))(
            means3D=p["means3D"], means2D=ssp, opacities=p["opacities"], shs=p["shs"], sh_objs=self.obj,
            scales=p["scales"], rotations=p["rotations"])
        pkg = {"render": color, "viewspace_points": ssp, "visibility_filter": radii > 0, "radii": radii,
               "render_object": objects}
        self.last_fn, self.last_radii = color.grad_fn, radii
        t.autograd.backward([pkg["render"]], [self.dev_dL[v]])

    def view_fwd_only(self, v):
        p = self.params
        color, radii = self.dgr.GaussianRasterizer(self.settings(v))(
            means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"],
            scales=p["scales"], rotations=p["rotations"])
        self.last_fn, self.last_radii, self.last_color = color.grad_fn, radii, color

    def step(self, dist=None, overlap=None):
        overlap = self.overlap if overlap is None else overlap
        if not (self.api == "views" and overlap):
            self.flat_grad.zero_()  # (the pipelined step below overwrites the gradient slices itself)
        if self.api == "views" and overlap:
            from diff_gaussian_rasterization import views
            if self.vp is None:
                self.vp = views.ViewPipeline(self.dev)
                self.settings_list = [self.settings(v) for v in range(len(self.cams))]
            p = self.params
            self.vp.step(self.settings_list, lambda img, k: self.dev_dL[k], self.grads, p["means3D"], p["opacities"],
                         shs=p["shs"], scales=p["scales"], rotations=p["rotations"], overwrite=True)
        else:
            for v in range(len(self.cams)):
                self.view_fwd_bwd(v)
        if dist is not None:
            dist.all_reduce(self.flat_grad)

    def pipeline(self):
        if self.pipe is None:
            from diff_gaussian_rasterization.pipeline import HostStepPipeline
            c = self.cams[0]
            self.pipe = HostStepPipeline(self.P, 16, self.W, self.H, len(self.cams), c["tanfovx"], c["tanfovy"],
                                         self.scene["sh_degree"], self.dev, bg=self.bg, depth=2)
            self.e2e_i = 0
            self.e2e_read = 0.0
        return self.pipe

    def step_e2e_pipelined(self, dist=None):
        """One step through HostStepPipeline; the host reads every gradient vector that completed."""
        pipe = self.pipeline()
        out = self.host_grad if self.e2e_i % 2 == 0 else self.host_grad2
        self.e2e_i += 1
        for g in pipe.submit(self.host_param, self.host_cam, self.host_dL, out):
            self.e2e_read += float(g[pipe.grad_slice[0]])  # this rank's slice of the summed gradient is on the host

    def finish_e2e(self):
        for g in self.pipe.flush():
            self.e2e_read += float(g[self.pipe.grad_slice[0]])

    def step_e2e(self, dist=None):
        t = self.torch
        self.flat_param.copy_(self.host_param, non_blocking=True)
        self.dev_cam.copy_(self.host_cam, non_blocking=True)
        self.flat_grad.zero_()
        for v in range(len(self.cams)):
            self.dev_dL[v].copy_(self.host_dL[v], non_blocking=True)
            self.view_fwd_bwd(v)
        if dist is not None:
            dist.all_reduce(self.flat_grad)
        self.host_grad.copy_(self.flat_grad, non_blocking=True)
        t.cuda.current_stream().synchronize()
        return float(self.host_grad[0])


def timed(torch, dist, fn, steps, warmup, dev, finish=None):
    """K steps between barrier + synchronize, CUDA events on the current stream, max over ranks (ms).
    `finish` (drains a pipeline) runs inside the timed region after the last step and after warm-up."""
    for _ in range(warmup):
        fn()
    if finish is not None and warmup:
        finish()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    if finish is not None:
        finish()
    b.record()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    ms = torch.tensor([a.elapsed_time(b)], device=dev)
    if dist is not None:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def fetch_u32(torch, wl, name):
    """Internal array of the last forward (gsr_debug_fetch), as a uint32 numpy array."""
    import ctypes as C
    from diff_gaussian_rasterization import _abi
    lib, fn = _abi.lib(), wl.last_fn
    geom, image = fn.saved_tensors[9], fn.saved_tensors[11]
    pv = lambda t: C.c_void_p(t.data_ptr())
    n = lib.gsr_debug_fetch(name.encode(), wl.P, wl.W, wl.H, fn.num_rendered, pv(geom), pv(image), None, None, None, 0, None)
    buf = torch.empty(int(n), dtype=torch.uint8, device=wl.dev)
    lib.gsr_debug_fetch(name.encode(), wl.P, wl.W, wl.H, fn.num_rendered, pv(geom), pv(image), None, None, pv(buf), n, None)
    torch.cuda.synchronize()
    return buf.cpu().numpy().view(np.uint32)


def fused_backward_leg(torch, wl, hbm_peak, reps=10):
    """The pipelined step (`value`) does the per-Gaussian backward once per group of views
    (gsr_backward_preprocess_views) instead of once per view: time that launch alone on the last group (CUDA events,
    one stream) and state the bytes it must move."""
    wl.step(None)  # a pipelined step: leaves the last group's accumulators / geometry in the pipeline's slots
    vp = wl.vp
    if vp is None or not getattr(vp, "last_group", None):
        return None
    radii = vp.replay_last_group(wl.grads)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        vp.replay_last_group(wl.grads)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = len(radii)
    live = torch.stack([r > 0 for r in radii])
    v_sum, v_any = int(live.sum().item()), int(live.any(0).sum().item())
    P = wl.P
    # per Gaussian some view rendered: mean 12 + rotation 16 + scale 12 + SH row 192 read once, gradient rows 236 written
    # once; per (Gaussian, view that rendered it): accumulator row 48 + conic 16 + clamp bits 1; radii of every view 4 n P
    own = v_any * (12 + 16 + 12 + 192 + 236) + v_sum * (48 + 16 + 1) + 4 * n * P + (P - v_any) * 236
    per_view_alg = 559 * (v_sum / n) + 300 * P  # SURVEY 8(d) formula of the per-view kernel it replaces
    dram = None
    try:
        dram = json.load(open(os.path.join(ROOT, "profiles", "r2c_traffic.json")))["preprocess_bwd_views"]
    except Exception:
        pass
    return {"ms_per_launch": ms, "views_per_launch": n, "ms_per_view": ms / n, "alg_own_bytes": int(own),
            "dram_bytes": dram.get("dram_bytes_per_launch") if dram else None,
            "dram_bytes_source": "profiles/r2c_traffic.json (ncu --set full, same workload)" if dram else None,
            "ncu_issue_active_pct": dram.get("issue_active_pct") if dram else None,
            "hbm_gbs": own / (ms * 1e-3) / 1e9, "hbm_frac": own / (ms * 1e-3) / 1e9 / hbm_peak,
            "bytes_of_the_per_view_kernels_it_replaces": int(per_view_alg * n),
            "what": "gsr_backward_preprocess_views: attributes read once and gradient rows written once per group of views"}


def fused_forward_leg(torch, wl, hbm_peak, reps=10):
    """The grouped preprocess of the pipelined step (gsr_forward_preprocess_views) timed alone on the last group."""
    vp = wl.vp
    if vp is None or not getattr(vp, "last_fwd_group", None):
        return None
    torch.cuda.synchronize()
    radii = vp.replay_last_preprocess()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        vp.replay_last_preprocess()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    n = len(radii)
    live = torch.stack([r > 0 for r in radii])
    v_sum, v_any = int(live.sum().item()), int(live.any(0).sum().item())
    P = wl.P
    # attributes once per Gaussian (mean 12, scale 12, rotation 16, opacity 4; SH row 192 if some view renders it);
    # per (Gaussian, view): radii 4 + tiles 4 + rect 8 + depth key 4 always, 69 more of geometry state when rendered
    own = P * 44 + v_any * 192 + n * P * 20 + v_sum * 69
    return {"ms_per_launch": ms, "views_per_launch": n, "ms_per_view": ms / n, "alg_own_bytes": int(own),
            "hbm_gbs": own / (ms * 1e-3) / 1e9, "hbm_frac": own / (ms * 1e-3) / 1e9 / hbm_peak,
            "bytes_of_the_per_view_kernels_it_replaces": int(311 * v_sum + 20 * (n * P - v_sum)),
            "what": "gsr_forward_preprocess_views: attributes read once per Gaussian for a group of views; per-view "
                    "outputs bit-identical to preprocess_fwd_kernel's"}


def stage_table(torch, wl, steps, hbm_peak, sm_max_mhz):
    """Per-stage average duration (CUDA events on the launching stream, gsr_profile_*) and roofline
    fractions from the algorithmic bytes of SURVEY.md section 8(d) / DESIGN.md."""
    import ctypes as C
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()
    lib.gsr_profile_enable(1)
    for _ in range(steps):
        wl.step(None, overlap=False)  # per-kernel durations: one stream, nothing running underneath
    ms = (C.c_float * len(_abi.STAGES))()
    calls = (C.c_int32 * len(_abi.STAGES))()
    lib.gsr_profile_read(ms, calls)
    lib.gsr_profile_enable(0)
    wl.view_fwd_only(len(wl.cams) - 1)  # keeps the saved state alive for the fetch below
    P, L = wl.P, wl.last_fn.num_rendered
    V = int((wl.last_radii > 0).sum().item())
    n_contrib = fetch_u32(torch, wl, "n_contrib")
    pairs = int(n_contrib.astype(np.int64).sum())
    tiles = ((wl.W + 15) // 16) * ((wl.H + 15) // 16)
    end_bit = 32 + max(1, int(tiles).bit_length())
    passes = (end_bit + 7) // 8
    HW = wl.W * wl.H
    # Bytes per launch (per view).  alg_ref = SURVEY.md 8(d): what the REFERENCE algorithm must move (64-bit keys, six
    # 8-bit sort passes, ...).  alg_own = what THIS design must move: Gaussians ordered by a
    # 32-bit depth sort of P pairs (histogram read + 4 passes), tile ids of L pairs in 2 passes with the histogram built
    # by the emitting kernel, 4-byte tile ids in the range pass.  hbm_frac is computed on alg_own.
    alg_ref = {
        "preprocess_fwd": 311 * V + 20 * (P - V),
        "duplicate_keys": 20 * P + 12 * L,
        "sort": (8 + 24 * passes) * L,
        "tile_ranges": 8 * L + 8 * tiles,
        "blend_fwd": 40 * L + 20 * HW,
        "blend_bwd": 40 * L + 20 * HW + 36 * L,
        "preprocess_bwd": 559 * V + 300 * P,
    }
    tile_passes = (max(1, int(tiles).bit_length()) + 7) // 8
    alg_own = dict(alg_ref)
    alg_own["duplicate_keys"] = (12 * P) + (12 * P + 8 * V + 8 * L)   # depth-order scan + emission of 8-byte pairs
    alg_own["sort"] = (4 + 16 * 4) * P + 16 * tile_passes * L            # depth sort + tile sort
    alg_own["tile_ranges"] = 4 * L + 8 * tiles
    cap = {}
    for nm in ("r2c_traffic.json", "r2b_traffic.json", "r2_traffic.json", "r1_traffic.json"):  # committed `ncu --set full` capture of one view (tools/ncu_traffic.py)
        try:
            cap = json.load(open(os.path.join(ROOT, "profiles", nm)))
            cap["_file"] = "profiles/" + nm
            break
        except Exception:
            continue
    # blend stages are instruction-issue bound: 148 SMs x 4 schedulers x clock warp-instructions / s
    ref_lane_instr = {"blend_fwd": 20, "blend_bwd": 45}  # lane-instructions per examined pair, reference algorithm
    fp32_peak = 148 * 128 * sm_max_mhz * 1e6   # lane-instructions / s
    issue_peak = 148 * 4 * sm_max_mhz * 1e6    # warp-instructions / s
    out, total = {}, 0.0
    views = max(int(calls[_abi.STAGES.index("blend_fwd")]), 1)
    for i, name in enumerate(_abi.STAGES):
        avg = float(ms[i]) / views  # a stage may be recorded more than once per view (two sorts)
        total += avg
        e = {"ms": avg, "calls": int(calls[i]), "alg_own_bytes": int(alg_own[name]), "alg_ref_bytes": int(alg_ref[name]),
             "hbm_gbs": alg_own[name] / (avg * 1e-3) / 1e9 if avg > 0 else None}
        e["hbm_frac"] = e["hbm_gbs"] / hbm_peak if e["hbm_gbs"] else None
        e["hbm_frac_on_ref_bytes"] = alg_ref[name] / (avg * 1e-3) / 1e9 / hbm_peak if avg > 0 else None
        c = cap.get({"sort": "sort_pass", "duplicate_keys": "duplicate_keys"}.get(name, name))
        if c and name != "sort":  # (the sort stage is six launches; its per-launch capture is under stages.sort.ncu)
            e["dram_bytes"] = c.get("dram_bytes_per_launch")
            e["dram_bytes_source"] = cap.get("_file")
        if name in ref_lane_instr:
            e["bound"] = "issue"
            e["pairs_lower_bound"] = pairs
            e["vs_reference_algorithm"] = pairs * ref_lane_instr[name] / (avg * 1e-3) / fp32_peak if avg > 0 else None
            if c and c.get("warp_inst_executed_per_launch"):
                e["warp_inst_per_launch"] = c["warp_inst_executed_per_launch"]
                e["issue_frac"] = c["warp_inst_executed_per_launch"] / (avg * 1e-3) / issue_peak if avg > 0 else None
                e["ncu"] = {k: c[k] for k in ("issue_active_pct", "l1tex_data_pipe_wavefronts_pct") if k in c}
                e["ncu"]["source"] = cap.get("_file")
        else:
            e["bound"] = "hbm"
            if c and name == "sort":
                e["ncu"] = {"per_pass": c, "source": cap.get("_file")}
        out[name] = e
    for e in out.values():
        e["share"] = e["ms"] / total if total > 0 else None
    return out, {"P": P, "V": V, "L": int(L), "sum_n_contrib": pairs, "tiles": tiles, "sort_passes_ref": passes,
                 "sort_passes_own": f"4 over P + {tile_passes} over L"}


def ref_cuda_views_per_sec(torch, wl, n_views):
    """The reference's own CUDA rasterizer (unmodified DGR/cuda_rasterizer sources compiled for sm_100a behind
    oracle/ref_shim.cu -> oracle/_ref/libdgr_ref.so) on the same B200, same inputs, fwd+bwd per view, including the
    zero-filled output / gradient tensors its torch glue allocates per call (rasterize_points.cu:68-78,151-159).
    A second baseline (north_star); None when the library did not travel."""
    import ctypes as C
    so = os.path.join(ROOT, "oracle", "_ref", "libdgr_ref.so")
    if not os.path.exists(so):
        return None
    L = C.CDLL(so)
    p = lambda t: C.c_void_p(t.data_ptr())
    P, W, H, pr = wl.P, wl.W, wl.H, wl.params
    det = {k: v.detach() for k, v in pr.items()}

    def one(v):
        c, cm = wl.cams[v], wl.dev_cam[v]
        out = torch.zeros((3, H, W), device=wl.dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=wl.dev)
        R = L.ref_forward(P, 3, 16, p(wl.bg), W, H, p(det["means3D"]), p(det["shs"]), None, p(det["opacities"]),
                          p(det["scales"]), C.c_float(1.0), p(det["rotations"]), None, p(cm[0:16]), p(cm[16:32]),
                          p(cm[32:35]), C.c_float(c["tanfovx"]), C.c_float(c["tanfovy"]), 0, p(out), p(radii), 0)
        assert R >= 0
        z = lambda *s: torch.zeros(s, device=wl.dev)
        g = [z(P, 3), z(P, 4), z(P, 1), z(P, 3), z(P, 3), z(P, 6), z(P, 16, 3), z(P, 3), z(P, 4)]
        r = L.ref_backward(P, 3, 16, R, p(wl.bg), W, H, p(det["means3D"]), p(det["shs"]), None, p(det["scales"]),
                           C.c_float(1.0), p(det["rotations"]), None, p(cm[0:16]), p(cm[16:32]), p(cm[32:35]),
                           C.c_float(c["tanfovx"]), C.c_float(c["tanfovy"]), p(radii), p(wl.dev_dL[v]),
                           *[p(t) for t in g], 0)
        assert r == 0

    nv = len(wl.cams)
    for v in range(min(2, nv)):
        one(v)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n_views):
        one(i % nv)
    b.record()
    torch.cuda.synchronize()
    return n_views / (a.elapsed_time(b) * 1e-3)


def c3_leg(args, torch, dist, rank, world, dev, wl_c2):
    """C3 (BASELINE configs[2]): 3M Gaussians SH3, 1920x1080, 64 views per PGD step sharded over the ranks, NCCL
    all-reduce of the attribute gradients.  Strong scaling: the step is fixed, each rank renders 64 / N views."""
    # release the C2 workload's device and pinned buffers first
    for k in list(vars(wl_c2)):
        setattr(wl_c2, k, None)
    torch.cuda.empty_cache()
    a3 = argparse.Namespace(**{**vars(args), "workload": "C3", "views": 64 // world, "P": None, "api": "views",
                               "no_overlap": False, "sh_objs": False})
    wl = Workload(a3, torch, rank, world, dev)
    steps = 2
    ms = timed(torch, dist, lambda: wl.step(dist), steps, 1, dev)
    out = {"metric": metric_name(wl.P, wl.W, wl.H), "unit": UNIT, "value": 64 * steps / (ms * 1e-3), "ms_per_step": ms / steps,
           "steps": steps, "warmup": 1, "views_per_step": 64, "views_per_rank": 64 // world, "scaling": "strong",
           "bucket_bytes": int(wl.flat_grad.numel() * 4)}
    ms_e = timed(torch, dist, lambda: wl.step_e2e_pipelined(dist), steps, 1, dev, finish=wl.finish_e2e)
    per_view_in = wl.pipe.h2d_bytes - 4 * (wl.pipe.grad_slice[1] - wl.pipe.grad_slice[0])
    out["e2e"] = {"value": 64 * steps / (ms_e * 1e-3), "unit": UNIT, "ms_per_step": ms_e / steps,
                  "h2d_bytes_per_step": int(4 * wl.pipe.n_param + world * per_view_in),
                  "d2h_bytes_per_step": int(4 * wl.pipe.n_param)}
    if dist is not None:
        reps = 5
        ms_ar = timed(torch, dist, lambda: dist.all_reduce(wl.flat_grad), reps, 2, dev) / reps
        nbytes = wl.flat_grad.numel() * 4
        out["all_reduce"] = {"ms": ms_ar, "bytes": int(nbytes),
                             "bus_gbs": 2 * (world - 1) / world * nbytes / (ms_ar * 1e-3) / 1e9}
    for k in list(vars(wl)):
        setattr(wl, k, None)
    torch.cuda.empty_cache()
    return out


def run_gsr(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()  # fails loudly if the CUDA library is missing
    from diff_gaussian_rasterization.pipeline import bind_host_to_gpu_numa
    numa_cpus = bind_host_to_gpu_numa(dev)  # pinned host buffers below are then local to this GPU's PCIe root
    wl = Workload(args, torch, rank, world, dev)
    hbm_peak, peak_src, sm_max = load_peaks()

    # ---- value: inputs resident in HBM
    sampler = ClockSampler(local) if rank == 0 else None
    for _ in range(args.warmup):
        wl.step(dist)
    torch.cuda.synchronize()
    lib.gsr_launch_count(1)
    if sampler:
        sampler.start()
    ms = timed(torch, dist, lambda: wl.step(dist), args.steps, 0, dev)
    clocks = sampler.stop() if sampler else None
    launches = int(lib.gsr_launch_count(1))
    views_total = world * len(wl.cams) * args.steps
    value = views_total / (ms * 1e-3)

    # ---- e2e: host buffers in, host gradients out, every step
    e2e = None
    if not args.no_e2e:
        ms_s = timed(torch, dist, lambda: wl.step_e2e(dist), max(2, args.steps // 4), 1, dev) / max(2, args.steps // 4)
        ms_e = timed(torch, dist, lambda: wl.step_e2e_pipelined(dist), args.steps, max(2, args.warmup - 1), dev,
                     finish=wl.finish_e2e)
        per_view_in = wl.pipe.h2d_bytes - 4 * (wl.pipe.grad_slice[1] - wl.pipe.grad_slice[0])  # cameras + dL/dimage of a rank
        e2e = {"value": views_total / (ms_e * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(4 * wl.pipe.n_param + world * per_view_in),  # whole job: attributes once + every rank's views
               "d2h_bytes_per_step": int(4 * wl.pipe.n_param),                        # whole job: the summed gradient once
               "per_rank": {"h2d_bytes_per_step": int(wl.pipe.h2d_bytes), "d2h_bytes_per_step": int(wl.pipe.d2h_bytes)},
               "ms_per_step": ms_e / args.steps,
               "how": "pipeline.HostStepPipeline depth 2: H2D of step i+1 and D2H of step i-1 overlap the compute of "
                      "step i; all copies, the pipeline fill and the drain are inside the timed region"
                      + ("; with N > 1 rank r uploads slice r of the attribute vector over its own PCIe link and an all-gather over "
                         "NVLink (own communicator, copy-in stream) completes it on every GPU, every rank uploads its own cameras "
                         "and dL/dimage, the gradient is reduce-scattered and rank r reads slice r of the sum back to its host "
                         "buffer" if world > 1 else ""),
               "host_cpus": numa_cpus,
               "serial": {"value": world * len(wl.cams) / (ms_s * 1e-3), "unit": UNIT, "ms_per_step": ms_s,
                          "how": "same step, copies and compute on one stream, nothing overlapped"}}

    # ---- the same step through the drop-in autograd module (what gaussian_renderer.render() calls), one stream
    drop_in = None
    if args.api == "views" and not args.no_drop_in:
        drop_in = {}
        n_d = max(2, args.steps // 4)
        for key, objs in (("autograd", False), ("autograd_sh_objs", True)):
            wl.api, wl.sh_objs = "autograd", objs
            ms_d = timed(torch, dist, lambda: wl.step(dist, overlap=False), n_d, 2, dev)
            drop_in[key] = {"value": world * len(wl.cams) * n_d / (ms_d * 1e-3), "unit": UNIT, "ms_per_step": ms_d / n_d}
        drop_in["how"] = ("GaussianRasterizer + torch autograd per view, .grad accumulation by torch, one stream; "
                          "autograd_sh_objs passes sh_objs (16 object channels rendered, colour-only loss) and builds the "
                          "screen-space tensor per call like gaussian_renderer.render()")
        wl.api, wl.sh_objs = args.api, bool(args.sh_objs)

    # ---- per-stage timing (rank 0 reports; every rank runs it to stay in lock-step)
    stages, counts = stage_table(torch, wl, max(1, min(args.steps, 5)), hbm_peak, sm_max)
    fused = fused_backward_leg(torch, wl, hbm_peak) if args.api == "views" and wl.overlap else None
    if fused:
        stages["preprocess_bwd"]["fused_views"] = fused
        ff = fused_forward_leg(torch, wl, hbm_peak)
        if ff:
            stages["preprocess_fwd"]["fused_views"] = ff
    if dist is not None:
        dist.barrier()

    # ---- BASELINE configs[2] as a second, smaller measurement in the same line (extra.c3): one DAGGER PGD step =
    # 64 camera views of a 3M-Gaussian scene, split over the ranks (strong scaling), one all-reduce of the 708 MB bucket
    do_c3 = args.workload == "C2" and args.P is None and not args.no_c3 and 64 % world == 0 and args.api == "views"
    if rank != 0:
        if do_c3:
            c3_leg(args, torch, dist, rank, world, dev, wl)
        if dist is not None:
            dist.destroy_process_group()
        return
    top = max(stages, key=lambda k: stages[k]["ms"])
    t = stages[top]
    if t["bound"] == "hbm":
        roofline = {"kernel": top, "bound": "hbm", "achieved": t["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s",
                    "frac": t["hbm_frac"], "traffic": t.get("dram_bytes"), "traffic_source": t.get("dram_bytes_source"),
                    "peak_source": peak_src}
    else:
        issue_peak_g = 148 * 4 * sm_max * 1e6 / 1e9
        roofline = {"kernel": top, "bound": "issue", "unit": "G warp-instr/s", "peak": issue_peak_g,
                    "achieved": (t["issue_frac"] * issue_peak_g) if t.get("issue_frac") else None,
                    "frac": t.get("issue_frac"), "traffic": t.get("dram_bytes"), "traffic_source": t.get("dram_bytes_source"),
                    "vs_reference_algorithm": t["vs_reference_algorithm"],
                    "hbm_view": {"achieved": t["hbm_gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": t["hbm_frac"]},
                    "peak_source": "148 SMs x 4 schedulers x sm_max_mhz (one warp-instruction per scheduler per clock); HBM " + peak_src,
                    "note": "frac = warp-instructions this kernel executes per launch (ncu smsp__inst_executed.sum of the committed "
                            "capture, traffic_source) / live CUDA-event time / issue peak.  vs_reference_algorithm = SURVEY 8(d) "
                            "lane-instructions of the REFERENCE algorithm (sum n_contrib x 45 bwd / 20 fwd) / time / (148 x 128 "
                            "lanes x clock): above 1 because culling and the one-scalar recurrence execute fewer instructions."}
        if t.get("ncu"):
            roofline["ncu"] = t["ncu"]
    c = wl.scene["cam"]
    line = {
        "metric": metric_name(wl.P, wl.W, wl.H), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {wl.P} Gaussians SH3, {c['image_width']}x{c['image_height']}, fwd+bwd",
                   "views_per_rank_per_step": len(wl.cams), "parallelism": f"view-sharded x{world}, replicated Gaussians"
                   + (", NCCL all-reduce of attribute gradients per step" if world > 1 else ""),
                   "l2": "inputs larger than L2 (236 MB of attributes + 25 MB dL/dpixel per view)",
                   "step": ("every camera: depth order, binning, blend_fwd, blend_bwd; the per-Gaussian preprocess and its "
                            "backward once per group of up to 16 cameras of the step (views.ViewPipeline)"
                            if args.api == "views" and wl.overlap else "every camera on its own, one stream"),
                   **counts},
        "e2e": e2e, "drop_in": drop_in, "extra": {"c3": None}, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "stages": stages,
    }
    if not args.no_cpu_baseline and world == 1:
        steps_c = 1
        vps, dt, cores, n = cpu_views_per_sec(wl.scene, wl.cams, steps_c, 0, args.cpu_sample_views)
        line["cpu_baseline"] = {"value": vps, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{n} full view(s) fwd+bwd of the same workload, {dt:.1f} s, OpenMP over all host cores"}
    if world == 1:
        rc = ref_cuda_views_per_sec(torch, wl, 2 * len(wl.cams))
        line["ref_cuda_baseline"] = None if rc is None else {
            "value": rc, "unit": UNIT, "what": "reference CUDA rasterizer (DGR sources, sm_100a build) on this B200, "
            "same inputs, legacy default stream, incl. its per-call zero-filled tensors"}
    if args.stages_out:
        json.dump({"stages": stages, "counts": counts}, open(args.stages_out, "w"), indent=1)
    if do_c3:  # last: it releases the C2 workload
        line["extra"]["c3"] = c3_leg(args, torch, dist, rank, world, dev, wl)
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_gsr(a)
