"""Two-GPU test of the host-buffer step pipeline with sharded host I/O (skipped on a single-GPU box): rank r uploads
slice r of the attribute vector, an all-gather completes it on both GPUs, the gradient is reduce-scattered and rank r's
host buffer receives slice r of the sum.  The concatenated slices must equal the single-process result over all views."""
import os
import socket

import numpy as np
import pytest
import torch

import helpers as h

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs")]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, P, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from diff_gaussian_rasterization.pipeline import HostStepPipeline
    from gsr_synth import ring_cameras
    s = h.make_scene("tiny", P=P)
    W, H, n_views = 160, 120, 4
    cams = ring_cameras(n_views, width=W, height=H)[rank::world]
    rng = np.random.default_rng(5)
    dL_all = (rng.standard_normal((n_views, 3, H, W)) / (3 * H * W)).astype(np.float32)
    pipe = HostStepPipeline(P, 16, W, H, len(cams), cams[0]["tanfovx"], cams[0]["tanfovy"], 3, f"cuda:{rank}", depth=2)
    host_param = torch.zeros(pipe.n_param).pin_memory()
    widths = dict(means3D=3, shs=48, scales=3, rotations=4, opacities=1)
    for n in widths:
        a, b = pipe.slices[n]
        host_param[a:b] = torch.from_numpy(s[n].reshape(-1))
    host_cam = torch.from_numpy(np.stack([np.concatenate([c["viewmatrix"].reshape(-1), c["projmatrix"].reshape(-1),
                                                          c["campos"], np.zeros(1, np.float32)]) for c in cams])).pin_memory()
    host_dL = torch.from_numpy(np.ascontiguousarray(dL_all[rank::world])).pin_memory()
    outs = [torch.full((pipe.n_param,), float("nan")).pin_memory() for _ in range(3)]
    done = []
    for o in outs:
        done += pipe.submit(host_param, host_cam, host_dL, o)
    done += pipe.flush()
    assert len(done) == 3
    lo, hi = pipe.grad_slice
    for o in done:
        assert torch.isnan(o[:lo]).all() and torch.isnan(o[hi:]).all()  # only this rank's slice is written
    np.save(os.path.join(out_dir, f"slice_{rank}.npy"), np.concatenate([[lo, hi], done[-1][lo:hi].numpy()]))
    dist.destroy_process_group()


@pytest.mark.parametrize("P", [2001])
def test_host_pipeline_shards_host_io_over_two_ranks(P, tmp_path):
    import torch.multiprocessing as mp
    from diff_gaussian_rasterization.dist import flat_layout
    from diff_gaussian_rasterization import views
    from gsr_synth import ring_cameras
    mp.spawn(_worker, args=(2, _free_port(), P, str(tmp_path)), nprocs=2, join=True)
    # single-process result over all four views
    s = h.make_scene("tiny", P=P)
    W, H = 160, 120
    rng = np.random.default_rng(5)
    dL_all = (rng.standard_normal((4, 3, H, W)) / (3 * H * W)).astype(np.float32)
    t = h.scene_tensors(s, requires_grad=False)
    want = {k: torch.zeros_like(v) for k, v in t.items()}
    for c, dL in zip(ring_cameras(4, width=W, height=H), dL_all):
        sc = dict(s); sc["cam"] = c
        _, _, st = views.forward_view(h.settings_for(sc), t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"],
                                      rotations=t["rotations"])
        views.backward_view(st, torch.from_numpy(dL).cuda(), want)
    widths = dict(means3D=3, shs=48, scales=3, rotations=4, opacities=1)
    slices, n_flat = flat_layout({n: P * w for n, w in widths.items()})
    full = np.zeros(n_flat, np.float32)
    covered = 0
    for r in range(2):
        z = np.load(os.path.join(str(tmp_path), f"slice_{r}.npy"))
        lo, hi = int(z[0]), int(z[1])
        full[lo:hi] = z[2:]
        covered += hi - lo
    assert covered == n_flat
    for n in widths:
        ok, worst = h.grad_close(full[slices[n][0]:slices[n][1]], want[n].cpu().numpy().reshape(-1))
        assert ok, (n, worst)
