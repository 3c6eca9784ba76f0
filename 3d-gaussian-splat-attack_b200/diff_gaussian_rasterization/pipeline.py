"""Host-buffer step pipeline: multi-view gradient steps whose inputs and outputs live in HOST memory.

A step = render `n_views` cameras forward + backward against one Gaussian set and return the summed
attribute gradients (the inner loop of attack.py:463-530 with the detector replaced by a given dL/dimage).
Inputs per step (pinned host memory): the flat attribute vector [means3D | shs | scales | rotations | opacities]
(236 B per Gaussian at SH degree 3; each slice starts on a 256-byte boundary, `HostStepPipeline.slices`, see
dist.flat_layout -- the same layout as the all-reduce bucket, dist.FlatGradBucket), one
36-float row per camera (viewmatrix 16, projmatrix 16, campos 3, pad) and one (3,H,W) dL/dimage per camera.
Output per step: the flat gradient vector in pinned host memory.

Three CUDA streams, two device buffer sets: while step i computes, the copy-in stream already moves step
i+1's inputs and the copy-out stream drains step i-1's gradients (PCIe is full duplex), so in steady state a
step costs max(compute, H2D, D2H) instead of their sum.  submit() is software-pipelined by one step because
the forward has one host sync per view (the duplicate count, as in rasterizer_impl.cu:281): the host thread
can only queue step i+1's copies before it starts issuing step i's kernels.
"""
from typing import List, Optional

import torch

from . import GaussianRasterizationSettings
from . import views as _views
from .dist import flat_layout

__all__ = ["HostStepPipeline", "ATTRS", "bind_host_to_gpu_numa", "shard_layout"]

# flat layout: name -> floats per Gaussian (SH width is 3*M, filled in per instance)
ATTRS = ("means3D", "shs", "scales", "rotations", "opacities")


def bind_host_to_gpu_numa(device) -> Optional[str]:
    """Pin the calling process to the CPUs of the NUMA node the GPU hangs off (sysfs local_cpulist of its PCI
    function), so that pinned host buffers allocated afterwards are local to that GPU's PCIe root.  With one
    process per GPU this keeps the host<->device copies of 8 ranks off the inter-socket link.  Returns the cpu
    list used, or None when the topology cannot be read (nothing is changed then)."""
    import os
    try:
        pr = torch.cuda.get_device_properties(torch.device(device))
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        cpulist = open(f"/sys/bus/pci/devices/{bdf}/local_cpulist").read().strip()
        cpus = set()
        for part in cpulist.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0) if cpus & os.sched_getaffinity(0) else cpus
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpulist
    except Exception:
        return None


def shard_layout(n_param: int, world: int, rank: int):
    """How the flat attribute / gradient vector of `n_param` floats is split over `world` ranks for host I/O:
    returns (n_pad, shard, (lo, hi)) -- device buffers hold n_pad = world * shard elements (equal slices for
    all_gather_into_tensor / reduce_scatter_tensor) and rank `rank` moves host elements [lo, hi) (the last slice is
    cut at n_param).  The slices of all ranks tile [0, n_param) exactly."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    n_pad = (n_param + world - 1) // world * world
    shard = n_pad // world
    lo = min(rank * shard, n_param)
    return n_pad, shard, (lo, min(lo + shard, n_param))


class _Slot:
    def __init__(self, n_param, n_views, H, W, dev, widths, P, n_alloc=None, shard=0):
        f32 = torch.float32
        n_alloc = n_param if n_alloc is None else n_alloc  # padded so that it splits into equal per-rank slices
        self.param = torch.empty(n_alloc, dtype=f32, device=dev)
        self.grad = torch.zeros(n_alloc, dtype=f32, device=dev)
        self.grad_shard = torch.empty(shard, dtype=f32, device=dev) if shard else None  # reduce-scatter output
        self.cam = torch.empty((n_views, 36), dtype=f32, device=dev)
        self.dL = torch.empty((n_views, 3, H, W), dtype=f32, device=dev)
        self.p, self.g = {}, {}
        slices, _ = flat_layout({n: P * w for n, w in widths})
        for n, w in widths:
            shape = (P, w // 3, 3) if n == "shs" else (P, w)
            a, b = slices[n]
            self.p[n] = self.param[a:b].view(shape)
            self.g[n] = self.grad[a:b].view(shape)
        self.ev_in = torch.cuda.Event()
        self.ev_compute = torch.cuda.Event()
        self.ev_out = torch.cuda.Event()
        self.pending_out = None  # host gradient tensor the slot's last D2H writes to
        self.used = False


class HostStepPipeline:
    def __init__(self, P: int, M: int, W: int, H: int, n_views: int, tanfovx: float, tanfovy: float, sh_degree: int,
                 device, bg: Optional[torch.Tensor] = None, depth: int = 2, group=None, overlap: bool = True,
                 share_over_nvlink: bool = True):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.P, self.M, self.W, self.H, self.n_views = P, M, W, H, n_views
        self.tanfovx, self.tanfovy, self.sh_degree = float(tanfovx), float(tanfovy), int(sh_degree)
        self.dev = torch.device(device)
        self.widths = (("means3D", 3), ("shs", 3 * M), ("scales", 3), ("rotations", 4), ("opacities", 1))
        # host and device share one layout: slices[name] = (begin, end) in floats, every begin 256-byte aligned
        self.slices, self.n_param = flat_layout({n: P * w for n, w in self.widths})
        self.bg = torch.zeros(3, device=self.dev) if bg is None else bg.to(self.dev).float()
        self.s_in = torch.cuda.Stream(self.dev)
        self.s_out = torch.cuda.Stream(self.dev)
        self.group = group
        # With several ranks on one box every rank would pull the same 236 B/Gaussian attribute vector over PCIe and
        # push the same reduced gradient back.  Instead the host traffic is SHARDED over the ranks' own PCIe links:
        # rank r uploads slice r of the attribute vector, an all-gather over NVLink completes it on every GPU; the
        # gradient is reduce-scattered and rank r copies slice r of the sum to its host buffer (the slices of all ranks
        # concatenate to the full gradient: `grad_slice`).  The input all-gather runs on its own process group /
        # communicator on the copy-in stream, so it never queues behind the gradient collective of the previous step.
        self.world, self.rank = 1, 0
        if group is not False:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.share = bool(share_over_nvlink) and self.world > 1
        self.in_group = None
        if self.share:
            import torch.distributed as dist
            ranks = dist.get_process_group_ranks(group) if group is not None else None
            self.in_group = dist.new_group(ranks=ranks)  # collective: every rank of `group` constructs the pipeline
        self.n_pad, self.shard, sl = shard_layout(self.n_param, self.world, self.rank)
        self.grad_slice = sl if self.share else (0, self.n_param)
        self.slots: List[_Slot] = [_Slot(self.n_param, n_views, H, W, self.dev, self.widths, P, n_alloc=self.n_pad,
                                         shard=self.shard if self.share else 0) for _ in range(depth)]
        self.views = _views.ViewPipeline(self.dev, overlap=overlap)
        self.n_submitted = 0
        self.staged = None  # slot whose inputs are in flight and whose compute has not been issued yet
        n_io = (self.grad_slice[1] - self.grad_slice[0])
        self.h2d_bytes = 4 * (n_io + n_views * 36 + n_views * 3 * H * W)  # per rank
        self.d2h_bytes = 4 * n_io                                         # per rank

    # -- stages ---------------------------------------------------------------------------------------------
    def _copy_in(self, slot, host_param, host_cam, host_dL):
        with torch.cuda.stream(self.s_in):
            if slot.used:
                self.s_in.wait_event(slot.ev_compute)  # the previous step in this slot has finished reading
            if not self.share:
                slot.param[:self.n_param].copy_(host_param, non_blocking=True)
            else:
                import torch.distributed as dist
                lo, hi = self.grad_slice
                mine = slot.param[self.rank * self.shard:(self.rank + 1) * self.shard]
                mine[:hi - lo].copy_(host_param[lo:hi], non_blocking=True)  # this rank's slice over its own PCIe link
                dist.all_gather_into_tensor(slot.param, mine, group=self.in_group)  # in place, over NVLink
            slot.cam.copy_(host_cam, non_blocking=True)
            slot.dL.copy_(host_dL, non_blocking=True)
            slot.ev_in.record(self.s_in)

    def _compute(self, slot):
        cur = torch.cuda.current_stream(self.dev)
        cur.wait_event(slot.ev_in)
        if slot.used:
            cur.wait_event(slot.ev_out)  # the previous gradients of this slot have left
        p = slot.p  # (no zeroing of slot.grad: the step below overwrites it; the padding between slices stays zero)
        settings = []
        for v in range(self.n_views):
            cm = slot.cam[v]
            settings.append(GaussianRasterizationSettings(
                image_height=self.H, image_width=self.W, tanfovx=self.tanfovx, tanfovy=self.tanfovy, bg=self.bg,
                scale_modifier=1.0, viewmatrix=cm[0:16].view(4, 4), projmatrix=cm[16:32].view(4, 4),
                sh_degree=self.sh_degree, campos=cm[32:35], prefiltered=False, debug=False))
        self.views.step(settings, lambda img, k: slot.dL[k], slot.g, p["means3D"], p["opacities"], shs=p["shs"],
                        scales=p["scales"], rotations=p["rotations"], overwrite=True)
        if self.world > 1:
            import torch.distributed as dist
            if self.share:  # rank r keeps slice r of the sum
                dist.reduce_scatter_tensor(slot.grad_shard, slot.grad, group=self.group)
            else:
                dist.all_reduce(slot.grad, group=self.group)
        slot.ev_compute.record(cur)

    def _copy_out(self, slot, host_grad):
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(slot.ev_compute)
            if not self.share:
                host_grad.copy_(slot.grad[:self.n_param], non_blocking=True)
            else:
                lo, hi = self.grad_slice
                host_grad[lo:hi].copy_(slot.grad_shard[:hi - lo], non_blocking=True)
            slot.ev_out.record(self.s_out)
        slot.pending_out = host_grad
        slot.used = True

    # -- public ---------------------------------------------------------------------------------------------
    def submit(self, host_param, host_cam, host_dL, host_grad):
        """Queue one step.  All four tensors are pinned host tensors (flat params, (n_views,36) cameras,
        (n_views,3,H,W) dL/dimage, flat gradient output).  The inputs must stay unchanged, and host_grad is not
        valid, until a later submit()/flush() returns it.  Returns the list of host_grad tensors that became
        valid during this call (steps complete in submission order).  With several ranks (share_over_nvlink) only
        host_grad[grad_slice[0]:grad_slice[1]] is written on this rank: the ranks' slices concatenate to the sum."""
        for t, n in ((host_param, "host_param"), (host_cam, "host_cam"), (host_dL, "host_dL"), (host_grad, "host_grad")):
            if t.is_cuda or not t.is_pinned():
                raise RuntimeError(f"{n} must be a pinned host tensor")
        if host_param.numel() != self.n_param or host_grad.numel() != self.n_param:
            raise RuntimeError(f"host_param / host_grad need {self.n_param} elements")
        done = []
        slot = self.slots[self.n_submitted % len(self.slots)]
        if self.staged is not None and slot is self.staged[0]:  # depth 1: nothing to overlap with
            done += self.flush()
        self._copy_in(slot, host_param, host_cam, host_dL)  # stream-ordered after the slot's previous compute
        prev, self.staged = self.staged, (slot, host_grad)
        self.n_submitted += 1
        if prev is not None:
            self._compute(prev[0])
            self._copy_out(prev[0], prev[1])
        # collect what has certainly had time to finish: every step older than the one just issued
        for sl in self.slots:
            if sl.pending_out is not None and (prev is None or sl is not prev[0]):
                sl.ev_out.synchronize()
                done.append(sl.pending_out)
                sl.pending_out = None
        return done

    def flush(self):
        """Run everything queued to completion; returns the host_grad tensors that became valid, in order."""
        if self.staged is not None:
            slot, host_grad = self.staged
            self.staged = None
            self._compute(slot)
            self._copy_out(slot, host_grad)
        done = []
        n, d = self.n_submitted, len(self.slots)
        for k in range(max(0, n - d), n):  # oldest first
            slot = self.slots[k % d]
            if slot.pending_out is not None:
                slot.ev_out.synchronize()
                done.append(slot.pending_out)
                slot.pending_out = None
        return done
