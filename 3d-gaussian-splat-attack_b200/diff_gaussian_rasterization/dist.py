"""View-sharded data parallelism for multi-view attack steps (new; no counterpart in the reference,
which loops over the cameras of a batch serially on one GPU, attack.py:481-484).

The path shards by camera view: every rank holds the full Gaussian set,
renders the views k with k % world == rank, accumulates attribute gradients locally and the ranks
sum them with ONE all-reduce per PGD step.  The gradient tensors of all attributes are views into
one flat fp32 buffer, which is the collective's bucket (no pack / unpack copies).  Works with any
torch.distributed backend: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist

__all__ = ["shard_views", "flat_layout", "FLAT_ALIGN", "FlatGradBucket", "sharded_step", "sharded_accumulate_step", "backprop_inputs"]


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Indices of the views rank `rank` renders: k = rank, rank + world, ... (round-robin keeps ring-neighbour
    cameras, whose cost is similar, on different ranks)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, n_views, world))


FLAT_ALIGN = 64  # elements: every slice of a flat buffer starts on a 256-byte boundary (include/gsr_b200.h)


def flat_layout(sizes: Dict[str, int], align: int = FLAT_ALIGN):
    """Offsets of named slices inside one flat buffer, each slice start rounded up to `align` elements, so that a
    slice handed to the kernels (16-byte vector loads / reductions on shs, rotations and their gradients) is aligned
    for ANY Gaussian count.  Returns ({name: (begin, end)}, total elements)."""
    slices, off = {}, 0
    for n, k in sizes.items():
        off = (off + align - 1) // align * align
        slices[n] = (off, off + int(k))
        off += int(k)
    return slices, (off + align - 1) // align * align


class FlatGradBucket:
    """One flat fp32 gradient buffer shared by all attribute tensors (slices 256-byte aligned, see flat_layout).

    `params` maps names to leaf tensors (xyz, SH, scale, rotation, opacity ...).  After construction each
    p.grad is a view into `self.flat`, so autograd accumulates per-view gradients straight into the bucket
    and `all_reduce()` is a single collective over it.
    """

    def __init__(self, params: Dict[str, torch.Tensor], bind: bool = True):
        """bind=False: only lay the buffer out after `params` (any tensors, e.g. activations); nothing is attached
        to .grad -- for the accumulate API (sharded_accumulate_step), which writes the slices itself."""
        if not params:
            raise ValueError("no parameters")
        ps = list(params.values())
        dev, dt = ps[0].device, ps[0].dtype
        self.bind = bind
        for n, p in params.items():
            if bind and (not p.is_leaf or not p.requires_grad):
                raise ValueError(f"{n}: must be a leaf tensor that requires grad")
            if p.device != dev or p.dtype != dt:
                raise ValueError(f"{n}: all parameters must share device and dtype")
        self.params = dict(params)
        self.slices, self.numel = flat_layout({n: p.numel() for n, p in self.params.items()})
        self.flat = torch.zeros(self.numel, dtype=dt, device=dev)  # padding stays zero: harmless in the all-reduce
        if bind:
            for n, p in self.params.items():
                a, b = self.slices[n]
                p.grad = self.flat[a:b].view(p.shape)

    def zero(self):
        self.flat.zero_()
        if not self.bind:
            return
        for n, p in self.params.items():  # re-bind in case an optimizer step replaced .grad (set_to_none=True)
            a, b = self.slices[n]
            if p.grad is None or p.grad.data_ptr() != self.flat.data_ptr() + a * self.flat.element_size():
                p.grad = self.flat[a:b].view(p.shape)

    def grad(self, name: str) -> torch.Tensor:
        a, b = self.slices[name]
        return self.flat[a:b].view(self.params[name].shape)

    def all_reduce(self, group=None, async_op: bool = False):
        """SUM over ranks, in place.  No-op without an initialised process group or with world size 1."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def global_norm(self) -> torch.Tensor:
        """L2 norm of the (already reduced) gradient: what the L2 PGD variants need (attack.py:25-173)."""
        return self.flat.norm()


def sharded_step(views: Sequence, render_backward: Callable[[object], Optional[torch.Tensor]], bucket: FlatGradBucket,
                 rank: Optional[int] = None, world: Optional[int] = None, group=None) -> torch.Tensor:
    """One gradient step over `views` (any per-view objects, e.g. cameras), sharded over the ranks.

    `render_backward(view)` must render the view, evaluate its loss and call backward (gradients land in the
    bucket); it may return the scalar loss.  Returns the loss summed over ALL views (all-reduced), after the
    bucket holds the gradient summed over all views -- identical on every rank, so applying the same PGD update
    keeps the replicas in sync without a broadcast.
    """
    on = dist.is_available() and dist.is_initialized()
    rank = (dist.get_rank(group) if on else 0) if rank is None else rank
    world = (dist.get_world_size(group) if on else 1) if world is None else world
    bucket.zero()
    total = torch.zeros((), dtype=bucket.flat.dtype, device=bucket.flat.device)
    for k in shard_views(len(views), rank, world):
        loss = render_backward(views[k])
        if loss is not None:
            total = total + loss.detach().to(total.dtype)
    bucket.all_reduce(group)
    if on and world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
    return total


_pipelines = {}  # one views.ViewPipeline per device, created on first use


def sharded_accumulate_step(settings: Sequence, inputs: Dict[str, torch.Tensor],
                            loss_fn: Callable[[torch.Tensor, int], torch.Tensor], bucket: FlatGradBucket,
                            rank: Optional[int] = None, world: Optional[int] = None, group=None,
                            forward_view=None, backward_view=None) -> torch.Tensor:
    """One multi-camera gradient step on the view-level accumulate API (views.forward_view / backward_view).

    `settings[k]` is the GaussianRasterizationSettings of camera k; `inputs` maps the rasterizer's argument names
    (means3D, opacities, shs | colors_precomp, scales + rotations | cov3D_precomp) to the tensors render() would
    pass (gaussian_renderer/__init__.py:86-95).  For each of this rank's cameras the view is rendered without an
    autograd graph, `loss_fn(image, k)` is differentiated w.r.t. the image only, and the rasterizer's backward adds
    the attribute gradients straight into the bucket (bucket.params must be keyed like `inputs`).  Then ONE
    all-reduce.  Returns the loss summed over all cameras and ranks.

    When the inputs are activations of raw parameters (sigmoid / exp / normalize, scene/gaussian_model.py:97-121)
    call `backprop_inputs(inputs, bucket)` afterwards: the activation backward then runs once per step instead of
    once per camera.
    """
    on = dist.is_available() and dist.is_initialized()
    rank = (dist.get_rank(group) if on else 0) if rank is None else rank
    world = (dist.get_world_size(group) if on else 1) if world is None else world
    if set(bucket.params) != set(inputs):
        raise ValueError(f"bucket is keyed {sorted(bucket.params)}, inputs {sorted(inputs)}")
    grads = {n: bucket.grad(n) for n in inputs}
    plain = {n: t.detach() for n, t in inputs.items()}
    total = torch.zeros((), dtype=bucket.flat.dtype, device=bucket.flat.device)
    mine = shard_views(len(settings), rank, world)
    extra = {n: t for n, t in plain.items() if n not in ("means3D", "opacities")}
    if forward_view is None and backward_view is None:
        # the three-stream pipeline: next camera's preprocess / sorts under this camera's blending kernels
        from . import views as _v
        dev = plain["means3D"].device
        vp = _pipelines.get(dev)
        if vp is None:
            vp = _pipelines[dev] = _v.ViewPipeline(dev)
        losses = []

        def dL_fn(image, i):
            image = image.requires_grad_(True)
            loss = loss_fn(image, mine[i])
            losses.append(loss.detach())
            return torch.autograd.grad(loss, image)[0]
        # overwrite: the step SETS the bucket slices to the sum over this rank's cameras (no zeroing pass)
        vp.step([settings[k] for k in mine], dL_fn, grads, plain["means3D"], plain["opacities"], overwrite=True, **extra)
        for l in losses:
            total = total + l.to(total.dtype)
    else:
        if forward_view is None or backward_view is None:
            from . import views as _v
            forward_view, backward_view = forward_view or _v.forward_view, backward_view or _v.backward_view
        bucket.zero()
        for k in mine:
            image, _radii, state = forward_view(settings[k], plain["means3D"], plain["opacities"], **extra)
            image = image.detach().requires_grad_(True)
            loss = loss_fn(image, k)
            (dL_dimage,) = torch.autograd.grad(loss, image)
            backward_view(state, dL_dimage, grads)
            total = total + loss.detach().to(total.dtype)
    bucket.all_reduce(group)
    if on and world > 1:
        dist.all_reduce(total, op=dist.ReduceOp.SUM, group=group)
    return total


def backprop_inputs(inputs: Dict[str, torch.Tensor], bucket: FlatGradBucket) -> None:
    """Chain the bucket's gradients (w.r.t. `inputs`) back to the leaves: inputs computed from raw parameters
    (sigmoid / exp / normalize ...) are back-propagated with ONE autograd call; inputs that are leaves themselves
    get .grad set to (or incremented by) their bucket slice."""
    ts, gs = [], []
    for n, t in inputs.items():
        if not t.requires_grad:
            continue
        if t.grad_fn is not None:
            ts.append(t)
            gs.append(bucket.grad(n))
        elif t.grad is None:
            t.grad = bucket.grad(n)
        elif t.grad.data_ptr() != bucket.grad(n).data_ptr():
            t.grad += bucket.grad(n)
    if ts:
        torch.autograd.backward(ts, gs)
