"""View-sharded mapping-window step (multi-GPU), the one place where the path has a real exchange step.

In Splat-SLAM one `map()` iteration renders every keyframe of the window against the SAME Gaussians, sums the
losses and calls `backward()` once (reference src/mapper.py:418-490).  The views are independent, so they shard
across GPUs: every rank holds a replica of the Gaussian parameters, renders its share of the views forward +
backward, and ONE all-reduce (sum, fp32) of the flat Gaussian-gradient buffer makes the gradients identical
everywhere.  Per-view quantities (pose gradients theta/rho, radii, n_touched) stay on the rank
that owns the view.  The reference itself is single-GPU; this module is new.

One process per GPU, `torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

import contextlib
from typing import Callable, Dict, List, Optional, Sequence

import torch
import torch.distributed as dist

# leaf parameters whose gradients are exchanged, in flat-buffer order:
# xyz 3 + sh 3*M + opacity 1 + scale 3 + rotation 4 = (11 + 3M) floats per Gaussian
PARAM_ORDER = ("means3D", "shs", "opacities", "scales", "rotations")


def partition_views(n_views: int, world_size: int, rank: int) -> List[int]:
    """Round-robin view ownership: view v belongs to rank v % world_size (10 views -> 5+5 / 3+3+2+2 / 2+2+1x6)."""
    return [v for v in range(n_views) if v % world_size == rank]


class FlatGrads:
    """One contiguous fp32 buffer holding the gradients of all Gaussian leaf parameters.

    The parameters' `.grad` fields are views into the buffer, so autograd accumulates the per-view gradients in
    place and the buffer can be handed to a single all-reduce without packing."""

    def __init__(self, params: Dict[str, torch.Tensor]):
        self.names = [n for n in PARAM_ORDER if params.get(n) is not None]
        self.params = params
        total = sum(params[n].numel() for n in self.names)
        ref = params[self.names[0]]
        self.flat = torch.zeros(total, dtype=torch.float32, device=ref.device)
        off = 0
        self.views = {}
        for n in self.names:
            p = params[n]
            v = self.flat[off:off + p.numel()].view_as(p)
            p.grad = v
            self.views[n] = v
            off += p.numel()

    def zero_(self):
        self.flat.zero_()
        for n in self.names:          # re-attach in case an optimizer replaced .grad
            if self.params[n].grad is not self.views[n]:
                self.params[n].grad = self.views[n]

    def all_reduce(self, group=None):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)


class ViewStreams:
    """Round-robin CUDA streams for the views of one window step.

    The views are independent until autograd accumulates their gradients, so issuing them on a few streams lets one
    view's latency-bound sort passes and its num_rendered hand-off (the one host sync of a forward) hide under another
    view's render kernels. `fork()` makes the side streams wait for the work queued so far, `join()` makes the
    caller's stream wait for all of them. With n <= 1 everything stays on the caller's stream.
    Measured on one B200 (bench workload): 1 stream 676 it/s, 3 streams 767, 5 streams with priority lanes 823."""

    def __init__(self, device, n: int = 5, priority_lanes: bool = True):
        self.device = torch.device(device)
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(n)] if n > 1 else []
        if self.streams:   # accumulation into the shared leaves deliberately crosses streams; autograd orders it
            torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
            if priority_lanes:
                # the ~20 short kernels of each forward go to a high-priority side stream so that they are not
                # queued behind the render CTAs of the other views in flight (library-wide switch)
                from . import _cabi
                _cabi.priority_lanes(True)

    def __len__(self):
        return len(self.streams)

    def fork(self):
        main = torch.cuda.current_stream(self.device) if self.streams else None
        for st in self.streams:
            st.wait_stream(main)

    def view(self, i: int):
        """Context manager: the stream view number i of the step runs on."""
        if not self.streams:
            return contextlib.nullcontext()
        return torch.cuda.stream(self.streams[i % len(self.streams)])

    def stream(self, i: int):
        return self.streams[i % len(self.streams)] if self.streams else torch.cuda.current_stream(self.device)

    def join(self):
        main = torch.cuda.current_stream(self.device) if self.streams else None
        for st in self.streams:
            main.wait_stream(st)


def window_step(render_view: Callable[[int], Sequence[torch.Tensor]],
                upstream: Callable[[int], Sequence[torch.Tensor]],
                flat: FlatGrads, n_views: int, group=None, views: Optional[List[int]] = None,
                streams: Optional[ViewStreams] = None) -> List[int]:
    """One mapping iteration over a window of `n_views` views.

    render_view(v) -> tensors to differentiate (e.g. (color, depth)) for global view index v;
    upstream(v)    -> matching upstream gradients (dL/dcolor, dL/ddepth).
    Gradients of the Gaussian parameters accumulate in `flat`; after the local views are done they are
    all-reduced over `group`. `streams` (CUDA only) spreads the local views over a few streams.
    Returns the list of view indices this rank rendered."""
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    rank = dist.get_rank(group) if world > 1 else 0
    mine = partition_views(n_views, world, rank) if views is None else views
    flat.zero_()
    if streams is not None and len(streams):
        streams.fork()
        for k, v in enumerate(mine):
            with streams.view(k):
                outs = render_view(v)
                torch.autograd.backward(list(outs), list(upstream(v)))
        streams.join()
    else:
        for v in mine:
            outs = render_view(v)
            torch.autograd.backward(list(outs), list(upstream(v)))
    flat.all_reduce(group)
    return mine
