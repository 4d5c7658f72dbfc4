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
import os
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
    place and the buffer can be handed to a single exchange without packing.

    `exchange` picks how `all_reduce()` sums the buffer over the ranks of `group`:
      "nccl"  `dist.all_reduce` (also what the gloo CPU tests use);
      "peer"  this library's own kernel over NVLink peer memory: the buffer is allocated as torch symmetric memory, every
              rank reduces its 1/world slice straight out of the other GPUs' buffers and writes the total back into all
              of them (`splat_raster_allreduce_multimem` through the NVSwitch when the allocation has a multicast
              address, else `splat_raster_allreduce_peer`), bracketed by two symmetric-memory barriers;
      "auto"  "peer" when every rank managed to set it up (one node, CUDA, world <= 8), else "nccl"."""

    def __init__(self, params: Dict[str, torch.Tensor], exchange: str = "nccl", group=None):
        self.names = [n for n in PARAM_ORDER if params.get(n) is not None]
        self.params = params
        total = sum(params[n].numel() for n in self.names)
        ref = params[self.names[0]]
        self.group = group
        self.exchange_kind = "nccl"
        self._peer = None
        padded = (total + 63) // 64 * 64          # whole 16-byte vectors for every rank's slice
        store = None
        if exchange in ("auto", "peer") and ref.is_cuda and dist.is_available() and dist.is_initialized() \
                and dist.get_world_size(group) > 1:
            store = self._try_symmetric(padded, ref.device, group, must=(exchange == "peer"))
        if store is None:
            store = torch.zeros(padded, dtype=torch.float32, device=ref.device)
        self._store = store
        self.flat = store[:total]
        off = 0
        self.views = {}
        for n in self.names:
            p = params[n]
            v = self.flat[off:off + p.numel()].view_as(p)
            p.grad = v
            self.views[n] = v
            off += p.numel()

    def _try_symmetric(self, padded, device, group, must):
        """Allocate the buffer as symmetric memory and exchange the peer pointers; every rank reports whether it
        worked and the peer path is only used when all of them did (otherwise all fall back to NCCL together)."""
        world = dist.get_world_size(group)
        store, err = None, None
        try:
            if world > 8:
                raise RuntimeError("peer exchange supports up to 8 ranks (one NVSwitch node)")
            import torch.distributed._symmetric_memory as symm_mem
            from . import _cabi
            _cabi.lib()
            store = symm_mem.empty(padded, dtype=torch.float32, device=device)
            store.zero_()
            hdl = symm_mem.rendezvous(store, group=(group if group is not None else dist.group.WORLD))
            ptrs = [int(x) for x in hdl.buffer_ptrs]
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            if os.environ.get("SPLAT_NO_MULTIMEM"):      # tests: the plain peer load / store kernel on a multicast-capable box
                mc = 0
            if len(ptrs) != world:
                raise RuntimeError("rendezvous returned %d buffers for %d ranks" % (len(ptrs), world))
            self._peer = dict(hdl=hdl, ptrs=ptrs, mc=mc, world=world, rank=dist.get_rank(group), count=padded)
        except Exception as ex:     # noqa: BLE001 - any failure means "use NCCL"
            err, store, self._peer = ex, None, None
        ok = torch.tensor([1 if self._peer is not None else 0], device=device, dtype=torch.int32)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) != 1:
            self._peer, store = None, None
            if must:
                raise RuntimeError(f"peer-memory gradient exchange unavailable: {err}")
            return None
        self.exchange_kind = "peer-multimem" if self._peer["mc"] else "peer"
        return store

    def zero_(self):
        self.flat.zero_()
        for n in self.names:          # re-attach in case an optimizer replaced .grad
            if self.params[n].grad is not self.views[n]:
                self.params[n].grad = self.views[n]

    def all_reduce(self, group=None):
        group = group if group is not None else self.group
        if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
            return
        if self._peer is None:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            return
        import ctypes
        from . import _cabi
        pr = self._peer
        L = _cabi.lib()
        stream = _cabi.stream_ptr(self.flat.device)
        pr["hdl"].barrier(channel=0)          # every rank's gradients are complete
        if pr["mc"]:
            _cabi.check(L.splat_raster_allreduce_multimem(pr["mc"], pr["world"], pr["rank"], pr["count"], stream),
                        "allreduce_multimem")
        else:
            arr = (ctypes.c_void_p * pr["world"])(*pr["ptrs"])
            _cabi.check(L.splat_raster_allreduce_peer(arr, pr["world"], pr["rank"], pr["count"], stream),
                        "allreduce_peer")
        pr["hdl"].barrier(channel=1)          # every rank's slice has landed in every buffer


class ViewStreams:
    """Round-robin CUDA streams for the views of one window step.

    The views are independent until autograd accumulates their gradients, so issuing them on a few streams lets one
    view's latency-bound sort passes and its num_rendered hand-off (the one host sync of a forward) hide under another
    view's render kernels. `fork()` makes the side streams wait for the work queued so far, `join()` makes the
    caller's stream wait for all of them. With n <= 1 everything stays on the caller's stream.
    Measured on one B200 (bench workload, round 2): 1 stream 800 it/s, 5 streams with priority lanes 960, one stream per
    view of the 10-view window 976."""

    def __init__(self, device, n: int = 10, priority_lanes: bool = True):
        self.device = torch.device(device)
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(n)] if n > 1 else []
        self._restore = []          # process-wide switches this object flipped, undone by close()
        if self.streams:   # accumulation into the shared leaves deliberately crosses streams; autograd orders it
            graph = torch.autograd.graph
            if hasattr(graph, "set_warn_on_accumulate_grad_stream_mismatch"):      # recent torch only
                graph.set_warn_on_accumulate_grad_stream_mismatch(False)    # advisory only; left off
            if priority_lanes:
                # the ~20 short kernels of each forward go to a high-priority side stream so that they are not
                # queued behind the render CTAs of the other views in flight (library-wide switch)
                from . import _cabi
                was = _cabi.priority_lanes(True)
                if not was:
                    self._restore.append(lambda: _cabi.priority_lanes(False))   # also frees the library's side streams

    def close(self):
        """Undo the library-wide switch this object set (priority lanes; also frees the library's side streams)."""
        while self._restore:
            self._restore.pop()()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

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
