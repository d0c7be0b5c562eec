"""Multi-GPU plumbing: one process per GPU (torchrun), tasks of a meta-batch sharded across ranks, ONE all-reduce
(sum) of the flat fp32 meta-gradient per outer step.  The reference is single-process
(fs_mol/utils/adaptive_dkt_utils.py:352-413 loops the tasks sequentially on one device); what is mirrored here is its
accumulation rule: grad = sum_t hypergrad_t / tasks_per_batch (:402-403), then clip and Adam on identical replicas.

Meta-test needs no collective at all: tasks are independent, each rank evaluates its shard.
"""
import os
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def init_from_env(device=None):
    """Initialise the default process group from torchrun's environment (NCCL on CUDA, gloo on CPU).
    Returns (rank, world_size).  A single process needs no group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if device is not None and torch.device(device).type == "cuda":
            dist.init_process_group("nccl", device_id=torch.device(device))
        else:
            dist.init_process_group("gloo")
    return rank, world


def world_size() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def shard_tasks(num_tasks: int, rank: int, world: int) -> List[int]:
    """Round-robin task shard of one meta-batch: rank r owns tasks r, r+R, ... (balanced to within one task)."""
    return list(range(rank, num_tasks, world))


def allreduce_sum_(flat: torch.Tensor) -> torch.Tensor:
    if world_size() > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    return flat


class FlatParams:
    """Feature-extractor parameters and gradients as views of two flat fp32 buffers, so that the all-reduce, the
    gradient-norm clip and Adam each touch ONE contiguous array (25.34 M floats = 101 MB for the CLI model).
    Every tensor starts on a 256-byte boundary (TMA tensor maps of the GEMM kernel need 16-byte aligned operands);
    the padding floats stay zero in both buffers, so norms and the all-reduced sum are unaffected."""

    ALIGN = 64  # floats

    def __init__(self, params: Sequence[torch.nn.Parameter]):
        self.params = list(params)
        self.sizes = [p.numel() for p in self.params]
        self.offsets = []
        off = 0
        for n in self.sizes:
            self.offsets.append(off)
            off += -(-n // self.ALIGN) * self.ALIGN
        self.total = off
        dev = self.params[0].device
        self.flat = torch.zeros(self.total, dtype=torch.float32, device=dev)
        # the gradient buffer carries ALIGN extra floats behind the parameters: tail[0] = number of tasks of this rank whose
        # hypergradient is valid -- it rides along in the ONE all-reduce of the step (no second collective)
        self.gbuf = torch.zeros(self.total + self.ALIGN, dtype=torch.float32, device=dev)
        self.grad = self.gbuf[:self.total]
        self.tail = self.gbuf[self.total:]
        for p, n, off in zip(self.params, self.sizes, self.offsets):
            self.flat[off:off + n].copy_(p.data.reshape(-1))
            p.data = self.flat[off:off + n].view_as(p.data)
            p.grad = self.grad[off:off + n].view_as(p.data)
        from . import ops
        ops.weights_changed()           # (every parameter moved: cached weight images are stale)

    def zero_grad(self):
        self.gbuf.zero_()
        for p, n, off in zip(self.params, self.sizes, self.offsets):   # re-attach the views if something replaced .grad
            if p.grad is None or p.grad.data_ptr() != self.grad[off:off + n].data_ptr():
                p.grad = self.grad[off:off + n].view_as(p.data)

    def release_grads(self):
        """Before a backward pass: zero the flat gradient and detach the per-parameter views, so that autograd hands
        over each freshly computed gradient instead of launching one `grad += g` kernel per tensor (375 launches)."""
        self.gbuf.zero_()
        for p in self.params:
            p.grad = None

    def gather_grads(self):
        """After the backward pass: one multi-tensor copy of the per-parameter gradients into the flat buffer
        (parameters that took no part keep their zeros, cauchy_hypergradient.py:144-146) and re-attach the views."""
        views, grads = [], []
        for p, n, off in zip(self.params, self.sizes, self.offsets):
            v = self.grad[off:off + n].view_as(p.data)
            if p.grad is not None:
                views.append(v)
                grads.append(p.grad)
            p.grad = v
        if views:
            torch._foreach_copy_(views, grads)

    def clip_(self, max_norm: Optional[float], pre_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        """grad <- grad * pre_scale, then torch.nn.utils.clip_grad_norm_(max_norm) -- both factors applied in ONE pass over
        the flat buffer: same global 2-norm, one reduction, no host sync.  pre_scale (0-dim tensor) renormalises the
        meta-batch mean when tasks were dropped (1.0 exactly otherwise: bit-identical to no scaling)."""
        norm = torch.linalg.vector_norm(self.grad, dtype=torch.float64).float()   # fp64 accumulation
        factor = torch.ones((), dtype=torch.float32, device=self.grad.device)
        if pre_scale is not None:
            norm = norm * pre_scale
            factor = factor * pre_scale
        if max_norm is not None:
            factor = factor * torch.clamp(max_norm / (norm + 1e-6), max=1.0)
        self.grad.mul_(factor)
        return norm


class BucketedReducer:
    """The all-reduce of the flat meta-gradient, cut into buckets and started WHILE the backward pass is still running.

    Parameters are registered bottom-up (init projection, GNN blocks 0..L-1, readout, fc) and the backward pass produces their
    gradients top-down, so buckets are contiguous slices of the flat buffer taken from its END: as soon as every parameter of
    a bucket has its gradient (post-accumulate-grad hooks), the gradients are copied into the bucket's slice and
    `all_reduce(slice, async_op=True)` is issued -- NCCL runs it on its own stream behind the work queued so far, the
    remaining backward kernels keep the compute stream busy.  `finish()` reduces whatever is left (parameters that took no
    part never fire a hook: their slices hold zeros, cauchy_hypergradient.py:144-146) together with the tail slot, and waits.
    Element-wise the result equals ONE all-reduce of the whole buffer: replicas stay bit-identical."""

    def __init__(self, flat: FlatParams, bucket_floats: int = 6_000_000):
        self.flat = flat
        self.buckets = []                      # (lo, hi, [param indices]) from the end of the buffer
        idx, hi = [], flat.total
        for i in range(len(flat.params) - 1, -1, -1):
            idx.append(i)
            lo = flat.offsets[i]
            if hi - lo >= bucket_floats or i == 0:
                self.buckets.append((lo, hi, list(idx)))
                idx, hi = [], lo
        self._bucket_of = {}
        for b, (_, _, members) in enumerate(self.buckets):
            for i in members:
                self._bucket_of[i] = b
        self._pending, self._works, self._done = [], [], []
        self._armed = False
        self._handles = [p.register_post_accumulate_grad_hook(self._make_hook(i)) for i, p in enumerate(flat.params)]

    def _make_hook(self, i):
        def hook(param):
            if not self._armed:
                return
            b = self._bucket_of[i]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._launch(b)
        return hook

    def _copy_bucket(self, b):
        flat = self.flat
        views, grads = [], []
        for i in self.buckets[b][2]:
            p, n, off = flat.params[i], flat.sizes[i], flat.offsets[i]
            v = flat.grad[off:off + n].view_as(p.data)
            if p.grad is not None and p.grad.data_ptr() != v.data_ptr():
                views.append(v)
                grads.append(p.grad)
            p.grad = v
        if views:
            torch._foreach_copy_(views, grads)

    def _launch(self, b):
        self._copy_bucket(b)
        lo, hi, _ = self.buckets[b]
        self._works.append(dist.all_reduce(self.flat.gbuf[lo:hi], op=dist.ReduceOp.SUM, async_op=True))
        self._done[b] = True

    def prepare(self):
        """Before the backward pass (after FlatParams.release_grads)."""
        self._pending = [len(m) for _, _, m in self.buckets]
        self._done = [False] * len(self.buckets)
        self._works = []
        self._armed = True

    def finish(self):
        """After the backward pass and after the caller wrote the tail slot: reduce the rest, wait for everything."""
        self._armed = False
        rest = [b for b, d in enumerate(self._done) if not d]
        for b in rest:
            self._copy_bucket(b)
        flat = self.flat
        for b in rest:                                      # the unreduced buckets need not be adjacent: one call each
            blo, bhi, _ = self.buckets[b]
            self._works.append(dist.all_reduce(flat.gbuf[blo:bhi], op=dist.ReduceOp.SUM, async_op=True))
        self._works.append(dist.all_reduce(flat.tail, op=dist.ReduceOp.SUM, async_op=True))
        for w in self._works:
            w.wait()
        self._works = []
