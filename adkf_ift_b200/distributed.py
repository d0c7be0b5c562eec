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
