"""world_size-2 gloo test of the multi-GPU path's host logic: round-robin task sharding, one
all-reduce (sum) of the flat meta-gradient, identical clip + Adam on every rank.  The per-task hypergradients are
injected (the CUDA kernels cannot run here); what is checked is that 2 ranks x 2 tasks reproduce 1 rank x 4 tasks
and that the replicas stay bit-identical."""
import os
import socket

import torch
import torch.multiprocessing as mp

from adkf_ift_b200 import distributed, synthetic
from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
from adkf_ift_b200.feature_extractor import cli_default_config

NUM_TASKS = 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model():
    cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(1), clip_value=1.0, learning_rate=1e-3)
    m = ADKTModelTrainer(cfg)
    synthetic.deterministic_init_(m, seed=0)
    return m


def _task_grad(task, total):
    g = torch.Generator().manual_seed(100 + task)
    return torch.randn(total, generator=g) * 0.01


def _steps(model, my_tasks, n_steps=3):
    flat = model._ensure_flat()
    for step in range(n_steps):
        flat.zero_grad()
        for t in my_tasks:
            flat.grad += _task_grad(t + 10 * step, flat.total) / NUM_TASKS   # what run_on_task_batch accumulates
        model.reduce_and_update()
    return flat.flat.clone()


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    r, w = distributed.init_from_env(device="cpu")
    assert (r, w) == (rank, world) and distributed.world_size() == world
    model = _model()
    out[rank] = _steps(model, distributed.shard_tasks(NUM_TASKS, rank, world))
    torch.distributed.destroy_process_group()


def test_two_ranks_equal_one_rank():
    assert distributed.shard_tasks(5, 0, 2) == [0, 2, 4] and distributed.shard_tasks(5, 1, 2) == [1, 3]
    assert distributed.shard_tasks(16, 3, 8) == [3, 11]
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    single = _steps(_model(), list(range(NUM_TASKS)))
    assert torch.equal(out[0], out[1])                       # replicas bit-identical
    # == sequential accumulation over all tasks (Adam divides by sqrt(v): where |g| ~ 0 the summation order shows)
    assert torch.allclose(out[0], single, rtol=0, atol=1e-5)
    assert float((out[0] - single).abs().mean()) < 1e-8
    assert not torch.equal(single, _model()._ensure_flat().flat)  # the steps did move the weights


def test_flat_params_views_and_clip():
    model = _model()
    flat = model._ensure_flat()
    p0 = model.feature_extractor_params()[0]
    assert p0.data_ptr() == flat.flat.data_ptr() and p0.grad.data_ptr() == flat.grad.data_ptr()
    flat.grad.fill_(1.0)
    norm = flat.clip_(1.0)
    assert abs(float(norm) - flat.total ** 0.5) < 1e-2
    assert abs(float(torch.linalg.vector_norm(flat.grad, dtype=torch.float64)) - 1.0) < 1e-4
    assert float(p0.grad.flatten()[0]) == float(flat.grad[0])
    assert "fc.0.weight" in model.state_dict()               # names survive the flattening


def test_release_and_gather_grads_equal_view_accumulation():
    """train_step hands the meta-gradient over through FlatParams.release_grads / gather_grads: autograd assigns fresh
    gradient tensors (no per-parameter `+=` launch) and ONE multi-tensor copy moves them into the flat buffer.  Same flat
    gradient as accumulating into the views; parameters that took no part keep zeros; views are re-attached."""
    model = _model()
    flat = model._ensure_flat()
    params = model.feature_extractor_params()
    used = params[: len(params) // 2]                         # the second half stays unused in this "step"

    def loss():
        return sum((p * p).sum() * (i + 1) for i, p in enumerate(used)) * 1e-3

    flat.zero_grad()
    loss().backward()
    ref = flat.grad.clone()
    assert float(ref.abs().sum()) > 0
    flat.grad.fill_(7.0)                                      # stale content must not survive
    flat.release_grads()
    assert all(p.grad is None for p in params)
    loss().backward()
    assert used[0].grad.data_ptr() != flat.grad.data_ptr()    # autograd handed over a fresh tensor
    flat.gather_grads()
    assert torch.equal(flat.grad, ref)
    for p, n, off in zip(flat.params, flat.sizes, flat.offsets):
        assert p.grad.data_ptr() == flat.grad[off:off + n].data_ptr()
    assert float(params[-1].grad.abs().max()) == 0.0


def _bucket_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    distributed.init_from_env(device="cpu")
    model = _model()
    flat = model._ensure_flat()
    params = model.feature_extractor_params()
    used = [p for i, p in enumerate(params) if i % 7 != 3]          # some parameters take no part (their hooks never fire)

    def loss():
        return sum((p * p).sum() * (i + 1 + rank) for i, p in enumerate(used)) * 1e-3

    # (a) one all-reduce of the whole buffer
    flat.release_grads()
    loss().backward()
    flat.gather_grads()
    flat.tail[0] = 3.0 + rank
    distributed.allreduce_sum_(flat.gbuf)
    ref = flat.gbuf.clone()
    # (b) bucketed, launched from the gradient hooks during the backward pass
    red = distributed.BucketedReducer(flat, bucket_floats=200_000)
    assert len(red.buckets) > 3 and red.buckets[0][1] == flat.total and red.buckets[-1][0] == 0
    flat.release_grads()
    red.prepare()
    loss().backward()
    flat.tail[0] = 3.0 + rank
    red.finish()
    out[rank] = (torch.equal(flat.gbuf, ref), float(flat.tail[0]), len(red.buckets),
                 all(p.grad.data_ptr() == flat.grad[o:o + n].data_ptr() for p, n, o in zip(flat.params, flat.sizes, flat.offsets)))
    torch.distributed.destroy_process_group()


def test_bucketed_overlapped_allreduce_equals_single_allreduce():
    """distributed.BucketedReducer (buckets all-reduced from the gradient hooks while the backward pass runs, last layers
    first) gives bit-for-bit the buffer of ONE all-reduce, carries the valid-task count in the tail slot and re-attaches the
    gradient views -- world size 2, gloo."""
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_bucket_worker, args=(2, port, out), nprocs=2, join=True)
    for r in (0, 1):
        same, tail, nb, views = out[r]
        assert same and tail == 7.0 and views, (r, out[r])
