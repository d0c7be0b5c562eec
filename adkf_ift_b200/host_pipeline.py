"""Background featurisation + collation of meta-training batches: the step BEFORE the hot path.

The reference feeds `train_loop` from worker PROCESSES: every worker reads its shard of the task files, draws the
support / query sample and featurises it, and the trainer pulls finished samples from a queue
(fs_mol/data/file_reader_iterable.py:21-60,132-172, fs_mol/data/dkt.py:156-207, fs_mol/data/fsmol_dataset.py:103-142).
Here one GPU step consumes `tasks_per_step` tasks in ~90 ms while drawing + DKT-batching + collating them takes ~150 ms of
Python on one core, so the same structure is kept, at the granularity of whole steps:

    worker w of P (process)   shard w of the task stream -> sample -> DKT batch -> collate T tasks -> HostTaskBatch (shared memory)
    staging thread (main)     round-robin over the workers (step k comes from worker k mod P: a deterministic order) ->
                              copy into a ring of PINNED host buffers (the GIL is released during the copies)
    task_batch.DevicePrefetcher   async H2D + CSR build of step k+1 on a side stream while step k computes
    train_step                loss read by the host one step late

`num_workers = 0` runs the SAME P shard generators in the calling process in the same round-robin order, so a run with and
without the pipeline sees identical batches (tests/test_gpu_trainer.py::test_train_loop_pipeline_identical_losses).
With several ranks, rank r uses shards r P .. r P + P - 1 of R P: every rank featurises only its own tasks.
"""
import itertools
import queue
import threading
from dataclasses import dataclass
from typing import Iterator, List, Optional

import torch

from .task_batch import HostTaskBatch, collate_tasks


@dataclass
class StreamSpec:
    """Everything a worker needs to produce its share of the meta-training batches (picklable)."""
    dataset: object
    data_fold: int
    support_size: int
    query_size: Optional[int]
    max_num_graphs: int
    use_numeric_labels: bool
    tasks_per_step: int          # tasks collated into one HostTaskBatch (this rank's share of the meta-batch)
    num_shards: int              # world_size x workers
    repeat: bool = True


class _ShardedDataset:
    """View of a dataset that yields every `num_shards`-th task starting at `shard`.  Datasets that can skip without
    loading expose `get_task_reading_iterable(fold, repeat=, shard=, num_shards=)`; for the others the stream is sliced."""

    def __init__(self, dataset, shard: int, num_shards: int):
        self._ds, self._shard, self._n = dataset, shard, num_shards

    def get_task_reading_iterable(self, data_fold, repeat: bool = False):
        if self._n == 1:
            return self._ds.get_task_reading_iterable(data_fold, repeat=repeat)
        try:
            return self._ds.get_task_reading_iterable(data_fold, repeat=repeat, shard=self._shard, num_shards=self._n)
        except TypeError:
            return itertools.islice(self._ds.get_task_reading_iterable(data_fold, repeat=repeat), self._shard, None, self._n)


def shard_batches(spec: StreamSpec, shard: int, pin: bool = False) -> Iterator[HostTaskBatch]:
    """The batches of one shard: get_dkt_task_sample_iterable (dkt.py:154-207) over the shard's tasks, `tasks_per_step`
    samples collated per batch.  A sample whose query does not fit one DKT batch is skipped (train_loop asserts it)."""
    from .dkt import get_dkt_task_sample_iterable
    it = get_dkt_task_sample_iterable(dataset=_ShardedDataset(spec.dataset, shard, spec.num_shards), data_fold=spec.data_fold,
                                      num_samples=1, max_num_graphs=spec.max_num_graphs, support_size=spec.support_size,
                                      query_size=spec.query_size, repeat=spec.repeat,
                                      filter_numeric_labels=spec.use_numeric_labels)
    group: List = []
    for smp in it:
        assert len(smp.batches) == 1, "a training sample must fit one batch (support + query <= batch_size)"
        group.append(smp.batches[0])
        if len(group) == spec.tasks_per_step:
            yield collate_tasks(group, spec.use_numeric_labels, pin=pin)
            group = []


def _worker_main(spec: StreamSpec, shard: int, out_q, stop_evt):
    torch.set_num_threads(1)
    try:
        for hb in shard_batches(spec, shard, pin=False):
            while not stop_evt.is_set():
                try:
                    out_q.put(hb, timeout=0.2)          # tensors travel through shared memory (torch.multiprocessing)
                    break
                except queue.Full:
                    continue
            if stop_evt.is_set():
                return
        out_q.put(None)
    except BaseException as e:  # noqa: BLE001  -- surfaced in the trainer process instead of a silent hang
        out_q.put(e)


class _PinnedRing:
    """`depth` reusable pinned copies of a HostTaskBatch (cudaHostAlloc per step would cost more than the copy)."""

    def __init__(self, depth: int):
        self._slots: List[Optional[dict]] = [None] * depth
        self._i = 0

    @staticmethod
    def _fit(buf: Optional[torch.Tensor], src: torch.Tensor) -> torch.Tensor:
        n = src.numel()
        if buf is None or buf.dtype != src.dtype or buf.numel() < n:
            buf = torch.empty(max(int(n * 1.25), 1), dtype=src.dtype)
            if torch.cuda.is_available():
                buf = buf.pin_memory()
        return buf

    def stage(self, hb: HostTaskBatch) -> HostTaskBatch:
        slot = self._slots[self._i] or {}
        out = {}
        for name in ("node_features", "node_to_graph", "fingerprints", "descriptors", "labels"):
            src = getattr(hb, name)
            slot[name] = self._fit(slot.get(name), src)
            dst = slot[name][:src.numel()].view(src.shape)
            dst.copy_(src)
            out[name] = dst
        adj = []
        for t, src in enumerate(hb.adjacency):
            key = f"adjacency{t}"
            slot[key] = self._fit(slot.get(key), src)
            dst = slot[key][:src.numel()].view(src.shape)
            dst.copy_(src)
            adj.append(dst)
        self._slots[self._i] = slot
        self._i = (self._i + 1) % len(self._slots)
        return HostTaskBatch(num_tasks=hb.num_tasks, n_sup=hb.n_sup, n_query=hb.n_query, num_graphs=hb.num_graphs,
                             num_nodes=hb.num_nodes, adjacency=adj, bool_labels=hb.bool_labels, max_degree=hb.max_degree, **out)


class HostBatchPipeline:
    """Iterator of pinned HostTaskBatches in a deterministic order (see the module docstring).

    ring_depth must cover the batches alive at once downstream: one being copied H2D, one staged, one in the hand-over
    queue and the one being filled (4)."""

    def __init__(self, spec: StreamSpec, num_workers: int = 4, emulate_workers: Optional[int] = None, rank: int = 0,
                 world: int = 1, queue_depth: int = 2, ring_depth: int = 5, mp_context: str = "spawn"):
        self._P = num_workers if num_workers > 0 else (emulate_workers or 1)
        if spec.num_shards != world * self._P:
            spec = StreamSpec(**{**spec.__dict__, "num_shards": world * self._P})
        self._spec, self._shards = spec, [rank * self._P + w for w in range(self._P)]
        self._procs, self._queues, self._stop = [], [], None
        self._ring = _PinnedRing(ring_depth)
        self._out: "queue.Queue" = queue.Queue(maxsize=max(1, ring_depth - 3))
        self._closed = False
        if num_workers > 0:
            import torch.multiprocessing as mp
            ctx = mp.get_context(mp_context)
            self._stop = ctx.Event()
            for s in self._shards:
                q = ctx.Queue(maxsize=queue_depth)
                p = ctx.Process(target=_worker_main, args=(spec, s, q, self._stop), daemon=True)
                p.start()
                self._procs.append(p)
                self._queues.append(q)
            self._thread = threading.Thread(target=self._stage_loop, daemon=True)
            self._thread.start()
        else:
            self._gens = [shard_batches(spec, s, pin=False) for s in self._shards]
            self._k = 0

    @property
    def num_workers(self) -> int:
        return len(self._procs)

    def _stage_loop(self):
        k = 0
        try:
            while not self._closed:
                item = self._queues[k % self._P].get()
                if item is None or isinstance(item, BaseException):
                    self._out.put(item)
                    return
                self._out.put(self._ring.stage(item))
                del item
                k += 1
        except BaseException as e:  # noqa: BLE001
            self._out.put(e)

    def __iter__(self):
        return self

    def __next__(self) -> HostTaskBatch:
        if self._procs:
            item = self._out.get()
            if item is None:
                raise StopIteration
            if isinstance(item, BaseException):
                raise RuntimeError("host pipeline worker failed") from item
            return item
        hb = next(self._gens[self._k % self._P])
        self._k += 1
        return self._ring.stage(hb)

    def close(self):
        self._closed = True
        if self._stop is not None:
            self._stop.set()
        for q in self._queues:
            try:
                while True:
                    q.get_nowait()
            except Exception:  # noqa: BLE001
                pass
        for p in self._procs:
            p.join(timeout=2)
            if p.is_alive():
                p.terminate()
        self._procs = []

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
