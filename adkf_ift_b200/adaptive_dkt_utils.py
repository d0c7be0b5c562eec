"""Task loops of ADKF-IFT on the batched B200 path -- surface of fs_mol/utils/adaptive_dkt_utils.py.

  run_on_batches(model, batches, batch_labels, batch_numeric_labels, train)   reference semantics (:70-116),
        one task, query batches looped; each step is the T=1 case of the batched kernels.
  run_on_task_batch(model, task_batch, train)                                  NEW: T tasks per launch.
        meta-test : embed -> sqdist -> fit -> predict                    (reference :87-104 per task)
        meta-train: embed -> sqdist -> fit -> IFT -> sqdist_bwd -> ONE backward through the extractor,
                    i.e. sum_t [dF_t/dtheta - v_t^T d2f_t/dphi dtheta] / T      (reference :361-404 + cauchy_hypergradient.py)
  ADKTModelTrainer                                                              (:208-458) with `train_step`.

Multi-GPU: one process per GPU; rank r takes tasks r::R of the meta-batch, the flat fp32 meta-gradient is
all-reduced (sum) once per outer step over NCCL, then every rank applies the identical clip + Adam update
(SURVEY.md 8(e)).  The reference has no distributed code.
"""
import math
import os
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import gp as gpk
from . import ops
from . import distributed
from .adaptive_dkt import ADKTModel, ADKTModelTrainerConfig
from .distributed import FlatParams
from .dkt import get_dkt_batcher, get_dkt_task_sample_iterable, task_sample_to_dkt_task_sample, torchify
from .eval_utils import DataFold, iter_eval_draws, make_eval_result, write_task_results
from .metrics import (avg_metrics_over_tasks, avg_numeric_metrics_over_tasks, compute_binary_task_metrics,
                      compute_numeric_task_metrics)
from .task_batch import DeviceTaskBatch, HostTaskBatch, collate_tasks, to_device


def fit_gpytorch_scipy(mll, **_ignored):
    """Name kept for call-site compatibility (adaptive_dkt_utils.py:91): runs the on-device Newton fit."""
    return mll.fit()


@dataclass
class TaskBatchResult:
    """Per-task outputs of one batched launch (device tensors; nothing is synced to the host here)."""
    phi: torch.Tensor                 # [T,3]
    status: torch.Tensor              # [T] int32
    aux: torch.Tensor                 # [T,32]
    mean: Optional[torch.Tensor] = None       # [rows] predictive mean (query rows valid)
    variance: Optional[torch.Tensor] = None
    preds: Optional[torch.Tensor] = None      # sigmoid(mean) | mean, query rows only, task-major
    query_mask: Optional[torch.Tensor] = None
    outer_loss: Optional[torch.Tensor] = None  # [T] F_t (training)
    loss_per_sample: Optional[torch.Tensor] = None


def _query_mask(layout: ops.TaskLayout, device) -> torch.Tensor:
    m = np.zeros(layout.total_rows, dtype=bool)
    off = 0
    for ns, nr in zip(layout.n_sup_host, layout.n_rows_host):
        m[off + ns:off + nr] = True
        off += nr
    return torch.from_numpy(m).to(device, non_blocking=True)


def _fit_or_fixed(g: gpk.BatchedExactGP, phi: Optional[torch.Tensor], max_iter: int, gtol: float) -> gpk.FitResult:
    if phi is None:
        return g.fit(max_iter=max_iter, gtol=gtol)
    init = g.init()   # median lengthscale -> prior loc, exactly as a fit would see it
    return g.evaluate(phi.to(init.phi.device, torch.float32).view(-1, 3).contiguous(), init.aux)


def run_on_task_batch(model: ADKTModel, tb: DeviceTaskBatch, train: bool = False, grad_scale: Optional[float] = None,
                      fit_max_iter: int = 50, fit_gtol: float = 2e-6,
                      phi: Optional[torch.Tensor] = None) -> TaskBatchResult:
    """All T tasks of `tb` in one pass.  train=True accumulates the IFT hypergradient, scaled by `grad_scale`
    (default 1/T = the reference's `/ tasks_per_batch`), into `.grad` of the feature-extractor parameters.
    `phi` [T,3] skips the inner fit and evaluates at the given raw GP parameters (fixed-phi parity checks)."""
    cfg = model.config
    if train:
        model.train()
        Z = model.embed(tb)
        g = gpk.BatchedExactGP(Z, tb.labels, tb.layout, cfg.gp_kernel, cfg.use_numeric_labels,
                               cfg.use_lengthscale_prior, qq_block=True)
        fit = _fit_or_fixed(g, phi, fit_max_iter, fit_gtol)
        dZ, aux, status = g.hypergradient(fit, mode=1 if cfg.ignore_grad_correction else 0)
        scale = (1.0 / tb.num_tasks) if grad_scale is None else grad_scale
        Z.backward(dZ * scale)
        F_t = aux[:, ops.AUX_FOUTER]
        n_q = ops.host_to_device([float(r - s) for r, s in zip(tb.layout.n_rows_host, tb.layout.n_sup_host)], torch.float32, Z.device)
        return TaskBatchResult(phi=fit.phi, status=status | fit.status, aux=aux, outer_loss=F_t,
                               loss_per_sample=F_t / n_q)
    model.eval()
    with torch.no_grad():
        Z = model.embed(tb)
        g = gpk.BatchedExactGP(Z, tb.labels, tb.layout, cfg.gp_kernel, cfg.use_numeric_labels,
                               cfg.use_lengthscale_prior, qq_block=False)
        fit = _fit_or_fixed(g, phi, fit_max_iter, fit_gtol)
        mean, var, st = g.predict(fit.phi)
        qm = _query_mask(tb.layout, Z.device)
        mq = mean[qm]
        preds = mq if cfg.use_numeric_labels else torch.sigmoid(mq)
    return TaskBatchResult(phi=fit.phi, status=fit.status | st, aux=fit.aux, mean=mean, variance=var, preds=preds,
                           query_mask=qm)


def run_on_batches_raw(model: ADKTModel, batches: Sequence, batch_labels: Sequence, batch_numeric_labels: Sequence,
                       train: bool = False):
    """Reference loop (adaptive_dkt_utils.py:70-104) over the query batches of ONE task.  The reference re-fits the GP for
    every query batch on the same support set; the fit is deterministic, so it is done once.
    Returns (predictions, labels) as numpy ((None, None) when train=True)."""
    if train:
        assert len(batches) == 1
    preds: List[np.ndarray] = []
    labels: List[np.ndarray] = []
    fitted = False
    for feats, bl, bnl in zip(batches, batch_labels, batch_numeric_labels):
        if not fitted:
            model.train()
            _ = model(feats, train_loss=True)
            fit_gpytorch_scipy(model.mll)
            fitted = True
        if not train:
            model.eval()
            with torch.no_grad():
                out = model(feats, train_loss=None)
            if model.config.use_numeric_labels:
                preds.append(out.mean.detach().cpu().numpy())
                labels.append(torch.as_tensor(bnl).detach().cpu().numpy())
            else:
                preds.append(torch.sigmoid(out.mean).detach().cpu().numpy())
                labels.append(torch.as_tensor(bl).detach().cpu().numpy())
    if train:
        return None, None
    return np.concatenate(preds, axis=0), np.concatenate(labels, axis=0)


def run_on_batches(model: ADKTModel, batches: Sequence, batch_labels: Sequence, batch_numeric_labels: Sequence,
                   train: bool = False):
    """adaptive_dkt_utils.py:70-116: returns the task's BinaryEvalMetrics / NumericEvalMetrics (None when train=True)."""
    preds, labels = run_on_batches_raw(model, batches, batch_labels, batch_numeric_labels, train)
    if train:
        return None
    if model.config.use_numeric_labels:
        return compute_numeric_task_metrics(predictions=preds, labels=labels)
    return compute_binary_task_metrics(predictions=preds, labels=labels)


def evaluate_adkt_model(model: ADKTModel, dataset, support_sizes: Sequence[int] = (16, 128), num_samples: int = 5, seed: int = 0,
                        batch_size: int = 320, query_size: Optional[int] = None, data_fold: DataFold = DataFold.TEST,
                        save_dir: Optional[str] = None, tasks_per_launch: int = 32) -> Dict[str, List]:
    """adaptive_dkt_utils.py:119-175 + eval_model (test_utils.py:236-350): the same (task, support size, run) draws in the
    same order with the same seeds, the same result objects and CSV files -- but the draws are evaluated `tasks_per_launch`
    at a time through the batched path (one extractor forward, one fit launch, one predict launch per group).  A draw whose
    query set needs more than one batch of `batch_size` graphs goes through `run_on_batches`."""
    numeric = model.config.use_numeric_labels
    batcher = get_dkt_batcher(max_num_graphs=batch_size)
    results: Dict[str, List] = {}
    pending: List = []                                    # (draw, host DKT sample)

    def flush():
        singles = [(d, h) for d, h in pending if len(h.batches) == 1]
        for d, h in pending:
            if len(h.batches) != 1:
                dev = torchify(h, model.device)
                m = run_on_batches(model, dev.batches, dev.batch_labels, dev.batch_numeric_labels, train=False)
                results[d.task_name].append((d, m))
        if singles:
            tb = to_device(collate_tasks([h.batches[0] for _, h in singles], numeric), model.device)
            res = run_on_task_batch(model, tb, train=False)
            preds = res.preds.detach().cpu().numpy()
            off = 0
            for (d, h), n_q in zip(singles, [r - s_ for r, s_ in zip(tb.layout.n_rows_host, tb.layout.n_sup_host)]):
                p = preds[off:off + n_q]
                off += n_q
                y = h.batch_numeric_labels[0] if numeric else h.batch_labels[0]
                m = compute_numeric_task_metrics(p, np.asarray(y)) if numeric else compute_binary_task_metrics(p, np.asarray(y))
                results[d.task_name].append((d, m))
        pending.clear()

    order: Dict[str, List] = {}
    for task, draws in iter_eval_draws(dataset, list(support_sizes), num_samples, 0.0, query_size, data_fold, seed, numeric):
        results[task.name] = []
        order[task.name] = draws
        for d in draws:
            pending.append((d, task_sample_to_dkt_task_sample(d.sample, batcher, numeric)))
            if len(pending) >= tasks_per_launch:
                flush()
    flush()
    out: Dict[str, List] = {}
    for name, draws in order.items():
        by_draw = {id(d): m for d, m in results[name]}
        out[name] = [make_eval_result(d, by_draw[id(d)], numeric) for d in draws]     # the reference's order
        write_task_results(save_dir, name, out[name], numeric)
    return out


def validate_by_finetuning_on_tasks(model: ADKTModel, dataset, seed: int = 0, aml_run=None,
                                    metric_to_use: str = "avg_precision") -> float:
    """adaptive_dkt_utils.py:178-211: mean over the validation tasks of the requested metric."""
    task_results = evaluate_adkt_model(model, dataset, support_sizes=model.config.validation_support_set_sizes,
                                       num_samples=model.config.validation_num_samples, seed=seed,
                                       batch_size=model.config.batch_size, query_size=model.config.validation_query_set_size,
                                       data_fold=DataFold.VALIDATION)
    mean_metrics = (avg_numeric_metrics_over_tasks if model.config.use_numeric_labels else avg_metrics_over_tasks)(task_results)
    if aml_run is not None:
        for name, (mean, _) in mean_metrics.items():
            aml_run.log(f"valid_task_test_{name}", float(mean))
    return float(mean_metrics[metric_to_use][0])


def linear_warmup(cur_step: int, warmup_steps: int = 0) -> float:      # fs_mol/utils/torch_utils.py
    if cur_step >= warmup_steps:
        return 1.0
    return cur_step / warmup_steps


# ----------------------------------------------------------------------------------------------------------
# trainer
# ----------------------------------------------------------------------------------------------------------
class ADKTModelTrainer(ADKTModel):
    def __init__(self, config: ADKTModelTrainerConfig):
        super().__init__(config)
        self.config = config
        self.optimizer = torch.optim.Adam(self.feature_extractor_params(), config.learning_rate)
        self.lr_scheduler = None
        self._flat: Optional[FlatParams] = None
        self._step = 0

    # ---- checkpoints (adaptive_dkt_utils.py:221-259, 305-327) -----------------------------------------
    def get_model_state(self) -> Dict[str, Any]:
        return {"model_config": self.config, "model_state_dict": self.state_dict()}

    def save_model(self, path: str, optimizer: Optional[torch.optim.Optimizer] = None, epoch: Optional[int] = None):
        data = self.get_model_state()
        if optimizer is not None:
            data["optimizer_state_dict"] = optimizer.state_dict()
        if epoch is not None:
            data["epoch"] = epoch
        torch.save(data, path)

    def load_model_weights(self, path: str, quiet: bool = False, device: Optional[torch.device] = None):
        """Copies tensors by key.  Reference checkpoints also carry gpytorch's `gp_*` / `mll.*` entries (several
        aliases of the same three tensors plus prior buffers): they are re-initialised per task anyway, so keys
        this model does not own are skipped."""
        ckpt = torch.load(path, map_location=device, weights_only=False)
        own = self.state_dict()
        skipped = []
        for name, value in ckpt["model_state_dict"].items():
            if isinstance(value, torch.nn.Parameter):
                value = value.data
            if name in own and own[name].shape == value.shape:
                own[name].copy_(value)
            elif name.startswith("gp_") or name.startswith("mll."):
                skipped.append(name)
            else:
                raise KeyError(f"checkpoint tensor {name} {tuple(value.shape)} has no counterpart in the model")
        return skipped

    @classmethod
    def build_from_model_file(cls, model_file: str, config_overrides: Dict[str, Any] = {}, quiet: bool = False,
                              device: Optional[torch.device] = None) -> "ADKTModelTrainer":
        ckpt = torch.load(model_file, map_location=device, weights_only=False)
        model = cls(ckpt["model_config"])
        if device is not None:
            model.to(device)
        model.load_model_weights(model_file, quiet=quiet, device=device)
        return model

    def load_model_gnn_weights(self, path: str, device: Optional[torch.device] = None):
        """adaptive_dkt_utils.py:261-303: initialise the graph feature extractor from a pre-trained GNNMultitask checkpoint
        (its names carry no `graph_feature_extractor.` prefix and call the readout `readout.`), then train the loaded part at
        a tenth of the learning rate with a 100-step linear warm-up."""
        from functools import partial
        gnn_state = torch.load(path, map_location=device, weights_only=False)["model_state_dict"]
        prefix = "graph_feature_extractor."
        gnn_params, other_params = [], []
        own_params = dict(self.named_parameters())
        with torch.no_grad():
            for name, tensor in self.state_dict().items():
                if name.startswith(prefix) and "final_norm_layer" not in name:
                    generic = name[len(prefix):]
                    if generic.startswith("readout_layer."):
                        generic = "readout" + generic[len("readout_layer"):]
                    tensor.copy_(gnn_state[generic])
                    if name in own_params:
                        gnn_params.append(own_params[name])
                elif name in own_params and not name.startswith("gp_"):
                    other_params.append(own_params[name])
        self._flat = None                                   # the optimiser below owns the (unflattened) parameters
        self.optimizer = torch.optim.Adam([{"params": other_params, "lr": self.config.learning_rate},
                                           {"params": gnn_params, "lr": self.config.learning_rate / 10}])
        self.lr_scheduler = torch.optim.lr_scheduler.LambdaLR(
            optimizer=self.optimizer, lr_lambda=[partial(linear_warmup, warmup_steps=0), partial(linear_warmup, warmup_steps=100)])
        self._grouped_optimizer = True

    def train_loop(self, out_dir: str, dataset, device: torch.device, aml_run=None):
        """adaptive_dkt_utils.py:329-458 on the batched path: every outer step draws `tasks_per_batch` task samples from
        the (repeating) training iterable, runs fit + IFT hypergradient of all of them in one pass, all-reduces, clips and
        steps; validates by fine-tuning every `validate_every_num_steps` and keeps `best_validation.pt` / `fully_trained.pt`.
        With several ranks each rank draws from its own shard of the iterable (`distributed.shard_tasks`)."""
        import logging
        log = logging.getLogger(__name__)
        os.makedirs(out_dir, exist_ok=True)
        self.to(device)
        cfg = self.config
        self.save_model(os.path.join(out_dir, "best_validation.pt"))
        it = iter(get_dkt_task_sample_iterable(dataset=dataset, data_fold=DataFold.TRAIN, num_samples=1,
                                               max_num_graphs=cfg.batch_size, support_size=cfg.support_set_size,
                                               query_size=cfg.query_set_size, repeat=True,
                                               filter_numeric_labels=cfg.use_numeric_labels))
        world, rank = distributed.world_size(), distributed.rank()
        best, metric = -np.inf, ("r2" if cfg.use_numeric_labels else "avg_precision")
        losses: List[float] = []
        for step in range(1, cfg.num_train_steps + 1):
            drawn = [next(it) for _ in range(cfg.tasks_per_batch)]
            mine = [drawn[i] for i in distributed.shard_tasks(len(drawn), rank, world)]
            for smp in mine:
                assert len(smp.batches) == 1, "a training sample must fit one batch (support + query <= batch_size)"
            tb = to_device(collate_tasks([smp.batches[0] for smp in mine], cfg.use_numeric_labels), device)
            res = self.train_step(tb, global_num_tasks=len(drawn))
            losses.append(float(res.loss_per_sample.mean().item()))
            if aml_run is not None:
                aml_run.log("loss", losses[-1])
            if step % cfg.validate_every_num_steps == 0:
                valid = validate_by_finetuning_on_tasks(self, dataset, aml_run=aml_run, metric_to_use=metric)
                log.info(f"Validated at train step [{step}/{cfg.num_train_steps}], valid {metric}: {valid:.3f}, "
                         f"mean loss {np.mean(losses[-cfg.validate_every_num_steps:]):.5f}")
                if valid > best:
                    best = valid
                    if rank == 0:
                        self.save_model(os.path.join(out_dir, "best_validation.pt"))
        if rank == 0:
            self.save_model(os.path.join(out_dir, "fully_trained.pt"))
        return losses

    # ---- one outer step -------------------------------------------------------------------------------
    def _ensure_flat(self) -> FlatParams:
        if self._flat is None:
            params = self.feature_extractor_params()
            self._flat = FlatParams(params)
            if getattr(self, "_grouped_optimizer", False):
                # load_model_gnn_weights built per-group learning rates: keep the groups, re-point them at the flat views
                groups = [{"params": list(g["params"]), "lr": g["lr"]} for g in self.optimizer.param_groups]
                self.optimizer = torch.optim.Adam(groups, fused=params[0].is_cuda)
                if self.lr_scheduler is not None:
                    from functools import partial
                    self.lr_scheduler = torch.optim.lr_scheduler.LambdaLR(
                        optimizer=self.optimizer,
                        lr_lambda=[partial(linear_warmup, warmup_steps=0), partial(linear_warmup, warmup_steps=100)])
            else:
                # one learning rate for everything: Adam runs on the FLAT buffer as a single tensor (one launch at HBM speed
                # instead of 11 multi-tensor launches over 375 small tensors: 1.07 -> 0.15 ms); element-wise identical, the
                # zero padding between tensors stays zero (g = m = v = 0 -> update 0)
                fp = torch.nn.Parameter(self._flat.flat)
                fp.grad = self._flat.grad
                self._flat_param = [fp]                      # kept in a list: not a registered parameter, state_dict unchanged
                self.optimizer = torch.optim.Adam([fp], self.config.learning_rate, fused=params[0].is_cuda)
        return self._flat

    def reduce_and_update(self):
        """All-reduce (sum) of the flat meta-gradient over the ranks, then clip + Adam -- identical on every rank, so
        the replicas stay bit-identical (adaptive_dkt_utils.py:406-413)."""
        flat = self._ensure_flat()
        distributed.allreduce_sum_(flat.grad)
        if self.config.clip_value is not None:
            flat.clip_(self.config.clip_value)
        self.optimizer.step()
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        self._step += 1

    def train_step(self, tb: DeviceTaskBatch, global_num_tasks: Optional[int] = None) -> TaskBatchResult:
        """One outer step on this rank's share of the meta-batch (adaptive_dkt_utils.py:352-413): hypergradients of
        the local tasks / global task count, all-reduce, clip, Adam."""
        flat = self._ensure_flat()
        flat.release_grads()
        total = global_num_tasks if global_num_tasks is not None else tb.num_tasks * distributed.world_size()
        res = run_on_task_batch(self, tb, train=True, grad_scale=1.0 / total)
        flat.gather_grads()
        self.reduce_and_update()
        return res
