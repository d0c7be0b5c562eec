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
    if getattr(cfg, "use_ard", False):
        return _run_on_task_batch_ard(model, tb, train, grad_scale, phi)
    if train:
        if not model.training:                      # (Module.train() walks and re-sets all ~560 submodules: 3 ms of host time per step)
            model.train()
        with ops.nvtx_range("adkf/embed"):
            Z = model.embed(tb)
        with ops.nvtx_range("adkf/gram"):
            g = gpk.BatchedExactGP(Z, tb.labels, tb.layout, cfg.gp_kernel, cfg.use_numeric_labels,
                                   cfg.use_lengthscale_prior, qq_block=True)
        with ops.nvtx_range("adkf/fit"):
            fit = _fit_or_fixed(g, phi, fit_max_iter, fit_gtol)
        with ops.nvtx_range("adkf/ift"):
            dZ, aux, status = g.hypergradient(fit, mode=1 if cfg.ignore_grad_correction else 0)
        scale = (1.0 / tb.num_tasks) if grad_scale is None else grad_scale
        with ops.nvtx_range("adkf/backward"):
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


def _run_on_task_batch_ard(model: ADKTModel, tb: DeviceTaskBatch, train: bool, grad_scale: Optional[float],
                           phi: Optional[torch.Tensor]) -> TaskBatchResult:
    """--use-ard (adaptive_dkt_train.py:107,151): d + 2 inner parameters per task, batched L-BFGS fit and matrix-free (CG)
    inverse-Hessian-vector product on the device (gp_ard.py) instead of the reference's dense 2050 x 2050 Hessian."""
    from .gp_ard import ARDFit, BatchedARDExactGP
    cfg = model.config
    model.train(train)
    with torch.set_grad_enabled(train):
        Z = model.embed(tb)
    g = BatchedARDExactGP(Z.detach(), tb.labels, tb.layout, cfg.gp_kernel, cfg.use_numeric_labels, cfg.use_lengthscale_prior)
    if phi is None:
        fit = g.fit()
    else:
        phi0, l0, loc = g.init_phi()
        p = phi.to(Z.device, torch.float32).view(tb.num_tasks, -1).contiguous()
        f, grad, st = g.value_and_grad(p, loc)
        fit = ARDFit(p, f, grad, loc, l0, 0, st)
    aux = torch.zeros(tb.num_tasks, ops.AUX, dtype=torch.float32, device=Z.device)
    if train:
        dZ, info = g.hypergradient(fit, ignore_grad_correction=cfg.ignore_grad_correction)
        scale = (1.0 / tb.num_tasks) if grad_scale is None else grad_scale
        Z.backward(dZ * scale)
        F_t = info["F"]
        n_q = ops.host_to_device([float(r - s) for r, s in zip(tb.layout.n_rows_host, tb.layout.n_sup_host)], torch.float32, Z.device)
        return TaskBatchResult(phi=fit.phi, status=fit.status | info["status"], aux=aux, outer_loss=F_t, loss_per_sample=F_t / n_q)
    mean, var, st = g.predict(fit.phi)
    qm = _query_mask(tb.layout, Z.device)
    mq = mean[qm]
    preds = mq if cfg.use_numeric_labels else torch.sigmoid(mq)
    return TaskBatchResult(phi=fit.phi, status=fit.status | st, aux=aux, mean=mean, variance=var, preds=preds, query_mask=qm)


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
            data["optimizer_state_dict"] = self._export_optimizer_state(optimizer)
        if epoch is not None:
            data["epoch"] = epoch
        torch.save(data, path)

    # ---- optimizer state across the flat-buffer optimiser ---------------------------------------------------
    def _export_optimizer_state(self, optimizer: torch.optim.Optimizer) -> Dict[str, Any]:
        """`optimizer.state_dict()` in the layout of the reference's `Adam(self.feature_extractor_params())` (one entry per
        parameter, registration order), also when the step runs on the flat buffer as ONE tensor: the moments are sliced
        back to the named parameters through the FlatParams offsets, so a checkpoint resumes in a fresh trainer -- or in
        the reference's own."""
        flat_param = getattr(self, "_flat_param", None)
        if self._flat is None or flat_param is None or optimizer is not self.optimizer:
            return optimizer.state_dict()
        sd = optimizer.state_dict()
        st = sd["state"].get(0)
        group = {k: v for k, v in sd["param_groups"][0].items() if k != "params"}
        flat = self._flat
        state = {}
        if st is not None:
            for i, (p, n, off) in enumerate(zip(flat.params, flat.sizes, flat.offsets)):
                state[i] = {"step": st["step"].detach().clone().cpu() if torch.is_tensor(st["step"]) else st["step"],
                            "exp_avg": st["exp_avg"][off:off + n].view_as(p).clone(),
                            "exp_avg_sq": st["exp_avg_sq"][off:off + n].view_as(p).clone()}
        group["params"] = list(range(len(flat.params)))
        group["fused"] = None
        return {"state": state, "param_groups": [group]}

    def _import_optimizer_state(self, sd: Dict[str, Any]):
        """Inverse of _export_optimizer_state: accepts the per-parameter layout (also a reference checkpoint's) and loads it
        into whichever optimiser is live -- the per-parameter one, or the flat one (moments scattered through the offsets)."""
        flat_param = getattr(self, "_flat_param", None)
        if self._flat is None or flat_param is None:
            self.optimizer.load_state_dict(sd)
            return
        flat = self._flat
        if len(sd["state"]) == 0:
            return
        exp_avg = torch.zeros_like(flat.flat)
        exp_avg_sq = torch.zeros_like(flat.flat)
        step = None
        for i, (p, n, off) in enumerate(zip(flat.params, flat.sizes, flat.offsets)):
            st = sd["state"].get(i)
            if st is None:
                continue
            exp_avg[off:off + n].copy_(st["exp_avg"].reshape(-1))
            exp_avg_sq[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
            step = st["step"]
        fp = flat_param[0]
        step_t = torch.as_tensor(float(step), dtype=torch.float32, device=fp.device if self.optimizer.param_groups[0].get("fused") else "cpu")
        self.optimizer.state[fp] = {"step": step_t, "exp_avg": exp_avg, "exp_avg_sq": exp_avg_sq}
        for k, v in sd["param_groups"][0].items():
            if k not in ("params", "fused", "foreach", "capturable", "differentiable"):
                self.optimizer.param_groups[0][k] = v

    def load_model_weights(self, path: str, quiet: bool = False, device: Optional[torch.device] = None):
        """Copies tensors by key.  Reference checkpoints also carry gpytorch's `gp_*` / `mll.*` entries (several
        aliases of the same three tensors plus prior buffers): they are re-initialised per task anyway, so keys
        this model does not own are skipped.  An `optimizer_state_dict` is restored as well (the reference tries to,
        adaptive_dkt_utils.py:255-258)."""
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
        opt_sd = ckpt.get("optimizer_state_dict")
        if opt_sd is not None:
            try:
                self._import_optimizer_state(opt_sd)
            except (ValueError, KeyError, RuntimeError) as e:       # a different parameter set: keep the fresh optimiser
                if not quiet:
                    import logging
                    logging.getLogger(__name__).warning(f"optimizer state of {path} not restored: {e}")
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

    def train_loop(self, out_dir: str, dataset, device: torch.device, aml_run=None, num_workers: Optional[int] = None,
                   emulate_workers: Optional[int] = None):
        """adaptive_dkt_utils.py:329-458 on the batched path: every outer step takes `tasks_per_batch` task samples from the
        (repeating) training stream, runs fit + IFT hypergradient of all of them in one pass, all-reduces, clips and steps;
        validates by fine-tuning every `validate_every_num_steps` and keeps `best_validation.pt` / `fully_trained.pt`.

        The stream is produced as the reference produces it -- by worker processes (host_pipeline.HostBatchPipeline:
        featurisation + collation off the trainer's core, pinned staging, async H2D one step ahead) -- and the loss of a step
        is read by the host one step late, so the GPU never waits for Python.  `num_workers=0` runs the same shard
        generators inline (identical batches, hence identical losses).  With R ranks every rank takes tasks_per_batch / R
        tasks per step from its own shards of the stream."""
        import logging
        from .host_pipeline import HostBatchPipeline, StreamSpec
        from .task_batch import DevicePrefetcher
        log = logging.getLogger(__name__)
        device = torch.device(device)
        if device.type == "cuda":
            torch.cuda.set_device(device)              # the C-ABI launches on the CURRENT device's stream
        cfg = self.config
        world, rank = distributed.world_size(), distributed.rank()
        if cfg.tasks_per_batch % world != 0:
            raise ValueError(f"tasks_per_batch ({cfg.tasks_per_batch}) must be a multiple of the number of ranks ({world}): "
                             "every rank takes an equal share of the meta-batch")
        os.makedirs(out_dir, exist_ok=True)
        self.to(device)
        if rank == 0:
            self.save_model(os.path.join(out_dir, "best_validation.pt"))
        if world > 1:
            torch.distributed.barrier()
        if num_workers is None:
            num_workers = max(1, min(4, (os.cpu_count() or 2) // max(world, 1) - 1))
        spec = StreamSpec(dataset=dataset, data_fold=int(DataFold.TRAIN), support_size=cfg.support_set_size,
                          query_size=cfg.query_set_size, max_num_graphs=cfg.batch_size,
                          use_numeric_labels=cfg.use_numeric_labels, tasks_per_step=cfg.tasks_per_batch // world,
                          num_shards=1, repeat=True)
        pipe = HostBatchPipeline(spec, num_workers=num_workers, emulate_workers=emulate_workers, rank=rank, world=world)
        best, metric = -np.inf, ("r2" if cfg.use_numeric_labels else "avg_precision")
        losses: List[float] = []
        readback = _LossReadback(device)
        try:
            batches = DevicePrefetcher(pipe, device) if device.type == "cuda" else (to_device(hb, device) for hb in pipe)
            for step in range(1, cfg.num_train_steps + 1):
                tb = next(batches)
                res = self.train_step(tb, global_num_tasks=cfg.tasks_per_batch)
                for rec in readback.push(res):
                    self._log_step(rec, losses, aml_run, log)
                if step % cfg.validate_every_num_steps == 0:
                    for rec in readback.flush():
                        self._log_step(rec, losses, aml_run, log)
                    valid = validate_by_finetuning_on_tasks(self, dataset, aml_run=aml_run, metric_to_use=metric)
                    log.info(f"Validated at train step [{step}/{cfg.num_train_steps}], valid {metric}: {valid:.3f}, "
                             f"mean loss {np.mean(losses[-cfg.validate_every_num_steps:]):.5f}")
                    if valid > best:
                        best = valid
                        if rank == 0:
                            self.save_model(os.path.join(out_dir, "best_validation.pt"), optimizer=self.optimizer, epoch=step)
            for rec in readback.flush():
                self._log_step(rec, losses, aml_run, log)
        finally:
            pipe.close()
        if rank == 0:
            self.save_model(os.path.join(out_dir, "fully_trained.pt"), optimizer=self.optimizer, epoch=cfg.num_train_steps)
        return losses

    @staticmethod
    def _log_step(rec, losses, aml_run, log):
        loss, dropped, not_converged = rec
        losses.append(loss)
        if aml_run is not None:
            aml_run.log("loss", loss)
        if dropped:
            log.warning(f"step {len(losses)}: {dropped} task(s) dropped from the meta-batch (covariance not positive definite after "
                        "the jitter ladder, or singular inner Hessian); the mean was renormalised over the remaining tasks")
        if not_converged:
            log.info(f"step {len(losses)}: inner fit of {not_converged} task(s) stopped at max_iter")

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

    def reduce_and_update(self, global_num_tasks: Optional[int] = None, already_reduced: bool = False):
        """All-reduce (sum) of the flat meta-gradient over the ranks, then clip + Adam -- identical on every rank, so
        the replicas stay bit-identical (adaptive_dkt_utils.py:406-413).  The tail slot of the gradient buffer carries the
        number of VALID tasks through the same all-reduce; when tasks were dropped the mean is renormalised over the rest."""
        flat = self._ensure_flat()
        if not already_reduced:
            with ops.nvtx_range("adkf/all-reduce"):
                distributed.allreduce_sum_(flat.gbuf)
        pre_scale = None
        if global_num_tasks is not None:
            pre_scale = float(global_num_tasks) / torch.clamp(flat.tail[0], min=1.0)     # == 1.0 exactly when nothing was dropped
        with ops.nvtx_range("adkf/clip+adam"):
            if self.config.clip_value is not None or pre_scale is not None:
                flat.clip_(self.config.clip_value, pre_scale)
            self.optimizer.step()
            ops.weights_changed()       # Adam wrote the parameters through the flat buffer: their version counters did not move
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        self._step += 1

    def train_step(self, tb: DeviceTaskBatch, global_num_tasks: Optional[int] = None) -> TaskBatchResult:
        """One outer step on this rank's share of the meta-batch (adaptive_dkt_utils.py:352-413): hypergradients of
        the local tasks / global task count, all-reduce, clip, Adam.  Tasks whose GP failed (ADKF_ST_NOT_PD /
        ADKF_ST_HESS_DAMPED: zero gradient, NaN loss) are left out of the mean -- the reference would raise."""
        flat = self._ensure_flat()
        flat.release_grads()
        total = global_num_tasks if global_num_tasks is not None else tb.num_tasks * distributed.world_size()
        overlap = distributed.world_size() > 1 and getattr(self, "overlap_allreduce", True)
        if overlap:
            if getattr(self, "_reducer", None) is None or self._reducer.flat is not flat:
                self._reducer = distributed.BucketedReducer(flat)
            self._reducer.prepare()             # buckets are all-reduced behind the backward pass, last layers first
        res = run_on_task_batch(self, tb, train=True, grad_scale=1.0 / total)
        valid = ((res.status & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED)) == 0).sum().to(torch.float32).reshape(1)
        if overlap:
            flat.tail[0:1].copy_(valid)
            with ops.nvtx_range("adkf/all-reduce (tail of the bucketed reduction)"):
                self._reducer.finish()
        else:
            flat.gather_grads()
            flat.tail[0:1].copy_(valid)
        self.reduce_and_update(total, already_reduced=overlap)
        return res


class _LossReadback:
    """Per-step (mean loss per sample over the valid tasks, #dropped tasks, #unconverged fits) read by the host ONE STEP LATE:
    an async copy into a pinned slot + an event per step, so that the host never drains the GPU inside the loop (a blocking
    `.item()` per step costs a pipeline refill, ~3 ms at C2)."""

    def __init__(self, device, depth: int = 4):
        self._cuda = torch.device(device).type == "cuda"
        self._pin = torch.empty(depth, 3, dtype=torch.float32)
        if self._cuda:
            self._pin = self._pin.pin_memory()
        self._pending: List = []
        self._i, self._depth = 0, depth

    def push(self, res: TaskBatchResult):
        bad = (res.status & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED)) != 0
        nc = (res.status & ops.ST_NOT_CONVERGED) != 0
        lps = torch.where(bad, torch.zeros_like(res.loss_per_sample), res.loss_per_sample)
        n_ok = torch.clamp((~bad).sum(), min=1).to(torch.float32)
        rec = torch.stack([lps.sum() / n_ok, bad.sum().to(torch.float32), nc.sum().to(torch.float32)])
        slot = self._i % self._depth
        self._i += 1
        self._pin[slot].copy_(rec, non_blocking=True)
        ev = None
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record()
        self._pending.append((ev, slot))
        out = []
        while len(self._pending) > 1:
            out.append(self._read(self._pending.pop(0)))
        return out

    def flush(self):
        out = [self._read(p) for p in self._pending]
        self._pending = []
        return out

    def _read(self, item):
        ev, slot = item
        if ev is not None:
            ev.synchronize()
        v = self._pin[slot]
        return float(v[0]), int(v[1]), int(v[2])
