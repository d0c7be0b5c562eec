"""Sibling GP models on the same sm_100a kernels: DKT, DKL and the single-task Tanimoto GP.

  DKTModel / DKTModelTrainerConfig     fs_mol/models/dkt.py:27-176, fs_mol/utils/dkt_utils.py:43-60
  DKLModel / DKLModelTrainerConfig     fs_mol/models/dkl.py:25-161
  gpst_test                            fs_mol/gpst_test.py:37-116 (GP-ST: ExactTanimotoGP on binary fingerprints)

They differ from ADKTModel only in what plays the role of the inner problem:

  * DKT meta-learns ONE shared phi = (raw_noise, raw_outputscale, raw_lengthscale) together with the extractor by
    minimising -MLL of the support AND query molecules of a task taken together (dkt.py:149-154, 168-170); at
    meta-test time phi is kept (or re-fitted from the saved values when `test_time_adaptation` is on, :158-164).
  * DKL trains extractor + phi on the support set of a single task (dkl.py:139-141) and predicts the query set.
  * Neither puts a prior on the noise (dkt.py:87, dkl.py:84); the optional lengthscale prior is LogNormal(0, 0.25)
    and the lengthscale then starts at the prior mean exp(0.03125) (dkt.py:93-100).

-MLL, its gradient w.r.t. phi and w.r.t. the features come out of ONE launch of adkf_gp_ift (mode 2) over all T
tasks of a batch; the posterior out of adkf_gp_predict.  No CPU fallback.
"""
import math
from copy import deepcopy
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.nn as nn

from . import gp as gpk
from . import gp_utils, metrics, ops
from .adaptive_dkt import ADKTModel, ADKTModelTrainerConfig
from .feature_extractor import GraphFeatureExtractorConfig, cli_default_config

DKT_LS_PRIOR_LOC = 0.0                                        # dkt.py:94-95
DKT_LS_PRIOR_MEAN = math.exp(DKT_LS_PRIOR_LOC + 0.5 * gpk.PRIOR_SCALE ** 2)   # LogNormal mean, dkt.py:100


@dataclass(frozen=True)
class DKTModelConfig:
    graph_feature_extractor_config: GraphFeatureExtractorConfig = field(default_factory=cli_default_config)
    used_features: str = "gnn+ecfp+fc"


@dataclass(frozen=True)
class DKTModelTrainerConfig(DKTModelConfig):
    """fs_mol/utils/dkt_utils.py:43-60."""
    batch_size: int = 256
    tasks_per_batch: int = 1
    support_set_size: int = 16
    query_set_size: int = 256
    num_train_steps: int = 10000
    validate_every_num_steps: int = 50
    validation_support_set_sizes: Tuple[int, ...] = (16, 128)
    validation_query_set_size: int = 256
    validation_num_samples: int = 5
    learning_rate: float = 0.001
    clip_value: Optional[float] = None
    use_ard: bool = False
    gp_kernel: str = "matern"
    use_lengthscale_prior: bool = False
    use_numeric_labels: bool = False


DKLModelConfig = DKTModelConfig


@dataclass(frozen=True)
class DKLModelTrainerConfig(DKTModelTrainerConfig):
    pass


class _SharedPhiNegMLL(torch.autograd.Function):
    """-MLL of every task of a layout at ONE shared phi, as a [T] vector differentiable in (Z, phi)."""

    @staticmethod
    def forward(ctx, Z, raw_noise, raw_outputscale, raw_lengthscale, y, layout, kernel, use_numeric, ls_loc):
        T = layout.T
        phi = torch.stack([raw_noise.reshape(()), raw_outputscale.reshape(()), raw_lengthscale.reshape(())])
        phi = phi.detach().float().view(1, 3).expand(T, 3).contiguous()
        g = gpk.BatchedExactGP(Z, y, layout, kernel, use_numeric, use_lengthscale_prior=ls_loc is not None,
                               qq_block=True)
        g.noise_loc = float("nan")                       # no noise prior (dkt.py:87)
        aux0 = torch.zeros(T, ops.AUX, dtype=torch.float32, device=Z.device)
        if ls_loc is not None:
            aux0[:, ops.AUX_LS_LOC] = ls_loc
        dZ, aux, status = g.hypergradient(gpk.FitResult(phi, aux0, None), mode=2)
        rows = ops.host_to_device(layout.n_rows_host, torch.int64, Z.device)
        ctx.save_for_backward(dZ, aux, rows)
        ctx.shapes = (raw_noise.shape, raw_outputscale.shape, raw_lengthscale.shape)
        ctx.mark_non_differentiable(status)
        return aux[:, ops.AUX_F].clone(), status

    @staticmethod
    def backward(ctx, go, _gstatus):
        dZ, aux, rows = ctx.saved_tensors
        gphi = (aux[:, ops.AUX_G:ops.AUX_G + 3] * go.unsqueeze(1)).sum(0)
        s = ctx.shapes
        w = torch.repeat_interleave(go, rows).unsqueeze(1)
        return (dZ * w, gphi[0].reshape(s[0]), gphi[1].reshape(s[1]), gphi[2].reshape(s[2]),
                None, None, None, None, None)


class PriorOutput:
    """What `model(batch)` returns in training mode: stands for the gpytorch MultivariateNormal of the GP prior at
    the training inputs; the only thing the reference does with it is `compute_loss` (dkt.py:168-170)."""

    def __init__(self, neg_mll: torch.Tensor, status: torch.Tensor):
        self.neg_mll, self.status = neg_mll, status


class DKTModel(ADKTModel):
    """fs_mol/models/dkt.py:37-176 on the batched kernels.  Reuses ADKTModel's extractor / fc / parameter holders
    (same state-dict names); the GP parameters are ordinary meta-learned parameters here."""

    _train_on_query_too = True

    def __init__(self, config: DKTModelTrainerConfig):
        super().__init__(config)
        self.test_time_adaptation = False

    def _create_tail_gp(self):
        super()._create_tail_gp()
        if self.config.use_lengthscale_prior:
            self.gp_model.covar_module.base_kernel.lengthscale = DKT_LS_PRIOR_MEAN
            self.gp_model.covar_module.base_kernel.lengthscale_prior_loc = DKT_LS_PRIOR_LOC

    def _ls_loc(self) -> Optional[float]:
        stationary = ops.KERNEL_STATS[self.config.gp_kernel] == ops.STAT_SQDIST
        return DKT_LS_PRIOR_LOC if (self.config.use_lengthscale_prior and stationary) else None

    def save_gp_params(self):
        self.gp_model_params = deepcopy(self.gp_model.state_dict())
        self.gp_likelihood_params = deepcopy(self.gp_likelihood.state_dict())

    def load_gp_params(self):
        self.gp_model.load_state_dict(self.gp_model_params)
        self.gp_likelihood.load_state_dict(self.gp_likelihood_params)

    # ---- batched surface: T tasks per launch -------------------------------------------------------------
    def neg_mll_on_rows(self, Z: torch.Tensor, y: torch.Tensor, n_rows: Sequence[int]):
        """[T] -MLL (mean over a task's rows, gpytorch convention) with autograd to Z and the shared phi."""
        layout = ops.TaskLayout.build(list(n_rows), [0] * len(n_rows), Z.device)
        p = self.gp_params()
        return _SharedPhiNegMLL.apply(Z, p[0], p[1], p[2], y.float().contiguous(), layout, self.config.gp_kernel,
                                      self.config.use_numeric_labels, self._ls_loc())

    def predict_rows(self, Z: torch.Tensor, y: torch.Tensor, layout: ops.TaskLayout, adapt: bool = False,
                     fit_max_iter: int = 100, fit_gtol: float = 2e-6):
        """Posterior (mean, variance incl. noise, phi used [T,3], status) on the query rows of every task."""
        cfg = self.config
        g = gpk.BatchedExactGP(Z, y.float(), layout, cfg.gp_kernel, cfg.use_numeric_labels,
                               use_lengthscale_prior=self._ls_loc() is not None, qq_block=False)
        g.noise_loc = float("nan")
        phi = self._read_phi().expand(layout.T, 3).contiguous()
        st = None
        if adapt:                                          # dkt.py:158-164: fit_gpytorch_scipy from the saved values
            aux0 = torch.zeros(layout.T, ops.AUX, dtype=torch.float32, device=Z.device)
            if self._ls_loc() is not None:
                aux0[:, ops.AUX_LS_LOC] = self._ls_loc()
            res = g.fit(start=gpk.FitResult(phi, aux0, None), max_iter=fit_max_iter, gtol=fit_gtol)
            phi, st = res.phi, res.status
        mean, var, st2 = g.predict(phi)
        return mean, var, phi, (st2 if st is None else st | st2)

    # ---- the reference's single-task surface ---------------------------------------------------------------
    def forward(self, input_batch, train: Optional[bool] = None):
        zs = self.embed(input_batch.support_features)
        ys = self._labels(input_batch, "support")
        n = zs.shape[0]
        if self.training:
            if self._train_on_query_too:
                zq = self.embed(input_batch.query_features)
                Z, y = torch.cat([zs, zq]), torch.cat([ys, self._labels(input_batch, "query")])
            else:
                assert train is None or train
                Z, y = zs, ys
            self.gp_model.set_train_data(inputs=Z, targets=y, strict=False)
            loss, status = self.neg_mll_on_rows(Z, y, [Z.shape[0]])
            return PriorOutput(loss[0], status)
        with torch.no_grad():
            zq = self.embed(input_batch.query_features)
            m = zq.shape[0]
            if self.test_time_adaptation:
                self.load_gp_params()
            layout = ops.TaskLayout.build([n], [m], zs.device)
            Z = torch.cat([zs, zq])
            y = torch.cat([ys, torch.zeros(m, device=ys.device)])
            mean, var, phi, _ = self.predict_rows(Z, y, layout, adapt=self.test_time_adaptation)
            if self.test_time_adaptation:
                self._write_phi(phi[0])
            self.gp_model.set_train_data(inputs=zs, targets=ys, strict=False)
            return gpk.PredictiveDistribution(mean[n:], var[n:])

    def compute_loss(self, logits: PriorOutput) -> torch.Tensor:
        assert self.training
        return logits.neg_mll


class DKLModel(DKTModel):
    """fs_mol/models/dkl.py:35-161: as DKT but trained on the support set of ONE task; `init_params` keeps the
    initial state so every task starts from the same weights (dkl.py:77)."""

    _train_on_query_too = False

    def __init__(self, config: DKLModelTrainerConfig):
        super().__init__(config)
        self.init_params = deepcopy(self.state_dict())

    def forward(self, input_batch, train: bool):
        if self.training and train:
            return super().forward(input_batch, True)
        assert not self.training and not train
        return super().forward(input_batch)


def dkt_run_on_task_batch(model: DKTModel, tb, train: bool = False, grad_scale: Optional[float] = None):
    """T tasks of a DeviceTaskBatch in one pass (batched form of dkt_utils.run_on_batches, dkt_utils.py:63-132).
    train=True: accumulates d(mean_t -MLL_t)/d(theta, phi) into .grad and returns the per-task losses;
    train=False: returns (mean, variance, query mask, predictions)."""
    if train:
        model.train()
        Z = model.embed(tb)
        loss, status = model.neg_mll_on_rows(Z, tb.labels, tb.layout.n_rows_host)
        scale = (1.0 / tb.num_tasks) if grad_scale is None else grad_scale
        (loss.sum() * scale).backward()
        return loss.detach(), status
    model.eval()
    with torch.no_grad():
        Z = model.embed(tb)
        mean, var, phi, st = model.predict_rows(Z, tb.labels, tb.layout, adapt=model.test_time_adaptation)
        qm = torch.ones(tb.layout.total_rows, dtype=torch.bool)
        for t in range(tb.layout.T):
            o = sum(tb.layout.n_rows_host[:t])
            qm[o:o + tb.layout.n_sup_host[t]] = False
        qm = qm.to(Z.device)
        mq = mean[qm]
        return mean, var, qm, (mq if model.config.use_numeric_labels else torch.sigmoid(mq))


class DKTModelTrainer(DKTModel):
    """fs_mol/utils/dkt_utils.py:283-431 reduced to the step itself: Adam over ALL parameters (extractor + the shared GP
    parameters, :296-303), optional clip_grad_norm_ (:383-384).  `train_step` runs the `tasks_per_batch` tasks of a collated
    DeviceTaskBatch in one pass (dkt_run_on_task_batch) instead of one run_on_batches call per task."""

    def __init__(self, config: DKTModelTrainerConfig):
        super().__init__(config)
        self.optimizer = torch.optim.Adam(self.parameters(), config.learning_rate)
        self.lr_scheduler = None

    def train_step(self, tb):
        """Returns the mean per-sample loss of the step, as the reference logs it (dkt_utils.py:95-100, 386)."""
        self.optimizer.zero_grad(set_to_none=True)
        losses, status = dkt_run_on_task_batch(self, tb, train=True)
        if self.config.clip_value is not None:
            torch.nn.utils.clip_grad_norm_(self.parameters(), self.config.clip_value)
        self.optimizer.step()
        ops.weights_changed()
        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        return losses.mean(), status

    def get_model_state(self):
        return {"model_config": self.config, "model_state_dict": self.state_dict()}

    def save_model(self, path: str, optimizer=None, epoch=None):
        data = self.get_model_state()
        if optimizer is not None:
            data["optimizer_state_dict"] = optimizer.state_dict()
        if epoch is not None:
            data["epoch"] = epoch
        torch.save(data, path)

    @classmethod
    def build_from_model_file(cls, model_file: str, config_overrides=None, quiet: bool = False, device=None):
        ck = torch.load(model_file, map_location=device, weights_only=False)
        model = cls(ck["model_config"])
        own = model.state_dict()
        model.load_state_dict({k: v for k, v in ck["model_state_dict"].items() if k in own}, strict=False)
        return model.to(device) if device is not None else model


def gpst_test(train_fp: torch.Tensor, y_train: torch.Tensor, test_fp: torch.Tensor, y_test,
              use_numeric_labels: bool, test_batch_size: int = 700):
    """GP-ST (fs_mol/gpst_test.py:37-116): single-task Tanimoto GP on binary fingerprints.  `train_fp` / `test_fp`
    are the [n,2048] binary fingerprints (the reference gets them from rdkit, chem_utils.py:12-23 -- out of
    scope); y_* raw labels (numeric: positive activities, log + standardise with the TRAIN mean / unbiased std,
    gpst_test.py:57-61; bool: 0/1 -> +-1).  Returns (metrics object, predictions, phi*)."""
    dev = train_fp.device
    X_train, X_test = train_fp.float(), test_fp.float()
    if use_numeric_labels:
        ly = torch.log(torch.as_tensor(y_train, dtype=torch.float32))
        mu, sd = ly.mean(), ly.std()
        y_tr = ((ly - mu) / sd).to(dev)
        y_te = ((torch.log(torch.as_tensor(y_test, dtype=torch.float32)) - mu) / sd).cpu().numpy()
    else:
        y_tr = torch.as_tensor(y_train, dtype=torch.float32).to(dev) * 2.0 - 1.0
        y_te = np.asarray(torch.as_tensor(y_test, dtype=torch.float32).cpu())
    likelihood = gp_utils.GaussianLikelihood().to(dev)
    model = gp_utils.ExactTanimotoGP(X_train, y_tr, likelihood, use_numeric_labels).to(dev)
    mll = gp_utils.ExactMarginalLogLikelihood(likelihood, model)
    model.train()
    likelihood.train()
    gp_utils.fit_gpytorch_scipy(mll)
    model.eval()
    likelihood.eval()
    preds: List[np.ndarray] = []
    with torch.no_grad():
        for i in range(0, X_test.shape[0], test_batch_size):
            dist = likelihood(model(X_test[i:i + test_batch_size]))
            p = dist.mean if use_numeric_labels else torch.sigmoid(dist.mean)
            preds.append(p.detach().cpu().numpy())
    pred = np.concatenate(preds, axis=0)
    if use_numeric_labels:
        m = metrics.compute_numeric_task_metrics(predictions=pred, labels=y_te)
    else:
        m = metrics.compute_binary_task_metrics(predictions=pred, labels=y_te)
    return m, pred, model._phi()


def gpst_run_on_tasks(train_fps: Sequence[torch.Tensor], train_ys: Sequence[torch.Tensor], test_fps: Sequence[torch.Tensor],
                      use_numeric_labels: bool, fit_max_iter: int = 100, fit_gtol: float = 2e-6):
    """GP-ST over T tasks in ONE launch per stage (Tanimoto pair statistics, fit, posterior): what `gpst_test` does one task
    at a time (fs_mol/gpst_test.py:68-104).  train_ys are the CONVERTED training targets (+-1, or standardised log-activities).
    Returns (list of per-task predictions on the test molecules -- sigmoid(mean) for classification --, phi* [T,3], status [T])."""
    dev = train_fps[0].device
    n_s = [int(x.shape[0]) for x in train_fps]
    n_q = [int(x.shape[0]) for x in test_fps]
    Z = torch.cat([t for pair in zip(train_fps, test_fps) for t in pair]).float().contiguous()
    y = torch.cat([t for ys, m in zip(train_ys, n_q) for t in (ys.float().to(dev), torch.zeros(m, device=dev))]).contiguous()
    layout = ops.TaskLayout.build(n_s, n_q, dev)
    g = gpk.BatchedExactGP(Z, y, layout, "tanimoto", use_numeric_labels, use_lengthscale_prior=False, qq_block=False,
                           freeze=0 if use_numeric_labels else ops.FREEZE_NOISE)        # gp_utils.py:143
    g.noise_loc = float("nan")                                                          # no priors in GP-ST
    T = len(n_s)
    noise0 = torch.tensor(gpk.noise_init(use_numeric_labels) - gpk.NOISE_FLOOR)
    phi0 = torch.zeros(T, 3, dtype=torch.float32)
    phi0[:, 0] = noise0 + torch.log(-torch.expm1(-noise0))                              # inverse softplus
    start = gpk.FitResult(phi0.to(dev), torch.zeros(T, ops.AUX, dtype=torch.float32, device=dev), None)
    res = g.fit(start=start, max_iter=fit_max_iter, gtol=fit_gtol)
    mean, _, st = g.predict(res.phi)
    preds, off = [], 0
    for a, b in zip(n_s, n_q):
        mq = mean[off + a: off + a + b]
        preds.append(mq if use_numeric_labels else torch.sigmoid(mq))
        off += a + b
    return preds, res.phi, res.status | st
