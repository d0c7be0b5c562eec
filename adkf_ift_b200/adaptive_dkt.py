"""ADKTModel on the B200 kernels -- same call surface as fs_mol/models/adaptive_dkt.py:26-209.

Kept from the reference: config fields, attribute names (`graph_feature_extractor`, `fc`, `use_fc`, `fc_out_dim`,
`gp_likelihood`, `gp_model`, `mll`, `normalizing_features`), `feature_extractor_params()` / `gp_params()` (names
not starting / starting with `gp_`, registration order), `reinit_gp_params`, `compute_median_lengthscale_init`,
the five branches of `forward`, and every non-GP state-dict key.  The three GP tensors keep their gpytorch names
(`gp_likelihood.noise_covar.raw_noise [1]`, `gp_model.covar_module.raw_outputscale []`,
`gp_model.covar_module.base_kernel.raw_lengthscale [1,1]`).

New (the reason this package exists): `embed(task_batch)` runs the extractor ONCE over all graphs of T tasks and
`adaptive_dkt_utils.run_on_task_batch` drives the batched GP kernels on the result.  The single-task `forward`
is the T=1 case of the same kernels -- there is no gpytorch, no torch_scatter and no CPU path behind it.
"""
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import gp as gpk
from . import ops
from .feature_extractor import GraphFeatureExtractor, GraphFeatureExtractorConfig, cli_default_config

FINGERPRINT_DIM = 2048
PHYS_CHEM_DESCRIPTORS_DIM = 42


@dataclass(frozen=True)
class ADKTModelConfig:
    graph_feature_extractor_config: GraphFeatureExtractorConfig = field(default_factory=cli_default_config)
    used_features: str = "gnn+ecfp+fc"


@dataclass(frozen=True)
class ADKTModelTrainerConfig(ADKTModelConfig):
    """fs_mol/utils/adaptive_dkt_utils.py:47-67 (ADKTModel reads gp_kernel / use_ard / use_numeric_labels /
    use_lengthscale_prior, which only exist here -- always construct the model with this class)."""
    batch_size: int = 256
    tasks_per_batch: int = 16
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
    use_lengthscale_prior: bool = True
    use_numeric_labels: bool = False
    ignore_grad_correction: bool = False


def _inv_softplus(x: torch.Tensor) -> torch.Tensor:
    return x + torch.log(-torch.expm1(-x))


# ---- parameter holders with gpytorch's attribute paths ------------------------------------------------
class _NoiseCovar(nn.Module):
    def __init__(self):
        super().__init__()
        self.raw_noise = nn.Parameter(torch.zeros(1))

    @property
    def noise(self):
        return F.softplus(self.raw_noise) + gpk.NOISE_FLOOR

    @noise.setter
    def noise(self, value):
        v = torch.as_tensor(value, dtype=self.raw_noise.dtype, device=self.raw_noise.device)
        with torch.no_grad():
            self.raw_noise.copy_(_inv_softplus(v - gpk.NOISE_FLOOR).expand_as(self.raw_noise))


class GaussianLikelihoodParams(nn.Module):
    def __init__(self):
        super().__init__()
        self.noise_covar = _NoiseCovar()

    @property
    def noise(self):
        return self.noise_covar.noise

    @noise.setter
    def noise(self, value):
        self.noise_covar.noise = value


class _BaseKernel(nn.Module):
    def __init__(self, ard_num_dims: Optional[int] = None):
        super().__init__()
        self.ard_num_dims = ard_num_dims
        self.raw_lengthscale = nn.Parameter(torch.zeros(1, 1 if ard_num_dims is None else ard_num_dims))
        self.lengthscale_prior_loc: Optional[float] = None

    @property
    def lengthscale(self):
        return F.softplus(self.raw_lengthscale)

    @lengthscale.setter
    def lengthscale(self, value):
        v = torch.as_tensor(value, dtype=self.raw_lengthscale.dtype, device=self.raw_lengthscale.device)
        with torch.no_grad():
            self.raw_lengthscale.copy_(_inv_softplus(v).expand_as(self.raw_lengthscale))


class _ScaleKernel(nn.Module):
    def __init__(self, ard_num_dims: Optional[int] = None):
        super().__init__()
        self.raw_outputscale = nn.Parameter(torch.zeros(()))
        self.base_kernel = _BaseKernel(ard_num_dims)

    @property
    def outputscale(self):
        return F.softplus(self.raw_outputscale)

    @outputscale.setter
    def outputscale(self, value):
        v = torch.as_tensor(value, dtype=self.raw_outputscale.dtype, device=self.raw_outputscale.device)
        with torch.no_grad():
            self.raw_outputscale.copy_(_inv_softplus(v))


class ExactGPLayer(nn.Module):
    """State of the exact GP bound to ONE task (fs_mol/utils/gp_utils.py:7-49): the three raw parameters plus
    the support features / labels.  All arithmetic happens in the batched kernels (gp.BatchedExactGP, T=1)."""

    def __init__(self, train_x=None, train_y=None, likelihood: Optional[GaussianLikelihoodParams] = None,
                 kernel: str = "matern", ard_num_dims=None, use_numeric_labels: bool = False):
        super().__init__()
        self.kernel = gpk.check_kernel(kernel)
        if ard_num_dims is not None and self.kernel not in ("matern", "rbf", "RBF"):
            ard_num_dims = None                          # gp_utils.py:22-43: only the RBF / Matern branches take ard_num_dims
        self.ard_num_dims = ard_num_dims
        self.use_numeric_labels = use_numeric_labels
        self.covar_module = _ScaleKernel(ard_num_dims)
        self.train_inputs: Optional[Tuple[torch.Tensor]] = None
        self.train_targets: Optional[torch.Tensor] = None
        if likelihood is not None:
            likelihood.noise = gpk.noise_init(use_numeric_labels)
        if train_x is not None:
            self.set_train_data(train_x, train_y)

    def set_train_data(self, inputs=None, targets=None, strict=False):
        if inputs is not None:
            self.train_inputs = (inputs,)
        if targets is not None:
            self.train_targets = targets


class MarginalLogLikelihood:
    """Handle returned as `model.mll`: what the reference passes to `fit_gpytorch_scipy`."""

    def __init__(self, model: "ADKTModel"):
        self._model = model

    @property
    def model(self):
        return self._model.gp_model

    @property
    def likelihood(self):
        return self._model.gp_likelihood

    def fit(self, max_iter: int = 50, gtol: float = 2e-6):
        return self._model.fit_gp(max_iter=max_iter, gtol=gtol)


def _bind_task_gp(Z, raws, y, layout, cfg, ls_loc):
    phi = torch.stack([r.detach().reshape(()) for r in raws]).view(1, 3).float().contiguous()
    g = gpk.BatchedExactGP(Z, y, layout, cfg.gp_kernel, cfg.use_numeric_labels,
                           use_lengthscale_prior=ls_loc is not None, qq_block=True)
    aux0 = torch.zeros(1, ops.AUX, dtype=torch.float32, device=Z.device)
    if ls_loc is not None:
        aux0[0, ops.AUX_LS_LOC] = ls_loc
    return g, gpk.FitResult(phi, aux0, None)


class _InnerLossGrad(torch.autograd.Function):
    """(df/dZ, df/draw_noise, df/draw_outputscale, df/draw_lengthscale) of the inner loss f = -MLL as a DIFFERENTIABLE
    function of (Z, phi).  Its backward is the second-order information the generic hypergradient routine asks for
    (cauchy_hypergradient.py:41-64,120-161): for a cotangent u on the phi-gradient it returns H u (w.r.t. phi, from the
    kernel's analytic 3x3 Hessian) and the mixed term d/deps df/dZ(phi + eps u) (w.r.t. Z, adkf_gp_ift mode 3).  A
    cotangent on df/dZ would need d2f/dZ2, which nothing on this path uses."""

    @staticmethod
    def forward(ctx, Z, raw_noise, raw_outputscale, raw_lengthscale, y, layout, cfg, ls_loc):
        g, fit = _bind_task_gp(Z, (raw_noise, raw_outputscale, raw_lengthscale), y, layout, cfg, ls_loc)
        dZ, aux, _ = g.hypergradient(fit, mode=2)
        ctx.gp, ctx.fit = g, gpk.FitResult(fit.phi, aux, None)
        ctx.shapes = (raw_noise.shape, raw_outputscale.shape, raw_lengthscale.shape)
        ctx.set_materialize_grads(False)
        gphi = aux[0, ops.AUX_G:ops.AUX_G + 3]
        s = ctx.shapes
        return dZ, gphi[0].reshape(s[0]), gphi[1].reshape(s[1]), gphi[2].reshape(s[2])

    @staticmethod
    def backward(ctx, cZ, c0, c1, c2):
        if cZ is not None:
            raise NotImplementedError("second derivative of the inner GP loss w.r.t. the features (d2f/dZ2) is not on the "
                                      "ADKF-IFT path: only grad_phi f is differentiated again")
        aux = ctx.fit.aux
        u = torch.stack([(torch.zeros((), device=aux.device) if c is None else c.reshape(()).float()) for c in (c0, c1, c2)])
        H = aux[0, ops.AUX_H:ops.AUX_H + 9].view(3, 3)
        Hu = H @ u
        dZ_mixed, _, _ = ctx.gp.inner_mixed_term(ctx.fit, u.view(1, 3))
        s = ctx.shapes
        return dZ_mixed, Hu[0].reshape(s[0]), Hu[1].reshape(s[1]), Hu[2].reshape(s[2]), None, None, None, None


class _TaskGPLoss(torch.autograd.Function):
    """-MLL (inner) or predictive joint NLL (outer) of one task as a differentiable scalar of (Z, phi):
    forward and first-order backward both come out of adkf_gp_ift (mode 2 / mode 1).  Under create_graph=True the
    backward of the INNER loss is itself differentiable (_InnerLossGrad) -- the reference's own training-step sequence
    (functional_call closures + the generic cauchy_hypergradient, adaptive_dkt_utils.py:361-404) runs on this model."""

    @staticmethod
    def forward(ctx, Z, raw_noise, raw_outputscale, raw_lengthscale, y, layout, cfg, ls_loc, outer: bool):
        g, fit = _bind_task_gp(Z, (raw_noise, raw_outputscale, raw_lengthscale), y, layout, cfg, ls_loc)
        dZ, aux, _ = g.hypergradient(fit, mode=1 if outer else 2)
        ctx.save_for_backward(dZ, aux, Z, raw_noise, raw_outputscale, raw_lengthscale)
        ctx.outer = outer
        ctx.rest = (y, layout, cfg, ls_loc)
        ctx.shapes = (raw_noise.shape, raw_outputscale.shape, raw_lengthscale.shape)
        return aux[0, ops.AUX_FOUTER if outer else ops.AUX_F].clone()

    @staticmethod
    def backward(ctx, go):
        dZ, aux, Z, rn, ro, rl = ctx.saved_tensors
        if torch.is_grad_enabled() and not ctx.outer:           # create_graph=True: keep the gradient differentiable
            gZ, g0, g1, g2 = _InnerLossGrad.apply(Z, rn, ro, rl, *ctx.rest)
            return gZ * go, g0 * go, g1 * go, g2 * go, None, None, None, None, None
        k = ops.AUX_DF_DPHI if ctx.outer else ops.AUX_G
        gphi = aux[0, k:k + 3] * go
        s = ctx.shapes
        return (dZ * go, gphi[0].reshape(s[0]), gphi[1].reshape(s[1]), gphi[2].reshape(s[2]),
                None, None, None, None, None)


class ADKTModel(nn.Module):
    def __init__(self, config: ADKTModelTrainerConfig):
        super().__init__()
        self.config = config
        if gpk.check_kernel(config.gp_kernel) == "tanimoto":   # ExactGPLayer has no Tanimoto branch (gp_utils.py:22-45)
            raise ValueError("[ERROR] the kernel 'tanimoto' is not supported! (use gp_utils.ExactTanimotoGP)")
        if config.used_features.startswith("gnn"):
            self.graph_feature_extractor = GraphFeatureExtractor(config.graph_feature_extractor_config)
        self.use_fc = config.used_features.endswith("+fc")
        if self.use_fc:
            self.fc_out_dim = 2048
            self._gnn_dim = (config.graph_feature_extractor_config.readout_config.output_dim
                             if "gnn" in config.used_features else 0)
            fc_in = self._gnn_dim
            if "ecfp" in config.used_features:
                fc_in += FINGERPRINT_DIM
            if "pc-descs" in config.used_features:
                fc_in += PHYS_CHEM_DESCRIPTORS_DIM
            self.fc = nn.Sequential(nn.Linear(fc_in, 2048), nn.ReLU(), nn.Linear(2048, self.fc_out_dim))
        self._create_tail_gp()
        self.normalizing_features = config.gp_kernel == "cossim"
        self._bound = None   # per-task state of the single-task surface

    # ---- reference helpers -------------------------------------------------------------------------
    def feature_extractor_params(self) -> List[nn.Parameter]:
        return [p for n, p in self.named_parameters() if not n.startswith("gp_")]

    def gp_params(self) -> List[nn.Parameter]:
        return [p for n, p in self.named_parameters() if n.startswith("gp_")]

    @property
    def device(self) -> torch.device:
        return next(self.parameters()).device

    def _create_tail_gp(self):
        dev = next(self.parameters()).device if any(True for _ in self.parameters()) else None
        self.gp_likelihood = GaussianLikelihoodParams()
        ard = getattr(self, "fc_out_dim", None) if getattr(self.config, "use_ard", False) else None   # adaptive_dkt.py:107-110
        if getattr(self.config, "use_ard", False) and ard is None:
            raise AttributeError("use_ard needs fc_out_dim, i.e. used_features ending in '+fc' (adaptive_dkt.py:107-108)")
        self.gp_model = ExactGPLayer(likelihood=self.gp_likelihood, kernel=self.config.gp_kernel, ard_num_dims=ard,
                                     use_numeric_labels=self.config.use_numeric_labels)
        if dev is not None:
            self.gp_likelihood.to(dev)
            self.gp_model.to(dev)
        self.mll = MarginalLogLikelihood(self)

    @property
    def use_ard(self) -> bool:
        return self.gp_model.ard_num_dims is not None

    def compute_median_lengthscale_init(self, gp_input: torch.Tensor) -> torch.Tensor:
        """sqrt(0.5 * lower-median of the positive pairwise squared distances) (adaptive_dkt.py:128-131), from
        the radix-select inside the fit kernel."""
        n = gp_input.shape[0]
        layout = ops.TaskLayout.build([n], [0], gp_input.device)
        g = gpk.BatchedExactGP(gp_input.float(), torch.zeros(n, device=gp_input.device), layout, self.config.gp_kernel)
        return g.init().aux[0, ops.AUX_L0]

    def reinit_gp_params(self, gp_input: torch.Tensor, use_lengthscale_prior: bool = False,
                         targets: Optional[torch.Tensor] = None):
        """Fresh GP parameters for a new task (adaptive_dkt.py:88-101): noise 0.1 | 0.01, outputscale softplus(0),
        lengthscale = median heuristic, LogNormal prior loc = log(l0) + 0.0625."""
        self._create_tail_gp()
        n = gp_input.shape[0]
        layout = ops.TaskLayout.build([n], [0], gp_input.device)
        y = targets if targets is not None else torch.zeros(n, device=gp_input.device)
        if self.use_ard:
            from .gp_ard import BatchedARDExactGP
            g = BatchedARDExactGP(gp_input.float(), y.float(), layout, self.config.gp_kernel, self.config.use_numeric_labels,
                                  use_lengthscale_prior)
            phi0, l0, loc = g.init_phi()
            self._write_phi(phi0[0])
            self.gp_model.covar_module.base_kernel.lengthscale_prior_loc = loc[0] if use_lengthscale_prior else None
            self._bound = dict(gp=g, ard=True, ls_loc=loc, use_prior=use_lengthscale_prior)
            return
        g = gpk.BatchedExactGP(gp_input.float(), y.float(), layout, self.config.gp_kernel,
                               self.config.use_numeric_labels, use_lengthscale_prior)
        init = g.init()
        self._write_phi(init.phi[0])
        self.gp_model.covar_module.base_kernel.lengthscale_prior_loc = (
            init.aux[0, ops.AUX_LS_LOC] if use_lengthscale_prior else None)
        self._bound = dict(gp=g, fit=init, use_prior=use_lengthscale_prior)

    def _write_phi(self, phi3: torch.Tensor):
        with torch.no_grad():
            self.gp_likelihood.noise_covar.raw_noise.copy_(phi3[0:1])
            self.gp_model.covar_module.raw_outputscale.copy_(phi3[1])
            self.gp_model.covar_module.base_kernel.raw_lengthscale.copy_(phi3[2:].view(1, -1))

    def _read_phi(self) -> torch.Tensor:
        return torch.cat([p.detach().reshape(-1) for p in self.gp_params()]).view(1, -1).float().contiguous()

    def fit_gp(self, max_iter: int = 50, gtol: float = 2e-6):
        """Replacement for `fit_gpytorch_scipy(model.mll)` (adaptive_dkt_utils.py:91): on-device Newton fit of the
        three raw parameters on the bound support set, starting from the current values."""
        if self._bound is None:
            raise RuntimeError("no task bound: call model(batch, train_loss=True) first")
        b = self._bound
        if b.get("ard"):
            res = b["gp"].fit(start=self._read_phi())
            self._write_phi(res.phi[0])
            b["fit"] = res
            return res
        start = gpk.FitResult(self._read_phi(), b["fit"].aux, b["fit"].status)
        res = b["gp"].fit(start=start, max_iter=max_iter, gtol=gtol)
        self._write_phi(res.phi[0])
        b["fit"] = res
        return res

    # ---- features ------------------------------------------------------------------------------------
    def _fc_forward(self, gnn_out: Optional[torch.Tensor], fingerprints: Optional[torch.Tensor],
                    descriptors: Optional[torch.Tensor]) -> torch.Tensor:
        """fc(cat(gnn, ecfp.float(), descs)) with the ~2%-dense int32 fingerprint block of fc[0] done by the
        sparse gather kernel instead of a dense [G,2048]x[2048,2048] GEMM."""
        w0, b0 = self.fc[0].weight, self.fc[0].bias
        col = 0
        base = None
        if gnn_out is not None:
            base = ops.linear_tc(gnn_out, w0[:, :self._gnn_dim], b0)
            col = self._gnn_dim
        fp_cols = None
        if fingerprints is not None:
            fp_cols = (col, col + FINGERPRINT_DIM)
            col += FINGERPRINT_DIM
        if descriptors is not None:
            d = F.linear(descriptors, w0[:, col:col + PHYS_CHEM_DESCRIPTORS_DIM], b0 if base is None else None)
            base = d if base is None else base + d
        if base is None:
            base = b0.unsqueeze(0).expand(fingerprints.shape[0], -1)
        if fp_cols is not None:
            base = ops.ecfp_linear(base, fingerprints, w0[:, fp_cols[0]:fp_cols[1]].t().contiguous())
        return ops.linear_tc(F.relu(base), self.fc[2].weight, self.fc[2].bias)

    def embed(self, batch) -> torch.Tensor:
        """GP inputs Z [G, d] for every graph of `batch` (an FSMolBatch-like with fingerprints/descriptors on the
        device, or a task_batch.DeviceTaskBatch covering many tasks)."""
        uf = self.config.used_features
        gnn_out = self.graph_feature_extractor(batch) if "gnn" in uf else None
        fp = batch.fingerprints if "ecfp" in uf else None
        desc = batch.descriptors if "pc-descs" in uf else None
        if self.use_fc:
            if fp is not None and fp.dtype != torch.int32:
                fp = fp.to(torch.int32)
            z = self._fc_forward(gnn_out, fp, desc)
        else:
            parts = [p for p in (gnn_out, fp.float() if fp is not None else None, desc) if p is not None]
            z = torch.cat(parts, dim=1)
        if self.normalizing_features:   # gp_kernel == "cossim" (adaptive_dkt.py:158-160)
            z = F.normalize(z, p=2, dim=1)
        return z

    # ---- the reference's forward ------------------------------------------------------------------------
    def _labels(self, batch, which: str) -> torch.Tensor:
        if self.config.use_numeric_labels:
            return getattr(batch, which + "_numeric_labels").float()
        return (getattr(batch, which + "_labels").float() - 0.5) * 2.0

    def forward(self, input_batch, train_loss: Optional[bool], predictive_val_loss: bool = False,
                is_functional_call: bool = False):
        zs = self.embed(input_batch.support_features)
        ys = self._labels(input_batch, "support")
        cfg = self.config
        n = zs.shape[0]
        if self.training:
            assert train_loss is not None
            if train_loss and not is_functional_call:
                self.reinit_gp_params(zs.detach(), cfg.use_lengthscale_prior, targets=ys.detach())
                self.gp_model.set_train_data(inputs=zs.detach(), targets=ys.detach(), strict=False)
                return None
            ls_loc = self.gp_model.covar_module.base_kernel.lengthscale_prior_loc
            phi = self.gp_params()
            if self.use_ard:
                raise NotImplementedError("use_ard: the differentiable single-task losses (functional_call surface) exist for the scalar "
                                          "lengthscale only; meta-training with ARD goes through run_on_task_batch / train_step")
            if train_loss:
                layout = ops.TaskLayout.build([n], [0], zs.device)
                return _TaskGPLoss.apply(zs, phi[0], phi[1], phi[2], ys, layout, cfg, ls_loc, False)
            assert is_functional_call
            zq = self.embed(input_batch.query_features)
            yq = self._labels(input_batch, "query")
            if predictive_val_loss:
                layout = ops.TaskLayout.build([n], [zq.shape[0]], zs.device)
                return _TaskGPLoss.apply(torch.cat([zs, zq]), phi[0], phi[1], phi[2], torch.cat([ys, yq]), layout,
                                         cfg, ls_loc, True)
            layout = ops.TaskLayout.build([zq.shape[0]], [0], zq.device)   # validation MLL on the query set
            return _TaskGPLoss.apply(zq, phi[0], phi[1], phi[2], yq, layout, cfg, ls_loc, False)
        assert train_loss is None
        with torch.no_grad():
            zq = self.embed(input_batch.query_features)
            m = zq.shape[0]
            layout = ops.TaskLayout.build([n], [m], zs.device)
            Z = torch.cat([zs.detach(), zq])
            if self.use_ard:
                from .gp_ard import BatchedARDExactGP
                ga = BatchedARDExactGP(Z, torch.cat([ys, torch.zeros(m, device=ys.device)]), layout, cfg.gp_kernel,
                                       cfg.use_numeric_labels, cfg.use_lengthscale_prior)
                mean, var, _ = ga.predict(self._read_phi())
                self.gp_model.set_train_data(inputs=zs.detach(), targets=ys, strict=False)
                return gpk.PredictiveDistribution(mean[n:], var[n:], None)
            g = gpk.BatchedExactGP(Z, torch.cat([ys, torch.zeros(m, device=ys.device)]), layout, cfg.gp_kernel,
                                   cfg.use_numeric_labels, cfg.use_lengthscale_prior, qq_block=True)
            phi = self._read_phi()
            mean, var, _ = g.predict(phi)
            self.gp_model.set_train_data(inputs=zs.detach(), targets=ys, strict=False)

            def nlp(yq):
                g.y = torch.cat([ys, yq.float()]).contiguous()
                aux0 = torch.zeros(1, ops.AUX, dtype=torch.float32, device=phi.device)
                return g.hypergradient(gpk.FitResult(phi, aux0, None), mode=1)[1][0, ops.AUX_FOUTER]

            return gpk.PredictiveDistribution(mean[n:], var[n:], nlp)
