"""--use-ard on the batched GP kernels: one lengthscale per feature dimension.

Reference: `ExactGPLayer(..., ard_num_dims=fc_out_dim)` (fs_mol/utils/gp_utils.py:14,27,30), configured by
`ADKTModel.__create_tail_GP` / `reinit_gp_params` with `use_ard=True` (fs_mol/models/adaptive_dkt.py:88-126): raw lengthscale
[1, d] initialised to the median heuristic in every dimension, LogNormal(log l0 + 1/16, 1/4) prior SUMMED over the d
lengthscales, noise prior as usual.  The inner problem has d + 2 = 2050 raw parameters per task; the reference fits them with
L-BFGS-B, builds the DENSE 2050 x 2050 Hessian by autograd (cauchy_hypergradient.py:41-64) and solves with it (:136).

Here, for T tasks at once, phi = [T, 2 + d] = (raw_noise, raw_outputscale, raw_lengthscale[0..d-1]):

  u_ab = sum_k w_k (z_ak - z_bk)^2, w = 1 / softplus(raw_lengthscale)^2          adkf_ard_wsqdist
  f, df/d(noise, outputscale), W = df/du                                          adkf_gp_ift mode 2 on u (scalar lengthscale 1)
  df/dl_k = (-2 / l_k^3) sum_ab W_ab (z_ak - z_bk)^2 - d log p(l_k) / n            adkf_ard_pairquad
  H x (Hessian-vector product, nothing (d+2)^2 is formed):
      du = wsqdist with the directional weights, dW and d(df/d(noise, outputscale)) along (x_noise, x_outputscale, du)   mode 4,
      then pairquad(dW)                                                          -> hvp()
  inner fit: batched L-BFGS (two-loop recursion, Armijo backtracking) on the device, all tasks in lock step       -> fit()
  inverse-Hessian-vector product v = H^-1 dF/dphi: conjugate gradients on hvp(), or the Neumann series
      v = a sum_j (I - a H)^j b (north star), both matrix-free                   -> ihvp()
  F, dF/dphi, dF/du from adkf_gp_ift mode 1 on the joint u; dL/dZ = dF/dZ - d/deps df/dZ(phi + eps v)            -> hypergradient()

The n x n arithmetic is that of the scalar-lengthscale path (same kernels, same tests); new code = the two contractions over d
(csrc/gp_ard.cu), mode 4 (csrc/gp_core.cu) and this orchestration.  No CPU fallback.
"""
import math
from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import gp as gpk
from . import ops

PRIOR_SCALE = gpk.PRIOR_SCALE
_INV_SP_ONE = math.log(math.e - 1.0)            # inv_softplus(1): raw value of the unit scalar lengthscale handed to the kernels


@dataclass
class ARDFit:
    phi: torch.Tensor          # [T, 2 + d] raw parameters
    f: torch.Tensor            # [T] inner loss
    grad: torch.Tensor         # [T, 2 + d]
    ls_loc: torch.Tensor       # [T] LogNormal loc of the lengthscale prior (log l0 + 1/16)
    l0: torch.Tensor           # [T] median-heuristic lengthscale
    iterations: int
    status: torch.Tensor       # [T] int32 ADKF_ST_*


class _CudaBackend:
    """The sm_100a kernels behind the ARD algebra (see the module docstring)."""

    def __init__(self, Z, y, layout: ops.TaskLayout, kernel: str, noise_loc: float):
        self.Z, self.y, self.layout, self.kernel, self.noise_loc = Z.contiguous(), y.contiguous(), layout, kernel, noise_loc
        self.T, self.d = layout.T, Z.shape[1]
        self._phi3 = None
        dev = Z.device
        self.rows_task = torch.repeat_interleave(torch.arange(layout.T, device=dev),
                                                 torch.tensor(layout.n_rows_host, device=dev))
        sup = []
        for ns, nr in zip(layout.n_sup_host, layout.n_rows_host):
            sup += [True] * ns + [False] * (nr - ns)
        self.support_rows = ops.host_to_device(sup, torch.bool, dev)
        self.n = ops.host_to_device([float(x) for x in layout.n_sup_host], torch.float32, dev)

    def _lib(self):
        from . import _cabi
        return _cabi.load()

    def wsqdist(self, c: torch.Tensor, support_only: bool) -> torch.Tensor:
        D2 = torch.zeros(max(self.layout.d2_total, 1), dtype=torch.float32, device=self.Z.device)
        L = self.layout
        ops.check(self._lib().adkf_ard_wsqdist(ops.dptr(self.Z, torch.float32), self.d, L.T, ops.dptr(L.z_off), ops.dptr(L.n_sup),
                                               ops.dptr(L.d2_off), L.max_n, ops.dptr(c.contiguous(), torch.float32),
                                               1 if support_only else 0, ops.dptr(D2), ops.stream_ptr()), "adkf_ard_wsqdist")
        return D2

    def pairquad(self, G: torch.Tensor, support_only: bool) -> torch.Tensor:
        q = torch.empty(self.T, self.d, dtype=torch.float32, device=self.Z.device)
        L = self.layout
        ops.check(self._lib().adkf_ard_pairquad(ops.dptr(self.Z, torch.float32), self.d, L.T, ops.dptr(L.z_off), ops.dptr(L.n_sup),
                                                ops.dptr(L.d2_off), ops.dptr(G.contiguous(), torch.float32), 1 if support_only else 0,
                                                ops.dptr(q), ops.stream_ptr()), "adkf_ard_pairquad")
        return q

    def _aux0(self):
        return torch.zeros(self.T, ops.AUX, dtype=torch.float32, device=self.Z.device)

    def phi3(self, phi2: torch.Tensor) -> torch.Tensor:
        return torch.cat([phi2, torch.full((self.T, 1), _INV_SP_ONE, dtype=torch.float32, device=phi2.device)], 1).contiguous()

    def inner(self, D2, phi2) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
        """-> (f incl. noise prior / n, raw gradient w.r.t. (noise, outputscale) [T,2], W = df/du, status)."""
        W, aux, st = ops.gp_ift(self.layout, D2, self.y, self.kernel, self.noise_loc, 0, self.phi3(phi2), self._aux0(), mode=2)
        return aux[:, ops.AUX_F], aux[:, ops.AUX_G:ops.AUX_G + 2], W, st

    def inner_dir(self, D2, phi2, u2, dD2) -> Tuple[torch.Tensor, torch.Tensor]:
        """Directional derivative of (W, raw gradient (noise, outputscale)) along (u2, dD2): adkf_gp_ift mode 4."""
        aux = self._aux0()
        aux[:, ops.AUX_V:ops.AUX_V + 2] = u2
        L = self.layout
        lib = self._lib()
        Wm = dD2.contiguous().clone()
        status = torch.zeros(L.T, dtype=torch.int32, device=D2.device)
        nws = L.ws_floats(2)
        ws = torch.empty(nws, dtype=torch.float32, device=D2.device)
        ops.check(lib.adkf_gp_ift(L.T, ops.dptr(L.z_off), ops.dptr(L.n_sup), ops.dptr(L.d2_off), ops.dptr(D2, torch.float32),
                                  ops.dptr(self.y, torch.float32), ops.kernel_arg(self.kernel), float(self.noise_loc), 0, 4,
                                  ops.dptr(self.phi3(phi2), torch.float32), ops.dptr(aux), ops.dptr(Wm), ops.dptr(status), ops.dptr(ws),
                                  nws, ops.stream_ptr()), "adkf_gp_ift")
        return Wm, aux[:, ops.AUX_DF_DPHI:ops.AUX_DF_DPHI + 2]

    def outer(self, D2_full, phi2):
        """-> (F [T], dF/d raw (noise, outputscale) [T,2], dF/du, status): adkf_gp_ift mode 1 on the joint statistics."""
        W, aux, st = ops.gp_ift(self.layout, D2_full, self.y, self.kernel, self.noise_loc, 0, self.phi3(phi2), self._aux0(), mode=1)
        return aux[:, ops.AUX_FOUTER], aux[:, ops.AUX_DF_DPHI:ops.AUX_DF_DPHI + 2], W, st

    def predict(self, D2_full, phi2):
        return ops.gp_predict(self.layout, D2_full, self.y, self.kernel, self.phi3(phi2))

    def dist_bwd(self, Wm) -> torch.Tensor:
        """dL/dZ for dL/du = Wm of the UNWEIGHTED squared distances (the caller scales the columns by w)."""
        mu = torch.zeros(self.T, self.d, dtype=torch.float32, device=self.Z.device)
        return ops.gp_sqdist_bwd(self.Z, self.layout, Wm, mu, 1.0)

    def median_l0(self) -> torch.Tensor:
        g = gpk.BatchedExactGP(self.Z, self.y, self.layout, self.kernel, False, True, qq_block=False)
        return g.init().aux[:, ops.AUX_L0]


def _lognormal(x, loc):
    """log p, d/dx, d2/dx2 of LogNormal(loc, PRIOR_SCALE) elementwise (loc broadcasts)."""
    it2 = 1.0 / (PRIOR_SCALE * PRIOR_SCALE)
    lx = torch.log(x)
    lp = -lx - math.log(PRIOR_SCALE) - 0.5 * math.log(2 * math.pi) - 0.5 * (lx - loc) ** 2 * it2
    q = (lx - loc) * it2
    return lp, -(1.0 + q) / x, (1.0 + q - it2) / (x * x)


class BatchedARDExactGP:
    """T tasks' ARD GP: fit / predict / hypergradient.  `backend` is the kernel layer (the CUDA one unless a test injects another)."""

    def __init__(self, Z: torch.Tensor, y: torch.Tensor, layout: ops.TaskLayout, kernel: str = "matern",
                 use_numeric_labels: bool = False, use_lengthscale_prior: bool = True, backend=None):
        if kernel not in ("matern", "rbf", "RBF"):
            raise ValueError(f"ard_num_dims only exists for the stationary kernels (gp_utils.py:27,30), not '{kernel}'")
        if backend is None:
            if not Z.is_cuda:
                raise RuntimeError("adkf_ift_b200 kernels need CUDA tensors (no CPU fallback)")
            backend = _CudaBackend(Z.detach().float(), y.float(), layout, kernel, gpk.noise_prior_loc(use_numeric_labels))
        self.be = backend
        self.T, self.d = backend.T, backend.d
        self.noise_init = gpk.noise_init(use_numeric_labels)
        self.use_ls_prior = use_lengthscale_prior
        self.dtype = backend.Z.dtype
        self._l0 = None

    # ---- parameters ---------------------------------------------------------------------------------------------
    def init_phi(self) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """reinit_gp_params with use_ard: (phi0 [T,2+d], l0 [T], prior loc [T])."""
        l0 = self.be.median_l0().to(self.dtype)
        raw_l = l0 + torch.log(-torch.expm1(-l0))
        rn = math.log(math.expm1(self.noise_init - gpk.NOISE_FLOOR))
        phi = torch.cat([torch.full((self.T, 1), rn, dtype=self.dtype, device=l0.device),
                         torch.zeros(self.T, 1, dtype=self.dtype, device=l0.device), raw_l[:, None].expand(self.T, self.d)], 1)
        return phi.contiguous(), l0, torch.log(l0) + PRIOR_SCALE ** 2

    # ---- f, grad f -------------------------------------------------------------------------------------------------
    def value_and_grad(self, phi, ls_loc, keep=False):
        ell = F.softplus(phi[:, 2:])
        w = 1.0 / (ell * ell)
        D2 = self.be.wsqdist(w, True)
        f3, g2, W, st = self.be.inner(D2, phi[:, :2].contiguous())
        q = self.be.pairquad(W, True)
        n = self.be.n.to(self.dtype)[:, None]
        g_ell = q * (-2.0 / ell ** 3)
        # the objective is carried in fp64: with d = 2048 the summed lengthscale prior is ~200 of f while a line-search step
        # changes f by ~1e-5 -- below fp32 resolution at that magnitude (the kernels' data term is O(1) and stays fp32)
        f = f3.double()
        if self.use_ls_prior:
            lp, d1, _ = _lognormal(ell.double(), ls_loc.double()[:, None])
            f = f - lp.sum(1) / n[:, 0].double()
            g_ell = g_ell - d1.to(self.dtype) / n
        t = torch.sigmoid(phi[:, 2:])
        grad = torch.cat([g2.to(self.dtype), t * g_ell], 1)
        if keep:
            return f, grad, dict(D2=D2, W=W, q=q, ell=ell, g_ell=g_ell, st=st)
        return f, grad, st

    # ---- Hessian-vector product ----------------------------------------------------------------------------------------
    def hvp(self, phi, ls_loc, x, cache=None):
        """H x for x [T, 2+d]; H = d2f/dphi2 in raw space.  `cache` = the dict of value_and_grad(..., keep=True) at phi."""
        if cache is None:
            _, _, cache = self.value_and_grad(phi, ls_loc, keep=True)
        ell, q, g_ell = cache["ell"], cache["q"], cache["g_ell"]
        t = torch.sigmoid(phi[:, 2:])
        c = t * (1.0 - t)
        dl = t * x[:, 2:]
        dw = -2.0 * dl / ell ** 3
        dD2 = self.be.wsqdist(dw, True)
        dW, dg2 = self.be.inner_dir(cache["D2"], phi[:, :2].contiguous(), x[:, :2].contiguous(), dD2)
        dq = self.be.pairquad(dW, True)
        n = self.be.n.to(self.dtype)[:, None]
        dg_ell = dq * (-2.0 / ell ** 3) + q * (6.0 * dl / ell ** 4)
        if self.use_ls_prior:
            _, _, d2 = _lognormal(ell, ls_loc[:, None])
            dg_ell = dg_ell - d2 * dl / n
        return torch.cat([dg2.to(self.dtype), t * dg_ell + c * x[:, 2:] * g_ell], 1)

    # ---- inner fit: batched L-BFGS ----------------------------------------------------------------------------------------
    def fit(self, max_iter: int = 200, gtol: float = 2e-5, history: int = 10, start: Optional[torch.Tensor] = None) -> ARDFit:
        phi0, l0, ls_loc = self.init_phi()
        phi = phi0 if start is None else start.to(self.dtype).clone()
        f, g, st = self.value_and_grad(phi, ls_loc)
        S, Y, RHO = [], [], []
        done = torch.zeros(self.T, dtype=torch.bool, device=phi.device)
        it = 0
        for it in range(1, max_iter + 1):
            # two-loop recursion, all tasks in lock step (per-task scalars)
            qv = g.clone()
            alphas = []
            for s_, y_, r_ in zip(reversed(S), reversed(Y), reversed(RHO)):
                a = r_ * (s_ * qv).sum(1)
                alphas.append(a)
                qv = qv - a[:, None] * y_
            if S:
                gamma = (S[-1] * Y[-1]).sum(1) / (Y[-1] * Y[-1]).sum(1).clamp_min(1e-30)
                qv = qv * gamma[:, None]
            for (s_, y_, r_), a in zip(zip(S, Y, RHO), reversed(alphas)):
                b = r_ * (y_ * qv).sum(1)
                qv = qv + (a - b)[:, None] * s_
            p = -qv
            gp_ = (g * p).sum(1)
            bad = gp_ >= 0                                                  # not a descent direction: steepest descent for that task
            p = torch.where(bad[:, None], -g, p)
            gp_ = torch.where(bad, -(g * g).sum(1), gp_)
            step = torch.ones(self.T, dtype=self.dtype, device=phi.device)
            gp_ = gp_.double()
            if not S:
                step = step / g.abs().sum(1).clamp_min(1.0)
            accepted = done.clone()
            f_new, g_new, phi_new = f.clone(), g.clone(), phi.clone()
            for _ in range(20):                                                # Armijo backtracking, per task
                trial = phi + step[:, None] * p
                ft, gt, _ = self.value_and_grad(trial, ls_loc)
                ok = (~accepted) & torch.isfinite(ft) & (ft <= f + 1e-4 * step.double() * gp_)
                f_new = torch.where(ok, ft, f_new)
                g_new = torch.where(ok[:, None], gt, g_new)
                phi_new = torch.where(ok[:, None], trial, phi_new)
                accepted = accepted | ok
                if bool(accepted.all()):
                    break
                step = torch.where(accepted, step, step * 0.5)
            s_k, y_k = phi_new - phi, g_new - g
            sy = (s_k * y_k).sum(1)
            good = sy > 1e-10 * (y_k * y_k).sum(1).clamp_min(1e-30)
            if bool(good.any()):
                S.append(torch.where(good[:, None], s_k, torch.zeros_like(s_k)))
                Y.append(torch.where(good[:, None], y_k, torch.zeros_like(y_k)))
                RHO.append(torch.where(good, 1.0 / sy.clamp_min(1e-30), torch.zeros_like(sy)))
                if len(S) > history:
                    S.pop(0); Y.pop(0); RHO.pop(0)
            stalled = ~accepted | ((f - f_new).abs() <= 1e-12 * f.abs().clamp_min(1.0))
            phi, f, g = phi_new, f_new, g_new
            done = done | (g.abs().max(1).values <= gtol) | stalled
            if bool(done.all()):
                break
        status = st | torch.where(g.abs().max(1).values <= max(gtol, 5e-5), 0, ops.ST_NOT_CONVERGED).to(torch.int32)
        return ARDFit(phi, f, g, ls_loc, l0, it, status)

    # ---- matrix-free inverse-Hessian-vector product ---------------------------------------------------------------------------
    def ihvp(self, phi, ls_loc, b, method: str = "cg", max_iter: int = 600, rtol: float = 1e-6, cache=None):
        """v with H v = b.  "cg": conjugate gradients (H is symmetric positive definite at an inner optimum: the priors bound
        its spectrum from below).  "neumann": v = a sum_j (I - a H)^j b with a = 1 / lambda_max (power iteration), the series of
        the north star -- linear convergence at rate 1 - lambda_min / lambda_max, so it is the slow option.
        Returns (v, iterations, relative residual [T])."""
        if cache is None:
            _, _, cache = self.value_and_grad(phi, ls_loc, keep=True)
        H = lambda x: self.hvp(phi, ls_loc, x, cache)          # noqa: E731
        bn = b.norm(dim=1).clamp_min(1e-30)
        if method == "cg":
            v = torch.zeros_like(b)
            r = b.clone()
            p = r.clone()
            rs = (r * r).sum(1)
            k = 0
            for k in range(1, max_iter + 1):
                Hp = H(p)
                alpha = rs / (p * Hp).sum(1).clamp_min(1e-30)
                v = v + alpha[:, None] * p
                r = r - alpha[:, None] * Hp
                rs_new = (r * r).sum(1)
                if k % 8 == 0 and bool((rs_new.sqrt() <= rtol * bn).all()):
                    rs = rs_new
                    break
                p = r + (rs_new / rs.clamp_min(1e-30))[:, None] * p
                rs = rs_new
            return v, k, rs.sqrt() / bn
        if method == "neumann":
            x = torch.randn(b.shape, generator=torch.Generator(device="cpu").manual_seed(0), dtype=torch.float32).to(b)
            lam = torch.ones(self.T, dtype=b.dtype, device=b.device)
            for _ in range(30):                                                  # power iteration for lambda_max
                hx = H(x)
                lam = hx.norm(dim=1) / x.norm(dim=1).clamp_min(1e-30)
                x = hx / hx.norm(dim=1, keepdim=True).clamp_min(1e-30)
            a = (1.0 / (1.05 * lam))[:, None]
            term = a * b
            v = term.clone()
            k = 0
            for k in range(1, max_iter + 1):
                term = term - a * H(term)
                v = v + term
                if k % 16 == 0 and bool((term.norm(dim=1) <= rtol * v.norm(dim=1)).all()):
                    break
            return v, k, (H(v) - b).norm(dim=1) / bn
        raise ValueError("method must be 'cg' or 'neumann'")

    # ---- outer loss and hypergradient w.r.t. the features ------------------------------------------------------------------------
    def hypergradient(self, fit: ARDFit, ignore_grad_correction: bool = False, method: str = "cg", max_iter: int = 600,
                      rtol: float = 1e-6):
        """(dL/dZ [rows,d], info): dL/dZ = dF/dZ - d/deps df/dZ(phi + eps v), v = H^-1 dF/dphi."""
        phi, ls_loc = fit.phi, fit.ls_loc
        ell = F.softplus(phi[:, 2:])
        w = 1.0 / (ell * ell)
        t = torch.sigmoid(phi[:, 2:])
        D2_full = self.be.wsqdist(w, False)
        Fv, dF2, WF, stF = self.be.outer(D2_full, phi[:, :2].contiguous())
        dF_ell = self.be.pairquad(WF, False) * (-2.0 / ell ** 3)
        dF_phi = torch.cat([dF2.to(self.dtype), t * dF_ell], 1)
        rows_w = w[self.be.rows_task]
        dZ = self.be.dist_bwd(WF).to(self.dtype) * rows_w
        info = dict(F=Fv, dF_dphi=dF_phi, status=stF, iterations=0)
        if not ignore_grad_correction:
            _, _, cache = self.value_and_grad(phi, ls_loc, keep=True)
            v, iters, res = self.ihvp(phi, ls_loc, dF_phi, method=method, max_iter=max_iter, rtol=rtol, cache=cache)
            dl = t * v[:, 2:]
            dw = -2.0 * dl / ell ** 3
            dD2 = self.be.wsqdist(dw, True)
            dW, _ = self.be.inner_dir(cache["D2"], phi[:, :2].contiguous(), v[:, :2].contiguous(), dD2)
            mixed = self.be.dist_bwd(dW).to(self.dtype) * rows_w + self.be.dist_bwd(cache["W"]).to(self.dtype) * dw[self.be.rows_task]
            dZ = dZ - mixed * self.be.support_rows[:, None]
            info.update(v=v, iterations=iters, residual=res)
        return dZ, info

    def predict(self, phi):
        ell = F.softplus(phi[:, 2:])
        D2_full = self.be.wsqdist(1.0 / (ell * ell), False)
        return self.be.predict(D2_full, phi[:, :2].contiguous())
