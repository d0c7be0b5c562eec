"""oracle/gp_ard_backend.py -- ORACLE / TEST INFRASTRUCTURE ONLY.

A plain-torch (fp64, autograd) stand-in for the kernel layer behind adkf_ift_b200/gp_ard.py (`_CudaBackend`): the same seven
operations written from their definitions, so that the HOST algebra of the ARD path (gradient assembly, Hessian-vector product,
L-BFGS, CG / Neumann inverse-Hessian-vector product, hypergradient) can be checked on the CPU against the golden produced by the
reference's own GP layer (tests/golden/gp_ard.npz), independently of the CUDA kernels.  The product never imports this."""
import math

import torch

NOISE_FLOOR = 1e-4
TAU = 0.25


def _kernel(D2, s, kind):
    if kind == "matern":
        r = torch.sqrt(D2.clamp_min(1e-30))
        return s * (1.0 + math.sqrt(5.0) * r + 5.0 / 3.0 * r * r) * torch.exp(-math.sqrt(5.0) * r)
    return s * torch.exp(-0.5 * D2)


def _nll(K, y):
    L = torch.linalg.cholesky(K)
    a = torch.cholesky_solve(y[:, None], L)[:, 0]
    return 0.5 * (y * a).sum() + torch.log(torch.diagonal(L)).sum() + 0.5 * len(y) * math.log(2 * math.pi)


class TorchBackend:
    def __init__(self, Z, y, n_sup, n_rows, kernel, noise_loc):
        self.Z, self.y, self.kernel, self.noise_loc = Z.double(), y.double(), ("rbf" if kernel in ("rbf", "RBF") else "matern"), noise_loc
        self.n_sup, self.n_rows = list(n_sup), list(n_rows)
        self.T, self.d = len(n_sup), Z.shape[1]
        self.z_off = [0]
        self.d2_off = [0]
        for r in self.n_rows:
            self.z_off.append(self.z_off[-1] + r)
            self.d2_off.append(self.d2_off[-1] + r * r)
        self.rows_task = torch.repeat_interleave(torch.arange(self.T), torch.tensor(self.n_rows))
        self.support_rows = torch.cat([torch.cat([torch.ones(s, dtype=torch.bool), torch.zeros(r - s, dtype=torch.bool)])
                                       for s, r in zip(self.n_sup, self.n_rows)])
        self.n = torch.tensor([float(s) for s in self.n_sup], dtype=torch.float64)

    def _zt(self, t):
        return self.Z[self.z_off[t]:self.z_off[t + 1]]

    def _mat(self, flat, t):
        N = self.n_rows[t]
        return flat[self.d2_off[t]:self.d2_off[t + 1]].view(N, N)

    def wsqdist(self, c, support_only):
        out = torch.zeros(self.d2_off[-1], dtype=torch.float64)
        for t in range(self.T):
            z = self._zt(t)
            R = self.n_sup[t] if support_only else self.n_rows[t]
            diff = z[:R, None, :] - z[None, :R, :]
            self._mat(out, t)[:R, :R] = (diff * diff * c[t][None, None, :]).sum(-1)
        return out

    def pairquad(self, G, support_only):
        q = torch.zeros(self.T, self.d, dtype=torch.float64)
        for t in range(self.T):
            z = self._zt(t)
            R = self.n_sup[t] if support_only else self.n_rows[t]
            diff = z[:R, None, :] - z[None, :R, :]
            q[t] = (self._mat(G, t)[:R, :R][:, :, None] * diff * diff).sum((0, 1))
        return q

    def _f3(self, D2ss, y, phi2):
        sig2 = torch.nn.functional.softplus(phi2[0]) + NOISE_FLOOR
        s = torch.nn.functional.softplus(phi2[1])
        n = len(y)
        K = _kernel(D2ss, s, self.kernel) + sig2 * torch.eye(n, dtype=torch.float64)
        lx = torch.log(sig2)
        lp = -lx - math.log(TAU) - 0.5 * math.log(2 * math.pi) - (lx - self.noise_loc) ** 2 / (2 * TAU * TAU)
        return (_nll(K, y) - lp) / n

    def inner(self, D2, phi2):
        f, g2, W = [], [], torch.zeros_like(D2)
        for t in range(self.T):
            n = self.n_sup[t]
            D = self._mat(D2, t)[:n, :n].clone().requires_grad_(True)
            p = phi2[t].clone().requires_grad_(True)
            val = self._f3(D, self.y[self.z_off[t]:self.z_off[t] + n], p)
            gp, gD = torch.autograd.grad(val, [p, D])
            f.append(val.detach()); g2.append(gp)
            self._mat(W, t)[:n, :n] = gD
        return torch.stack(f), torch.stack(g2), W, torch.zeros(self.T, dtype=torch.int32)

    def inner_dir(self, D2, phi2, u2, dD2):
        dW, dg2 = torch.zeros_like(D2), []
        for t in range(self.T):
            n = self.n_sup[t]
            y = self.y[self.z_off[t]:self.z_off[t] + n]
            D0 = self._mat(D2, t)[:n, :n].clone()
            Dd = self._mat(dD2, t)[:n, :n].clone()

            def grads(p, D):
                val = self._f3(D, y, p)
                return torch.autograd.grad(val, [p, D], create_graph=True)

            _, (tg, tW) = torch.autograd.functional.jvp(lambda p, D: grads(p, D), (phi2[t].clone(), D0), (u2[t].clone(), Dd))
            dg2.append(tg)
            self._mat(dW, t)[:n, :n] = tW
        return dW, torch.stack(dg2)

    def outer(self, D2_full, phi2):
        Fv, dF2, W = [], [], torch.zeros_like(D2_full)
        for t in range(self.T):
            n, N = self.n_sup[t], self.n_rows[t]
            D = self._mat(D2_full, t).clone().requires_grad_(True)
            p = phi2[t].clone().requires_grad_(True)
            sig2 = torch.nn.functional.softplus(p[0]) + NOISE_FLOOR
            s = torch.nn.functional.softplus(p[1])
            y = self.y[self.z_off[t]:self.z_off[t + 1]]
            J = _kernel(D, s, self.kernel) + sig2 * torch.eye(N, dtype=torch.float64)
            val = _nll(J, y) - _nll(J[:n, :n], y[:n])
            gp, gD = torch.autograd.grad(val, [p, D])
            Fv.append(val.detach()); dF2.append(gp)
            self._mat(W, t)[:, :] = gD
        return torch.stack(Fv), torch.stack(dF2), W, torch.zeros(self.T, dtype=torch.int32)

    def predict(self, D2_full, phi2):
        mean = torch.zeros(self.z_off[-1], dtype=torch.float64)
        var = torch.zeros_like(mean)
        for t in range(self.T):
            n, N = self.n_sup[t], self.n_rows[t]
            D = self._mat(D2_full, t)
            sig2 = torch.nn.functional.softplus(phi2[t, 0]) + NOISE_FLOOR
            s = torch.nn.functional.softplus(phi2[t, 1])
            K = _kernel(D, s, self.kernel)
            A = K[:n, :n] + sig2 * torch.eye(n, dtype=torch.float64)
            y = self.y[self.z_off[t]:self.z_off[t] + n]
            sol = torch.linalg.solve(A, torch.cat([y[:, None], K[:n, n:]], 1))
            mean[self.z_off[t] + n:self.z_off[t + 1]] = K[n:, :n] @ sol[:, 0]
            var[self.z_off[t] + n:self.z_off[t + 1]] = s - (K[n:, :n] * sol[:, 1:].T).sum(1) + sig2
        return mean, var, torch.zeros(self.T, dtype=torch.int32)

    def dist_bwd(self, Wm):
        out = torch.zeros_like(self.Z)
        for t in range(self.T):
            z = self._zt(t)
            W = self._mat(Wm, t)
            out[self.z_off[t]:self.z_off[t + 1]] = 4.0 * (W.sum(1)[:, None] * z - W @ z)
        return out

    def median_l0(self):
        out = []
        for t in range(self.T):
            z = self._zt(t)[:self.n_sup[t]]
            d2 = torch.triu(torch.cdist(z, z) ** 2, diagonal=1)
            out.append(torch.sqrt(0.5 * torch.median(d2[d2 > 0.0])))
        return torch.stack(out)
