"""oracle/gp_closed_form.py -- ORACLE / TEST INFRASTRUCTURE ONLY.

numpy (float64) restatement of the CLOSED-FORM algebra the fused CUDA GP kernels implement
(SURVEY.md App. B): -MLL f, its gradient and 3x3 Hessian w.r.t. the raw parameters, the predictive
joint NLL F through the joint-covariance identity  F = NLL([y_s;y_q]; J) - NLL(y_s; A), dF/dphi,
v = H^-1 dF/dphi and the symmetric weight matrix  Wm = dF/dD2 - d/deps[df/dD2(phi + eps v)].
Checked against the autograd oracle (oracle/adkf_oracle.py) in tests/test_oracle.py; the CUDA kernels
are checked against both.  Also holds a numpy mirror of the on-device damped-Newton inner fit.
"""
import math

import numpy as np

SQRT5 = math.sqrt(5.0)
TAU = 0.25  # LogNormal prior scale (fs_mol/models/adaptive_dkt.py:95,112,116)


def softplus(x):
    return np.log1p(np.exp(-abs(x))) + max(x, 0.0)


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def inv_softplus(x):
    return x + np.log(-np.expm1(-x))


def sq_dist_centered(Z, n_sup):
    mu = Z[:n_sup].mean(0, keepdims=True)
    Zc = Z - mu
    nrm = (Zc * Zc).sum(1)
    D2 = nrm[:, None] + nrm[None, :] - 2.0 * Zc @ Zc.T
    D2 = np.maximum(D2, 0.0)
    np.fill_diagonal(D2, 0.0)
    return D2


def kernel_parts(D2, s, l, kind):
    """K, K_u = dK/dD2, K_l = dK/dl, K_ll, K_ul  (elementwise)."""
    if kind == "matern":
        r = np.sqrt(np.maximum(D2 / (l * l), 1e-30))
        e = np.exp(-SQRT5 * r)
        h = 1.0 + SQRT5 * r
        K = s * (h + 5.0 / 3.0 * r * r) * e
        Ku = -s * (5.0 / (6.0 * l * l)) * h * e
        Kl = s * (5.0 / 3.0) * (r * r / l) * h * e
        Kll = -(5.0 * s / (3.0 * l * l)) * e * r * r * (3.0 + 3.0 * SQRT5 * r - 5.0 * r * r)
        Kul = (5.0 * s / (6.0 * l ** 3)) * e * (2.0 + 2.0 * SQRT5 * r - 5.0 * r * r)
    elif kind == "rbf":
        k = np.exp(-D2 / (2.0 * l * l))
        K = s * k
        Ku = -K / (2.0 * l * l)
        Kl = K * D2 / l ** 3
        Kll = K * (D2 * D2 / l ** 6 - 3.0 * D2 / l ** 4)
        Kul = (K / l ** 3) * (1.0 - D2 / (2.0 * l * l))
    else:
        raise ValueError(kind)
    return K, Ku, Kl, Kll, Kul


def lognormal_terms(x, loc):
    """log p, d/dx, d2/dx2 of LogNormal(loc, TAU)."""
    lx = math.log(x)
    lp = -lx - math.log(TAU) - 0.5 * math.log(2 * math.pi) - (lx - loc) ** 2 / (2 * TAU * TAU)
    q = (lx - loc) / (TAU * TAU)
    d1 = -(1.0 + q) / x
    d2 = (1.0 + q - 1.0 / (TAU * TAU)) / (x * x)
    return lp, d1, d2


def transforms(phi):
    sig2 = softplus(phi[0]) + 1e-4
    s = softplus(phi[1])
    l = softplus(phi[2])
    t = np.array([sigmoid(phi[0]), sigmoid(phi[1]), sigmoid(phi[2])])
    c = t * (1.0 - t)
    return sig2, s, l, t, c


def inner_all(D2ss, y, phi, kind, noise_loc, ls_loc, want_hess=True):
    """f, grad (raw), Hessian (raw) and the pieces (P, alpha, ...) of the inner objective."""
    n = len(y)
    sig2, s, l, t, c = transforms(phi)
    K, Ku, S, S2, Kul = kernel_parts(D2ss, s, l, kind)
    A = K + sig2 * np.eye(n)
    L = np.linalg.cholesky(A)
    P = np.linalg.inv(A)
    alpha = P @ y
    logdet = 2.0 * np.log(np.diag(L)).sum()
    nll = 0.5 * y @ alpha + 0.5 * logdet + 0.5 * n * math.log(2 * math.pi)
    lp_n, d1_n, d2_n = lognormal_terms(sig2, noise_loc)
    if ls_loc is not None:
        lp_l, d1_l, d2_l = lognormal_terms(l, ls_loc)
    else:
        lp_l = d1_l = d2_l = 0.0
    f = (nll - lp_n - lp_l) / n
    trP = np.trace(P)
    aa = alpha @ alpha
    ya = y @ alpha
    u = S @ alpha
    w = P @ alpha
    trPS = (P * S).sum()
    aSa = alpha @ u
    g_psi = np.array([
        0.5 * trP - 0.5 * aa - d1_n,
        0.5 * (n - sig2 * trP) / s - 0.5 * (ya - sig2 * aa) / s,
        0.5 * trPS - 0.5 * aSa - d1_l,
    ]) / n
    out = dict(f=f, g=t * g_psi, P=P, alpha=alpha, S=S, Ku=Ku, Kul=Kul, K=K, sig2=sig2, s=s, l=l, t=t,
               nll=nll, logdet=logdet)
    if want_hess:
        PS = P @ S
        trPP = (P * P).sum()
        aPa = alpha @ w
        trPPS = (PS * P).sum()
        trPSPS = (PS * PS.T).sum()
        trPS2 = (P * S2).sum()
        wu = w @ u
        uPu = u @ P @ u
        aS2a = alpha @ S2 @ alpha
        Hp = np.zeros((3, 3))
        Hp[0, 0] = -0.5 * trPP + aPa - d2_n
        Hp[0, 1] = Hp[1, 0] = -0.5 * (trP - sig2 * trPP) / s + (aa - sig2 * aPa) / s
        Hp[0, 2] = Hp[2, 0] = -0.5 * trPPS + wu
        Hp[1, 1] = -0.5 * (n - 2 * sig2 * trP + sig2 ** 2 * trPP) / s ** 2 + (ya - 2 * sig2 * aa + sig2 ** 2 * aPa) / s ** 2
        Hp[1, 2] = Hp[2, 1] = 0.5 * sig2 * trPPS / s + (0.5 * aSa - sig2 * wu) / s
        Hp[2, 2] = -0.5 * trPSPS + 0.5 * trPS2 + uPu - 0.5 * aS2a - d2_l
        Hp /= n
        out["H"] = Hp * np.outer(t, t) + np.diag(c * g_psi)
    return out


def predict(D2, n, y_s, phi, kind):
    sig2, s, l, _, _ = transforms(phi)
    K = kernel_parts(D2, s, l, kind)[0]
    A = K[:n, :n] + sig2 * np.eye(n)
    Kqs = K[n:, :n]
    alpha = np.linalg.solve(A, y_s)
    mean = Kqs @ alpha
    V = np.linalg.solve(np.linalg.cholesky(A), Kqs.T)
    var = s + sig2 - (V * V).sum(0)
    return mean, var


def ift_all(D2, n, y, phi, kind, noise_loc, ls_loc, ignore_grad_correction=False):
    """F, dF/dphi (raw), H, v and Wm over the joint N x N layout (support rows first)."""
    N = len(y)
    inner = inner_all(D2[:n, :n], y[:n], phi, kind, noise_loc, ls_loc)
    sig2, s, l, t = inner["sig2"], inner["s"], inner["l"], inner["t"]
    P, alpha = inner["P"], inner["alpha"]
    K, Ku, Kl, _, Kul = kernel_parts(D2, s, l, kind)
    J = K + sig2 * np.eye(N)
    Jinv = np.linalg.inv(J)
    beta = Jinv @ y
    logdetJ = np.linalg.slogdet(J)[1]
    nllJ = 0.5 * y @ beta + 0.5 * logdetJ + 0.5 * N * math.log(2 * math.pi)
    F = nllJ - inner["nll"]
    Gam = 0.5 * (P - np.outer(alpha, alpha))
    G = 0.5 * (Jinv - np.outer(beta, beta))
    G[:n, :n] -= Gam
    dF_psi = np.array([np.trace(G), (G * K).sum() / s, (G * Kl).sum()])
    dF = t * dF_psi
    H = inner["H"]
    Wm = G * Ku
    v = np.linalg.solve(H, dF)
    if not ignore_grad_correction:
        pd = t * v  # direction in (sigma^2, s, l) space
        Adot = pd[0] * np.eye(n) + pd[1] * K[:n, :n] / s + pd[2] * Kl[:n, :n]
        X = P @ Adot
        Pdot = -X @ P
        adot = -X @ alpha
        Gamdot = 0.5 * (Pdot - np.outer(adot, alpha) - np.outer(alpha, adot))
        Kudot = pd[1] * Ku[:n, :n] / s + pd[2] * Kul[:n, :n]
        Wm[:n, :n] -= (Gamdot * Ku[:n, :n] + Gam * Kudot) / n
    return dict(F=F, dF=dF, H=H, v=v, Wm=Wm, f=inner["f"], g=inner["g"])


def dz_from_wm(Z, n_sup, Wm):
    mu = Z[:n_sup].mean(0, keepdims=True)
    Zc = Z - mu
    return 4.0 * (Wm.sum(1)[:, None] * Zc - Wm @ Zc)


def median_lengthscale(D2ss):
    n = D2ss.shape[0]
    iu = np.triu_indices(n, 1)
    vals = np.sort(D2ss[iu][D2ss[iu] > 0.0])
    med = vals[(len(vals) - 1) // 2]  # torch.median = lower median
    return math.sqrt(0.5 * med)


def newton_fit(D2ss, y, kind, noise_init, noise_loc, use_ls_prior=True, max_iter=50, gtol=1e-6,
               dtype=np.float64, phi0=None):
    """numpy mirror of the on-device damped Newton fit (adkf_gp_fit).  Returns phi*, info."""
    l0 = median_lengthscale(D2ss)
    ls_loc = math.log(l0) + TAU * TAU if use_ls_prior else None
    phi = np.array([inv_softplus(noise_init - 1e-4), 0.0, inv_softplus(l0)]) if phi0 is None else np.array(phi0, float)
    lam = 0.0
    nfev = 0
    cur = inner_all(D2ss, y, phi, kind, noise_loc, ls_loc)
    nfev += 1
    it = 0
    for it in range(max_iter):
        g, H = cur["g"], cur["H"]
        if np.abs(g).max() <= gtol:
            break
        accepted = False
        for _ in range(12):
            Hd = H + lam * np.eye(3)
            try:
                np.linalg.cholesky(Hd)
            except np.linalg.LinAlgError:
                lam = max(4.0 * lam, 1e-3 * max(1.0, np.abs(np.diag(H)).max()))
                continue
            p = -np.linalg.solve(Hd, g)
            mx = np.abs(p).max()
            if mx > 2.0:
                p *= 2.0 / mx
            try:
                trial = inner_all(D2ss, y, phi + p, kind, noise_loc, ls_loc)
                nfev += 1
                ok = np.isfinite(trial["f"]) and trial["f"] <= cur["f"] + 1e-4 * (g @ p)
            except np.linalg.LinAlgError:
                ok = False
            if ok:
                phi = phi + p
                cur = trial
                lam = lam * 0.25 if lam > 1e-8 else 0.0
                accepted = True
                break
            lam = max(4.0 * lam, 1e-3 * max(1.0, np.abs(np.diag(H)).max()))
        if not accepted:
            break
    return phi, dict(f=cur["f"], g=cur["g"], H=cur["H"], iters=it, nfev=nfev, l0=l0, ls_loc=ls_loc)
