"""-m gpu parity tests of the batched GP kernels (through the C-ABI) against the CPU oracle and the golden
fixtures generated from the reference.  Tolerances: 1e-4 relative (north_star) at FIXED phi."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import gp_closed_form as cf

pytestmark = pytest.mark.gpu

FIXTURES = [("c1_classification", "matern"), ("c1_regression", "matern"), ("s64_rbf_l2", "rbf")]


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def load_task(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    Zs, Zq = g["z_support"].astype(np.float32), g["z_query"].astype(np.float32)
    Z = np.concatenate([Zs, Zq])
    y = np.concatenate([g["y_support"], g["y_query"]]).astype(np.float32)
    return g, Z, y, len(Zs), len(Zq)


def stack_tasks(tasks, device):
    from adkf_ift_b200 import ops
    layout = ops.TaskLayout.build([t[2] for t in tasks], [t[3] for t in tasks], device)
    Z = torch.from_numpy(np.concatenate([t[0] for t in tasks])).to(device)
    y = torch.from_numpy(np.concatenate([t[1] for t in tasks])).to(device)
    return layout, Z, y


@pytest.mark.parametrize("name,kind", FIXTURES)
def test_sqdist_matches_oracle(cuda_device, golden_dir, name, kind):
    from adkf_ift_b200 import ops
    g, Z, y, n, m = load_task(golden_dir, name)
    layout, Zd, yd = stack_tasks([(Z, y, n, m)], cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=True)
    D2 = sq.D2.cpu().numpy().reshape(n + m, n + m)
    ref = cf.sq_dist_centered(Z.astype(np.float64), n)
    assert rel(D2, ref) < 2e-6
    assert np.all(np.diag(D2) == 0.0)
    assert np.array_equal(D2, D2.T)


@pytest.mark.parametrize("name,kind", FIXTURES)
def test_fixed_phi_inner_and_ift(cuda_device, golden_dir, name, kind):
    """f, grad, Hessian, F, dF/dphi, v and dL/dZ at the reference's phi* (golden: reference autograd)."""
    from adkf_ift_b200 import ops
    g, Z, y, n, m = load_task(golden_dir, name)
    numeric = bool(g["meta"][3])
    layout, Zd, yd = stack_tasks([(Z, y, n, m)], cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=True)
    phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3).to(cuda_device)
    locs = g["fixed_f64_prior_locs"]
    phi_o, aux, st = ops.gp_fit(layout, sq.D2, yd, kind, 0.01 if numeric else 0.1, float(locs[0]), ls_prior=1,
                                init_mode=0, max_iter=0, phi=phi)
    a = aux.cpu().numpy()[0]
    assert int(st.item()) & ops.ST_NOT_PD == 0
    # every tolerance below is max(1e-4, 3 x the reference's own fp32 error against its fp64 run) -- the north star's 1e-4
    # wherever the reference itself achieves it
    def tol(key):
        return max(1e-4, 3 * rel(np.atleast_1d(g["fixed_f32_" + key]), np.atleast_1d(g["fixed_f64_" + key])))
    assert abs(a[ops.AUX_F] - g["fixed_f64_f"]) < 1e-4 * abs(g["fixed_f64_f"])
    e_H = rel(a[ops.AUX_H:ops.AUX_H + 9].reshape(3, 3), g["fixed_f64_H"])
    print(name, "inner Hessian rel err", e_H, "tol", tol("H"))
    assert e_H < tol("H")
    assert abs(a[ops.AUX_LS_LOC] - locs[1]) < 1e-5
    Wm, aux2, st2 = ops.gp_ift(layout, sq.D2, yd, kind, float(locs[0]), 1, phi, aux)
    b = aux2.cpu().numpy()[0]
    assert abs(b[ops.AUX_FOUTER] - g["fixed_f64_F"]) < 1e-4 * abs(g["fixed_f64_F"])
    e_dF = rel(b[ops.AUX_DF_DPHI:ops.AUX_DF_DPHI + 3], g["fixed_f64_dF_dphi"])
    e_v = rel(b[ops.AUX_V:ops.AUX_V + 3], g["fixed_f64_v"])
    print(name, "dF/dphi rel err", e_dF, "tol", tol("dF_dphi"), " v = H^-1 dF/dphi rel err", e_v, "tol", tol("v"))
    assert e_dF < tol("dF_dphi")
    assert e_v < tol("v")
    dZ = ops.gp_sqdist_bwd(Zd, layout, Wm, sq.mu, 1.0).cpu().numpy()
    ref = np.concatenate([g["fixed_f64_dL_dzs"], g["fixed_f64_dL_dzq"]])
    ref32 = np.concatenate([g["fixed_f32_dL_dzs"], g["fixed_f32_dL_dzq"]])
    err, base = rel(dZ, ref), rel(ref32, ref)
    print(name, "dL/dZ rel err cuda", err, "reference-fp32-vs-fp64", base)
    assert err < max(1e-4, 3 * base)


@pytest.mark.parametrize("name,kind", FIXTURES)
def test_predict_fixed_phi(cuda_device, golden_dir, name, kind):
    from adkf_ift_b200 import ops
    g, Z, y, n, m = load_task(golden_dir, name)
    layout, Zd, yd = stack_tasks([(Z, y, n, m)], cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=False)
    phi = torch.from_numpy(g["test_phi_star"].astype(np.float32)).view(1, 3).to(cuda_device)
    mean, var, st = ops.gp_predict(layout, sq.D2, yd, kind, phi)
    assert rel(mean.cpu().numpy()[n:], g["test_mean"]) < 1e-4
    assert rel(var.cpu().numpy()[n:], g["test_variance"]) < 1e-4


@pytest.mark.parametrize("name,kind", FIXTURES)
def test_fit_reaches_reference_optimum(cuda_device, golden_dir, name, kind):
    """After-fit parity at each side's own optimum: <=1e-3 on phi*, and a smaller inner gradient than SciPy."""
    from adkf_ift_b200 import ops
    g, Z, y, n, m = load_task(golden_dir, name)
    numeric = bool(g["meta"][3])
    layout, Zd, yd = stack_tasks([(Z, y, n, m)], cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=False)
    locs = g["fixed_f64_prior_locs"]
    phi, aux, st = ops.gp_fit(layout, sq.D2, yd, kind, 0.01 if numeric else 0.1, float(locs[0]))
    a = aux.cpu().numpy()[0]
    p = phi.cpu().numpy()[0]
    print(name, "phi*", p, "ref", g["test_phi_star"], "iters", a[ops.AUX_ITERS], "nfev", a[ops.AUX_NFEV],
          "|g|", np.abs(a[ops.AUX_G:ops.AUX_G + 3]).max(), "status", int(st.item()))
    assert int(st.item()) & (ops.ST_NOT_PD | ops.ST_NOT_CONVERGED) == 0
    assert np.abs(p - g["test_phi_star"]).max() < 2e-3
    assert np.abs(a[ops.AUX_G:ops.AUX_G + 3]).max() < np.abs(g["fixed_f64_g_inner"]).max() + 5e-6
    assert abs(a[ops.AUX_L0] - math.exp(locs[1] - 0.0625)) < 1e-5 * a[ops.AUX_L0]
    mean, var, _ = ops.gp_predict(layout, sq.D2, yd, kind, phi)
    assert rel(mean.cpu().numpy()[n:], g["test_mean"]) < 2e-3
    assert rel(var.cpu().numpy()[n:], g["test_variance"]) < 2e-3


def test_batched_ragged_tasks(cuda_device):
    """T ragged tasks in one launch == the closed-form oracle per task (incl. n = 128 and a global-ws size)."""
    from adkf_ift_b200 import ops
    rng = np.random.default_rng(0)
    sizes = [(16, 32), (64, 256), (128, 192), (200, 40), (24, 7)]
    tasks = []
    for (n, m) in sizes:
        c = rng.normal(size=(3, 96))
        lab = rng.integers(0, 3, size=n + m)
        Z = (c[lab] + 0.7 * rng.normal(size=(n + m, 96))).astype(np.float32)
        y = (np.array([1.0, -1.0, 0.2])[lab] + 0.3 * rng.normal(size=n + m)).astype(np.float32)
        tasks.append((Z, y, n, m))
    layout, Zd, yd = stack_tasks(tasks, cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=True)
    noise_loc = math.log(0.1) + 0.0625
    phi, aux, st = ops.gp_fit(layout, sq.D2, yd, "matern", 0.1, noise_loc)
    Wm, aux2, st2 = ops.gp_ift(layout, sq.D2, yd, "matern", noise_loc, 1, phi, aux)
    dZ = ops.gp_sqdist_bwd(Zd, layout, Wm, sq.mu, 1.0).cpu().numpy()
    mean, var, _ = ops.gp_predict(layout, sq.D2, yd, "matern", phi)
    phi_h, aux_h, mean_h, var_h = phi.cpu().numpy(), aux2.cpu().numpy(), mean.cpu().numpy(), var.cpu().numpy()
    off = 0
    for t, (Z, y, n, m) in enumerate(tasks):
        assert int(st[t].item()) & (ops.ST_NOT_PD | ops.ST_NOT_CONVERGED) == 0, (t, int(st[t].item()))
        D2 = cf.sq_dist_centered(Z.astype(np.float64), n)
        p_ref, info = cf.newton_fit(D2[:n, :n], y[:n].astype(np.float64), "matern", 0.1, noise_loc)
        assert np.abs(phi_h[t] - p_ref).max() < 2e-3, (t, phi_h[t], p_ref)
        r = cf.ift_all(D2, n, y.astype(np.float64), phi_h[t].astype(np.float64), "matern", noise_loc, info["ls_loc"])
        assert abs(aux_h[t, ops.AUX_FOUTER] - r["F"]) < 2e-4 * abs(r["F"]), (t, aux_h[t, ops.AUX_FOUTER], r["F"])
        ref_dz = cf.dz_from_wm(Z.astype(np.float64), n, r["Wm"])
        e = rel(dZ[off:off + n + m], ref_dz)
        print("task", t, (n, m), "dZ rel", e, "H cond", np.linalg.cond(r["H"]))
        assert e < 2e-3
        mr, vr = cf.predict(D2, n, y[:n].astype(np.float64), phi_h[t].astype(np.float64), "matern")
        assert rel(mean_h[off + n:off + n + m], mr) < 2e-4
        assert rel(var_h[off + n:off + n + m], vr) < 2e-4
        off += n + m


KERNEL_FAMILIES = ["matern", "rbf", "linear", "poli1", "poli2", "cossim", "tanimoto"]


def _family_inputs(g, kern, device):
    from adkf_ift_b200 import ops
    n = int(g["n"])
    if kern == "tanimoto":
        X = g["fp"].astype(np.float32)
    else:
        X = g["X"].astype(np.float32)
        if kern == "cossim":   # adaptive_dkt.py:158-160: features are L2-normalised for cossim
            X = X / np.maximum(np.linalg.norm(X, axis=1, keepdims=True), 1e-12)
    y = g["y"].astype(np.float32)
    m = len(y) - n
    layout = ops.TaskLayout.build([n], [m], device)
    phi = np.zeros(3, np.float32)
    p = g[f"{kern}_phi"].astype(np.float32)
    phi[:len(p)] = p
    return X, y, n, m, layout, torch.from_numpy(X).to(device), torch.from_numpy(y).to(device), phi


@pytest.mark.parametrize("kern", KERNEL_FAMILIES)
@pytest.mark.parametrize("impl", ["tcgen05", "simt"])
def test_kernel_families_fixed_phi(cuda_device, golden_dir, kern, impl):
    """Every kernel family of fs_mol/utils/gp_utils.py:22-43,106-153 at the fixture's fixed hyper-parameters: -MLL, its
    gradient, predictive mean / variance and the joint predictive NLL against the reference's own ExactGPLayer /
    ExactTanimotoGP outputs (tests/golden/gp_kernels.npz)."""
    from adkf_ift_b200 import ops
    g = np.load(os.path.join(golden_dir, "gp_kernels.npz"))
    X, y, n, m, layout, Xd, yd, phi = _family_inputs(g, kern, cuda_device)
    stat = ops.KERNEL_STATS[kern]
    noise_loc = float("nan") if kern == "tanimoto" else math.log(0.1) + 0.0625   # the Tanimoto fixture has no noise prior
    noise_init = 0.01 if kern == "tanimoto" else 0.1
    sq = ops.gp_sqdist(Xd, layout, qq_block=True, impl=impl, stat=stat)
    phid = torch.from_numpy(phi).view(1, 3).to(cuda_device)
    _, aux, st = ops.gp_fit(layout, sq.D2, yd, kern, noise_init, noise_loc, ls_prior=0, init_mode=0, max_iter=0, phi=phid)
    a = aux.cpu().numpy()[0]
    assert int(st.item()) & ops.ST_NOT_PD == 0
    nfree = len(g[f"{kern}_phi"])
    assert abs(a[ops.AUX_F] - g[f"{kern}_loss"]) < 1e-4 * abs(g[f"{kern}_loss"]), (a[ops.AUX_F], g[f"{kern}_loss"])
    assert rel(a[ops.AUX_G:ops.AUX_G + nfree], g[f"{kern}_grad"]) < 2e-3
    if nfree == 2:
        assert a[ops.AUX_G + 2] == 0.0          # no third parameter: frozen
    mean, var, _ = ops.gp_predict(layout, sq.D2, yd, kern, phid)
    assert rel(mean.cpu().numpy()[n:], g[f"{kern}_mean"]) < 1e-4
    assert rel(var.cpu().numpy()[n:], g[f"{kern}_var"]) < 1e-4
    # meta-test layout (no query x query block): the diagonal needed by the variance must still be there
    sq2 = ops.gp_sqdist(Xd, layout, qq_block=False, impl=impl, stat=stat)
    mean2, var2, _ = ops.gp_predict(layout, sq2.D2, yd, kern, phid)
    assert rel(var2.cpu().numpy()[n:], g[f"{kern}_var"]) < 1e-4
    if f"{kern}_nll" in g.files:
        _, aux2, _ = ops.gp_ift(layout, sq.D2, yd, kern, noise_loc, 0, phid, aux, mode=1)
        F = aux2.cpu().numpy()[0, ops.AUX_FOUTER]
        assert abs(F - g[f"{kern}_nll"]) < 2e-4 * abs(g[f"{kern}_nll"]), (F, g[f"{kern}_nll"])


@pytest.mark.parametrize("kern", ["linear", "poli1", "poli2", "cossim"])
def test_kernel_families_fit(cuda_device, golden_dir, kern):
    """Inner fit for the dot-product families: lands where fit_gpytorch_scipy lands (same loss to 1e-4, predictions to
    2e-3; these objectives are flat along outputscale x variance, so phi itself is not compared)."""
    from adkf_ift_b200 import ops
    g = np.load(os.path.join(golden_dir, "gp_kernels.npz"))
    X, y, n, m, layout, Xd, yd, phi = _family_inputs(g, kern, cuda_device)
    noise_loc = math.log(0.1) + 0.0625
    sq = ops.gp_sqdist(Xd, layout, qq_block=False, stat=ops.KERNEL_STATS[kern])
    phid = torch.from_numpy(phi).view(1, 3).to(cuda_device)
    phi_fit, aux, st = ops.gp_fit(layout, sq.D2, yd, kern, 0.1, noise_loc, ls_prior=0, init_mode=0, max_iter=100, phi=phid)
    a = aux.cpu().numpy()[0]
    print(kern, "phi", phi_fit.cpu().numpy()[0], "ref", g[f"{kern}_phi_fit"], "f", a[ops.AUX_F], "ref", g[f"{kern}_loss_fit"],
          "iters", a[ops.AUX_ITERS], "status", int(st.item()))
    assert int(st.item()) & ops.ST_NOT_PD == 0
    assert a[ops.AUX_F] <= g[f"{kern}_loss_fit"] + 1e-4 * abs(g[f"{kern}_loss_fit"])
    if a[ops.AUX_F] < g[f"{kern}_loss_fit"] - 1e-3 * abs(g[f"{kern}_loss_fit"]):
        # poli2: SciPy's L-BFGS-B run of the reference collapses the outputscale (raw -31.7, loss 1.862, predictions ~1e-9);
        # the Newton fit reaches a far lower objective (0.447) -- a different optimum, predictions are not comparable
        assert kern == "poli2"
        return
    mean, var, _ = ops.gp_predict(layout, sq.D2, yd, kern, phi_fit)
    assert rel(mean.cpu().numpy()[n:], g[f"{kern}_mean_fit"]) < 5e-3
    assert rel(var.cpu().numpy()[n:], g[f"{kern}_var_fit"]) < 5e-3


@pytest.mark.parametrize("kern", ["linear", "poli1", "poli2"])
def test_dot_family_hypergradient_vs_autograd_oracle(cuda_device, kern):
    """IFT hypergradient w.r.t. the features for a dot-product family at fixed phi against the torch-autograd oracle
    (fp64): exercises Ku / Kul of the new families and the DOT backward of the Gram kernel."""
    from adkf_ift_b200 import ops
    from oracle import adkf_oracle as orc
    rng = np.random.default_rng(3)
    n, m, d = 20, 28, 32
    X = (rng.normal(size=(n + m, d)) * 0.5).astype(np.float32)
    y = np.sign(rng.normal(size=n + m)).astype(np.float32)
    cfg = orc.OracleConfig(num_layers=1, gp_kernel=kern, use_lengthscale_prior=False)
    tail = orc.GPTail(cfg, torch.float64)
    phi = np.array([-2.0, 0.3, -0.4])
    with torch.no_grad():
        for p, v in zip(tail.params(), phi):
            p.fill_(float(v))
    Xt, yt = torch.from_numpy(X).double(), torch.from_numpy(y).double()
    ref = orc.hypergrad_wrt_features(tail, Xt[:n], yt[:n], Xt[n:], yt[n:])
    layout = ops.TaskLayout.build([n], [m], cuda_device)
    Xd, yd = torch.from_numpy(X).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    sq = ops.gp_sqdist(Xd, layout, qq_block=True, stat=ops.STAT_DOT)
    phid = torch.tensor(phi, dtype=torch.float32, device=cuda_device).view(1, 3)
    noise_loc = math.log(0.1) + 0.0625
    _, aux, _ = ops.gp_fit(layout, sq.D2, yd, kern, 0.1, noise_loc, ls_prior=0, init_mode=0, max_iter=0, phi=phid)
    Wm, aux2, st = ops.gp_ift(layout, sq.D2, yd, kern, noise_loc, 0, phid, aux)
    dZ = ops.gp_sqdist_bwd(Xd, layout, Wm, sq.mu, 1.0, stat=ops.STAT_DOT).cpu().numpy()
    b = aux2.cpu().numpy()[0]
    assert abs(b[ops.AUX_FOUTER] - float(ref["F"])) < 1e-4 * abs(float(ref["F"]))
    assert rel(b[ops.AUX_H:ops.AUX_H + 9].reshape(3, 3), ref["H"].numpy()) < 2e-3
    # linear: outputscale x variance is a degenerate direction (K = s v u), H is near-singular there -> v is ill-conditioned
    tol = 3e-2 if kern == "linear" else 5e-3
    assert rel(b[ops.AUX_V:ops.AUX_V + 3], ref["v"].numpy()) < tol
    want = np.concatenate([ref["dL_dzs"].numpy(), ref["dL_dzq"].numpy()])
    e = rel(dZ, want)
    print(kern, "dL/dZ rel err", e, "cond(H)", float(torch.linalg.cond(ref["H"])))
    assert e < tol


def test_undersized_workspace_is_flagged_not_overrun(cuda_device):
    """adkf_gp_fit / _predict / _ift lay out per-task workspace slices from device-side sizes; a caller-owned buffer
    that is too small (e.g. reused after the batch grew) must be refused ON THE DEVICE: every task gets ADKF_ST_BADWS and
    nothing is written (compute-sanitizer is not available on this pool, so the bound is checked explicitly)."""
    from adkf_ift_b200 import _cabi, ops
    lib = _cabi.load()
    rng = np.random.default_rng(0)
    n, m, d = 40, 30, 16
    Z = torch.from_numpy(rng.normal(size=(n + m, d)).astype(np.float32)).to(cuda_device)
    y = torch.from_numpy(rng.normal(size=n + m).astype(np.float32)).to(cuda_device)
    layout = ops.TaskLayout.build([n], [m], cuda_device)
    sq = ops.gp_sqdist(Z, layout, qq_block=True)
    need = layout.ws_floats(2)
    phi = torch.zeros(1, 3, device=cuda_device)
    aux = torch.zeros(1, ops.AUX, device=cuda_device)
    Wm = torch.full_like(sq.D2, 7.0)
    status = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    small = need // 2
    guard = torch.full((small + 4096,), -3.0, device=cuda_device)          # the tail past `small` must stay untouched
    rc = lib.adkf_gp_ift(1, ops.dptr(layout.z_off), ops.dptr(layout.n_sup), ops.dptr(layout.d2_off), ops.dptr(sq.D2), ops.dptr(y),
                         ops.kernel_arg("matern"), float(math.log(0.1) + 0.0625), 1, 0, ops.dptr(phi), ops.dptr(aux), ops.dptr(Wm),
                         ops.dptr(status), ops.dptr(guard), small, ops.stream_ptr())
    assert rc == 0
    assert int(status.item()) == ops.ST_BADWS
    assert bool((guard[small:] == -3.0).all()) and bool((Wm == 7.0).all())
    # with the right size the same call succeeds
    ws = torch.empty(need, device=cuda_device)
    rc = lib.adkf_gp_ift(1, ops.dptr(layout.z_off), ops.dptr(layout.n_sup), ops.dptr(layout.d2_off), ops.dptr(sq.D2), ops.dptr(y),
                         ops.kernel_arg("matern"), float(math.log(0.1) + 0.0625), 1, 0, ops.dptr(phi), ops.dptr(aux), ops.dptr(Wm),
                         ops.dptr(status), ops.dptr(ws), need, ops.stream_ptr())
    assert rc == 0 and int(status.item()) & ops.ST_BADWS == 0
