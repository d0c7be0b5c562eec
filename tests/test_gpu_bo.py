"""-m gpu parity tests of BASELINE config 5 (bayes_opt dockstring-shaped pool, blocked Cholesky path) against
tests/golden/c5_bayes_opt.npz, which holds the outputs of the reference's own create_gp / CustomKernelGP
(bayes_opt/bo_utils.py:400-461) + fit_gpytorch_scipy + ExpectedImprovement on the same seeded pool.
Tolerances: 1e-4 relative at fixed phi (north_star), 2e-3 after each side's own fit (SciPy's stopping error)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def _setup(g, kernel_type, device):
    from adkf_ift_b200 import synthetic
    d = 64 if kernel_type == "matern" else 256
    x_all, y_all = synthetic.bo_pool(seed=5, pool=2312, dim=d, binary=(kernel_type == "tanimoto"))
    tr = g[f"{kernel_type}_train_idx"]
    te = np.array([i for i in range(len(y_all)) if i not in set(tr.tolist())])
    X, Y = torch.from_numpy(x_all).to(device), torch.from_numpy(y_all).to(device)
    y_tr = Y[tr]
    mu, sd = y_tr.mean(), y_tr.std()
    return X, (Y - mu) / sd, torch.from_numpy(tr).to(device), torch.from_numpy(te).to(device)


@pytest.mark.parametrize("kernel_type", ["matern", "tanimoto"])
def test_create_gp_fit_predict_ei_nll(cuda_device, golden_dir, kernel_type):
    from adkf_ift_b200 import bo_utils, ops
    from adkf_ift_b200.gp_utils import fit_gpytorch_scipy
    g = np.load(os.path.join(golden_dir, "c5_bayes_opt.npz"))
    k32, k64 = f"{kernel_type}_f32_", f"{kernel_type}_f64_"
    X, Ys, tr, te = _setup(g, kernel_type, cuda_device)
    likelihood, model, mll = bo_utils.create_gp(X[tr], Ys[tr], kernel_type, cuda_device, 0.01, True)
    # initial parameters: noise 0.01, outputscale softplus(0), median lengthscale (+ its prior loc)
    phi0 = model._phi().cpu().numpy()[0]
    nfree = len(g[k64 + "phi0"])
    assert np.abs(phi0[:nfree] - g[k64 + "phi0"]).max() < 1e-4
    if kernel_type == "matern":
        assert abs(model.covar_module.base_kernel.lengthscale_prior.loc - float(g[k64 + "ls_prior_loc"])) < 1e-5

    # ---- fixed phi (the reference's fp64 optimum): posterior over the pool, EI, joint NLL of the 2248 test molecules ----
    phi_ref = np.zeros(3, np.float32)
    phi_ref[:nfree] = g[k64 + "phi"]
    model._write_phi(torch.from_numpy(phi_ref).to(cuda_device))
    model.eval(); likelihood.eval()
    pool = likelihood(model(X))
    assert rel(pool.mean.cpu().numpy(), g[k64 + "pool_mean"]) < 1e-4
    assert rel(pool.variance.cpu().numpy(), g[k64 + "pool_var"]) < 1e-4
    ei = ops.ei_pool(pool.mean, pool.variance, float(likelihood.noise.item()), float(g[k64 + "best_f"])).cpu().numpy()
    ref_ei = g[k64 + "ei"]
    # EI spans many decades; compare where it is not denormal-small, relative to the largest value
    assert np.abs(ei - ref_ei).max() < 2e-4 * ref_ei.max()
    assert int(np.argmax(ei)) == int(np.argmax(ref_ei))
    pred = likelihood(model(X[te]))
    nll = float(-pred.log_prob(Ys[te]).item() / len(te))
    ref_nll = float(g[k64 + "nll_per_point"])
    err32 = abs(float(g[k32 + "nll_per_point"]) - ref_nll)
    print(kernel_type, "nll/m", nll, "ref f64", ref_nll, "ref f32 err", err32)
    assert abs(nll - ref_nll) < max(1e-4 * abs(ref_nll), 3 * err32)

    # ---- own fit from the reference's initialisation ----
    likelihood, model, mll = bo_utils.create_gp(X[tr], Ys[tr], kernel_type, cuda_device, 0.01, True)
    model.train(); likelihood.train()
    _, info = fit_gpytorch_scipy(mll)
    phi = model._phi().cpu().numpy()[0]
    print(kernel_type, "phi*", phi, "ref", g[k64 + "phi"], "iters", float(info["iterations"]))
    assert np.abs(phi[:nfree] - g[k64 + "phi"]).max() < 3e-3
    model.eval(); likelihood.eval()
    pred = likelihood(model(X[te]))
    assert rel(pred.mean.cpu().numpy(), g[k64 + "test_mean"]) < 2e-3
    assert rel(pred.variance.cpu().numpy(), g[k64 + "test_var"]) < 2e-3


@pytest.mark.parametrize("N,n", [(700, 64), (2376, 64), (1024, 128), (130, 16)])
def test_blocked_cholesky_joint_nll_vs_numpy(cuda_device, N, n):
    """adkf_gp_joint_nll_blocked against a numpy fp64 Cholesky of the same covariance (support NLL, joint NLL, log det),
    and against the one-CTA gp_ift path where both apply."""
    from adkf_ift_b200 import ops
    from oracle import gp_closed_form as cf
    rng = np.random.default_rng(N)
    d = 32
    Z = (np.tanh(rng.normal(size=(N, 3)) @ rng.normal(size=(3, d))) + 0.2 * rng.normal(size=(N, d))).astype(np.float32)
    y = rng.normal(size=N).astype(np.float32)
    phi = np.array([-2.5, 0.4, 1.8], np.float32)
    layout = ops.TaskLayout.build([n], [N - n], cuda_device)
    Zd, yd = torch.from_numpy(Z).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=True)
    phid = torch.from_numpy(phi).to(cuda_device)
    out, status = ops.gp_joint_nll_blocked(sq.D2, N, n, yd, "matern", phid)
    out = out.cpu().numpy()
    assert int(status.item()) & ops.ST_NOT_PD == 0
    D2 = cf.sq_dist_centered(Z.astype(np.float64), n)
    sig2, s, l, _, _ = cf.transforms(phi.astype(np.float64))
    K = cf.kernel_parts(D2, s, l, "matern")[0] + sig2 * np.eye(N)

    def nll(A, yy):
        L = np.linalg.cholesky(A)
        w = np.linalg.solve(L, yy)
        return 0.5 * w @ w + np.log(np.diag(L)).sum() + 0.5 * len(yy) * np.log(2 * np.pi), 2 * np.log(np.diag(L)).sum()

    nll_all, logdet = nll(K, y.astype(np.float64))
    nll_sup, _ = nll(K[:n, :n], y[:n].astype(np.float64))
    print(N, n, "blocked", out, "ref", nll_all, nll_sup, nll_all - nll_sup, logdet)
    assert abs(out[0] - nll_all) < 1e-4 * abs(nll_all)
    assert abs(out[1] - nll_sup) < 1e-4 * abs(nll_sup)
    assert abs(out[2] - (nll_all - nll_sup)) < 1e-4 * abs(nll_all - nll_sup) + 1e-4 * abs(nll_all)
    assert abs(out[3] - logdet) < 1e-4 * abs(logdet)
    if N <= 1024:
        aux0 = torch.zeros(1, ops.AUX, dtype=torch.float32, device=cuda_device)
        _, aux, _ = ops.gp_ift(layout, sq.D2, yd, "matern", 0.0, 0, phid.view(1, 3), aux0, mode=1)
        F = float(aux[0, ops.AUX_FOUTER].item())
        assert abs(F - out[2]) < 2e-4 * abs(out[2]) + 1e-4 * abs(nll_all)


def test_run_gp_ei_bo_loop_improves(cuda_device):
    """The BO loop end to end on a 600-molecule pool: records are valid pool indices, nothing is queried twice, and the
    best rank found improves on the initial design (the pool is sorted, index == rank)."""
    from adkf_ift_b200 import bo_utils, synthetic
    x_all, y_all = synthetic.bo_pool(seed=11, pool=600, dim=64)
    X, Y = torch.from_numpy(x_all).to(cuda_device), torch.from_numpy(y_all).to(cuda_device)
    np.random.seed(0)
    rec = bo_utils.run_gp_ei_bo(list(range(600)), X, Y, num_init_points=16, query_batch_size=4, num_bo_iters=6,
                                kernel_type="matern", device=cuda_device, init_from=300, noise_init=0.01, noise_prior=True)
    assert len(rec) == 1 + 6 * 4 and all(0 <= i < 600 for i in rec)
    assert len(set(rec[1:])) == len(rec[1:])
    best = bo_utils.min_so_far(rec)
    print("BO record", rec, "best-so-far", best[-1])
    assert best[-1] < rec[0] and best[-1] < 300


@pytest.mark.parametrize("outputscale_raw,duplicates", [(0.4, False), (30.0, True), (3000.0, True)])
def test_joint_nll_device_jitter_ladder_equals_host_ladder(cuda_device, outputscale_raw, duplicates):
    """adkf_gp_joint_nll_blocked_ladder (psd_safe_cholesky's 0 / 1e-6 / 1e-5 / 1e-4 ladder run on the device through gated
    launches, no host synchronisation) == the same ladder driven from the host with the single-attempt entry point: same
    status word, bit-identical outputs -- for a well-conditioned covariance, one that needs jitter and one that stays non-PD."""
    import ctypes
    from adkf_ift_b200 import _cabi, ops
    lib = _cabi.load()
    rng = np.random.default_rng(3)
    N, n, d = 600, 48, 24
    Z = rng.normal(size=(N, d)).astype(np.float32)
    if duplicates:
        Z[200:400] = Z[:200]                                  # exact duplicates: K is singular up to the noise
    y = rng.normal(size=N).astype(np.float32)
    phi = np.array([-30.0, outputscale_raw, 1.5], np.float32)   # noise at its 1e-4 floor
    layout = ops.TaskLayout.build([n], [N - n], cuda_device)
    Zd, yd = torch.from_numpy(Z).to(cuda_device), torch.from_numpy(y).to(cuda_device)
    sq = ops.gp_sqdist(Zd, layout, qq_block=True)
    phid = torch.from_numpy(phi).to(cuda_device)
    out, status = ops.gp_joint_nll_blocked(sq.D2, N, n, yd, "matern", phid)
    # host-driven reference ladder
    nws = int(lib.adkf_chol_blocked_ws_floats(N))
    ws = torch.empty(nws, dtype=torch.float32, device=cuda_device)
    want = torch.zeros(4, dtype=torch.float32, device=cuda_device)
    st = torch.zeros(1, dtype=torch.int32, device=cuda_device)
    want_status = ops.ST_NOT_PD
    for jitter in (0.0, 1e-6, 1e-5, 1e-4):
        rc = lib.adkf_gp_joint_nll_blocked(ops.dptr(sq.D2), N, n, ops.dptr(yd), ops.kernel_arg("matern") & 0xff, ops.dptr(phid),
                                           ctypes.c_float(jitter), ops.dptr(ws), nws, ops.dptr(want), ops.dptr(st), ops.stream_ptr())
        assert rc == 0
        if int(st.item()) & ops.ST_NOT_PD == 0:
            want_status = ops.ST_JITTER if jitter > 0 else 0
            break
    print(f"outputscale raw {outputscale_raw}, duplicates {duplicates}: ladder status {int(status.item())}, host ladder {want_status}")
    assert int(status.item()) == want_status
    if want_status != ops.ST_NOT_PD:
        assert torch.equal(out, want)
