"""--use-ard: adkf_ift_b200/gp_ard.py against tests/golden/gp_ard.npz, which comes from the reference's OWN GP layer
`ExactGPLayer(..., ard_num_dims=d)` set up as ADKTModel does with use_ard=True (tests/golden/make_golden_ard.py; dense Hessian +
torch.linalg.solve as in cauchy_hypergradient.py:41-64,136).

  CPU : the host algebra (gradient assembly, matrix-free Hessian-vector product, CG and Neumann inverse-Hessian-vector products,
        hypergradient, batched L-BFGS) on a plain-torch fp64 kernel layer (oracle/gp_ard_backend.py): 1e-6.
  GPU : the same through the sm_100a kernels (adkf_ard_wsqdist / adkf_ard_pairquad / adkf_gp_ift modes 1, 2, 4): the north
        star's 1e-4 on v and dL/dZ at fixed phi, iteration counts and residuals printed.
"""
import math
import os

import numpy as np
import pytest
import torch

CASES = ["d64", "d64rbf", "d2048"]


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def load(golden_dir, tag):
    g = np.load(os.path.join(golden_dir, "gp_ard.npz"))
    n, m, d = (int(x) for x in g[f"{tag}_meta"])
    return g, n, m, d, str(g[f"{tag}_kernel"])


def check_fixed_phi(gp, g, tag, n, m, d, tol, dtype, device):
    from adkf_ift_b200.gp_ard import ARDFit
    phi = torch.from_numpy(g[f"{tag}_phi"]).to(dtype).to(device).view(1, -1)
    ls_loc = torch.tensor([float(g[f"{tag}_ls_loc"])], dtype=dtype, device=device)
    f, grad, cache = gp.value_and_grad(phi, ls_loc, keep=True)
    e_f = abs(float(f[0]) - float(g[f"{tag}_f"])) / abs(float(g[f"{tag}_f"]))
    e_g = rel(grad[0].cpu().numpy(), g[f"{tag}_g"])
    probes = torch.from_numpy(g[f"{tag}_probes"]).to(dtype).to(device)
    e_h = max(rel(gp.hvp(phi, ls_loc, probes[i:i + 1], cache)[0].cpu().numpy(), g[f"{tag}_Hu"][i]) for i in range(len(probes)))
    fit = ARDFit(phi, f, grad, ls_loc, torch.tensor([float(g[f"{tag}_l0"])], dtype=dtype, device=device), 0,
                 torch.zeros(1, dtype=torch.int32, device=device))
    dZ, info = gp.hypergradient(fit, method="cg", rtol=1e-7 if dtype == torch.float64 else 3e-6)
    e_F = abs(float(info["F"][0]) - float(g[f"{tag}_F"])) / abs(float(g[f"{tag}_F"]))
    e_dF = rel(info["dF_dphi"][0].cpu().numpy(), g[f"{tag}_dF_dphi"])
    e_v = rel(info["v"][0].cpu().numpy(), g[f"{tag}_v"])
    want = np.concatenate([g[f"{tag}_dL_dzs"], g[f"{tag}_dL_dzq"]])
    e_z = rel(dZ.cpu().numpy(), want)
    mean, var, _ = gp.predict(phi)
    e_m = rel(mean.cpu().numpy()[n:], g[f"{tag}_mean"])
    e_s = rel(var.cpu().numpy()[n:], g[f"{tag}_var"])
    print(f"ARD {tag} fixed phi: f {e_f:.1e} grad {e_g:.1e} Hu {e_h:.1e} F {e_F:.1e} dF/dphi {e_dF:.1e} | CG {info['iterations']} its, "
          f"residual {float(info['residual'][0]):.1e} | v {e_v:.1e} dL/dZ {e_z:.1e} | mean {e_m:.1e} var {e_s:.1e}")
    for name, e in (("f", e_f), ("grad", e_g), ("Hu", e_h), ("F", e_F), ("dF", e_dF), ("v", e_v), ("dZ", e_z), ("mean", e_m), ("var", e_s)):
        assert e < tol, (tag, name, e)
    return fit, cache


@pytest.mark.parametrize("tag", CASES)
def test_ard_host_algebra_vs_reference_gp_layer(golden_dir, tag):
    from adkf_ift_b200.gp_ard import BatchedARDExactGP
    from oracle.gp_ard_backend import TorchBackend
    g, n, m, d, kernel = load(golden_dir, tag)
    be = TorchBackend(torch.from_numpy(g[f"{tag}_X"]), torch.from_numpy(g[f"{tag}_y"]), [n], [n + m], kernel, math.log(0.1) + 0.0625)
    gp = BatchedARDExactGP(be.Z, be.y, None, kernel, backend=be)
    fit, cache = check_fixed_phi(gp, g, tag, n, m, d, 2e-6, torch.float64, "cpu")
    if tag == "d64":
        # the Neumann series of the north star reaches the same v (slowly: rate 1 - lambda_min / lambda_max)
        b = torch.from_numpy(g[f"{tag}_dF_dphi"]).view(1, -1)
        v, its, res = gp.ihvp(fit.phi, fit.ls_loc, b, method="neumann", max_iter=4000, rtol=1e-9, cache=cache)
        e = rel(v[0].numpy(), g[f"{tag}_v"])
        print(f"    Neumann series: {its} terms, residual {float(res[0]):.1e}, v err {e:.1e}; eig(H) {g[f'{tag}_eig_minmax']}")
        assert e < 1e-4
        # initialisation = the reference's reinit_gp_params with use_ard
        phi0, l0, loc = gp.init_phi()
        assert rel(phi0[0].numpy(), g[f"{tag}_phi_init"]) < 1e-9 and abs(float(loc[0]) - float(g[f"{tag}_ls_loc"])) < 1e-9
        # batched L-BFGS lands at (or below) the reference's L-BFGS-B optimum
        res = gp.fit(max_iter=300, gtol=1e-6)
        print(f"    L-BFGS: {res.iterations} iterations, f {float(res.f[0]):.8f} (reference {float(g[f'{tag}_f_fit']):.8f}), "
              f"|g|inf {float(res.grad.abs().max()):.1e}")
        assert float(res.f[0]) <= float(g[f"{tag}_f_fit"]) + 1e-6
        assert rel(res.phi[0].numpy(), g[f"{tag}_phi_fit"]) < 5e-3


@pytest.mark.gpu
@pytest.mark.parametrize("tag", CASES)
def test_ard_cuda_vs_reference_gp_layer(cuda_device, golden_dir, tag):
    from adkf_ift_b200 import ops
    from adkf_ift_b200.gp_ard import BatchedARDExactGP
    g, n, m, d, kernel = load(golden_dir, tag)
    Z = torch.from_numpy(g[f"{tag}_X"]).to(cuda_device)
    y = torch.from_numpy(g[f"{tag}_y"]).to(cuda_device)
    layout = ops.TaskLayout.build([n], [m], cuda_device)
    gp = BatchedARDExactGP(Z, y, layout, kernel)
    fit, cache = check_fixed_phi(gp, g, tag, n, m, d, 1e-4, torch.float32, cuda_device)
    phi0, l0, loc = gp.init_phi()
    assert rel(phi0[0].cpu().numpy(), g[f"{tag}_phi_init"]) < 1e-5
    res = gp.fit(max_iter=300, gtol=2e-5)
    print(f"    L-BFGS on device: {res.iterations} iterations, f {float(res.f[0]):.6f} (reference L-BFGS-B {float(g[f'{tag}_f_fit']):.6f}), "
          f"|g|inf {float(res.grad.abs().max()):.1e}, status {int(res.status[0])}")
    assert float(res.f[0]) <= float(g[f"{tag}_f_fit"]) + 2e-5 * abs(float(g[f"{tag}_f_fit"]))
    if tag == "d64":
        b = torch.from_numpy(g[f"{tag}_dF_dphi"]).float().to(cuda_device).view(1, -1)
        v, its, r = gp.ihvp(fit.phi, fit.ls_loc, b, method="neumann", max_iter=4000, rtol=1e-7, cache=cache)
        e = rel(v[0].cpu().numpy(), g[f"{tag}_v"])
        print(f"    Neumann series on device: {its} terms, residual {float(r[0]):.1e}, v err {e:.1e}")
        assert e < 1e-3


@pytest.mark.gpu
def test_ard_batched_tasks_equal_single(cuda_device, golden_dir):
    """Two ragged tasks in one launch == each alone (gradient, Hessian-vector product)."""
    from adkf_ift_b200 import ops
    from adkf_ift_b200.gp_ard import BatchedARDExactGP
    g = np.load(os.path.join(golden_dir, "gp_ard.npz"))
    parts = []
    for tag in ("d64", "d64rbf"):
        n, m, d = (int(x) for x in g[f"{tag}_meta"])
        parts.append((g[f"{tag}_X"], g[f"{tag}_y"], n, m, g[f"{tag}_phi"].astype(np.float32), float(g[f"{tag}_ls_loc"])))
    parts[1] = (parts[1][0][:50], parts[1][1][:50], 24, 26, parts[1][4], parts[1][5])      # ragged: 24 + 26 rows
    Z = torch.from_numpy(np.concatenate([p[0] for p in parts])).to(cuda_device)
    y = torch.from_numpy(np.concatenate([p[1] for p in parts])).to(cuda_device)
    layout = ops.TaskLayout.build([p[2] for p in parts], [p[3] for p in parts], cuda_device)
    gp = BatchedARDExactGP(Z, y, layout, "matern")
    phi = torch.from_numpy(np.stack([p[4] for p in parts])).to(cuda_device)
    loc = torch.tensor([p[5] for p in parts], device=cuda_device)
    f, grad, cache = gp.value_and_grad(phi, loc, keep=True)
    x = torch.randn(2, 66, generator=torch.Generator().manual_seed(1)).to(cuda_device)
    hx = gp.hvp(phi, loc, x, cache)
    off = 0
    for t, p in enumerate(parts):
        rows = p[2] + p[3]
        lt = ops.TaskLayout.build([p[2]], [p[3]], cuda_device)
        single = BatchedARDExactGP(Z[off:off + rows].contiguous(), y[off:off + rows].contiguous(), lt, "matern")
        f1, g1, c1 = single.value_and_grad(phi[t:t + 1], loc[t:t + 1], keep=True)
        h1 = single.hvp(phi[t:t + 1], loc[t:t + 1], x[t:t + 1], c1)
        assert rel(grad[t].cpu().numpy(), g1[0].cpu().numpy()) < 1e-5 and rel(hx[t].cpu().numpy(), h1[0].cpu().numpy()) < 1e-5
        off += rows


@pytest.mark.gpu
def test_ard_model_surface(cuda_device):
    """ADKTModel with use_ard=True through the batched task path: (i) with all d lengthscales equal the outer loss and its
    feature gradient are those of the scalar-lengthscale kernels at the same value; (ii) a meta-training step and a meta-test
    pass run end to end (fit converges, finite losses, gradients reach the extractor)."""
    from adkf_ift_b200 import dkt, gp as gpk, ops, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer, run_on_task_batch
    from adkf_ift_b200.feature_extractor import cli_default_config
    from adkf_ift_b200.gp_ard import ARDFit, BatchedARDExactGP
    model = ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(2), use_ard=True,
                                                    tasks_per_batch=2, clip_value=1.0, learning_rate=1e-4))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(cuda_device)
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(800 + t, 16, 24), dkt.get_dkt_batcher(max_num_graphs=96),
                                                False).batches[0] for t in range(2)]
    tb = task_batch.to_device(task_batch.collate_tasks(hosts, False), cuda_device)
    with torch.no_grad():
        Z = model.embed(tb)
    # (i) isotropic special case against the scalar path
    g_iso = gpk.BatchedExactGP(Z, tb.labels, tb.layout, "matern", False, True, qq_block=True)
    fit_iso = g_iso.fit()
    dZ_iso, aux_iso, _ = g_iso.hypergradient(fit_iso, mode=1)                 # direct term only
    g_ard = BatchedARDExactGP(Z, tb.labels, tb.layout, "matern")
    phi = torch.cat([fit_iso.phi[:, :2], fit_iso.phi[:, 2:3].expand(2, Z.shape[1])], 1).contiguous()
    _, l0, loc = g_ard.init_phi()
    f, grad, st = g_ard.value_and_grad(phi, loc)
    dZ_ard, info = g_ard.hypergradient(ARDFit(phi, f, grad, loc, l0, 0, st), ignore_grad_correction=True)
    assert rel(info["F"].cpu().numpy(), aux_iso[:, ops.AUX_FOUTER].cpu().numpy()) < 1e-4
    e = rel(dZ_ard.cpu().numpy(), dZ_iso.cpu().numpy())
    print("ARD with equal lengthscales vs scalar path: dF/dZ rel err", e)
    assert e < 1e-4
    # the sum over dimensions of the ARD lengthscale gradient is the scalar lengthscale gradient (same prior location)
    # (ii) end to end
    res = run_on_task_batch(model, tb, train=False)
    assert torch.isfinite(res.preds).all() and int(res.status.max()) & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED) == 0
    assert tuple(res.phi.shape) == (2, 2 + Z.shape[1])
    before = torch.cat([p.detach().reshape(-1).clone() for p in model.feature_extractor_params()])
    out = model.train_step(tb)
    torch.cuda.synchronize()
    after = torch.cat([p.detach().reshape(-1) for p in model.feature_extractor_params()])
    print("ARD train step: F", out.outer_loss.cpu().numpy(), "status", out.status.cpu().numpy())
    assert torch.isfinite(out.outer_loss).all() and torch.isfinite(after).all() and float((after - before).abs().max()) > 0
