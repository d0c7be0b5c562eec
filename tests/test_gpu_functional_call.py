"""the reference's OWN meta-training call sequence on this package's ADKTModel.

The body of the task loop of fs_mol/utils/adaptive_dkt_utils.py:361-404 is run literally -- run_on_batches(train=True)
(reinit + inner fit), f_inner / f_outer closures over functional_call (fs_mol/utils/_stateless.py:79-119), the generic
cauchy_hypergradient (fs_mol/utils/cauchy_hypergradient.py:5-164) -- on adkf_ift_b200.ADKTModel, i.e. through the
differentiable single-task losses (double backward of the inner loss = H u and the mixed term of adkf_gp_ift mode 3).
It must reproduce (a) the fp64 gradient of the reference's own routine and (b) the fused batched `train_step` path."""
import os
import zlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


@pytest.mark.parametrize("name,kernel", [("s64_rbf_l2", "rbf"), ("c1_classification", "matern")])
def test_reference_train_step_body_on_this_model(cuda_device, golden_dir, name, kernel):
    from adkf_ift_b200 import dkt
    from adkf_ift_b200.adaptive_dkt_utils import run_on_batches, run_on_task_batch
    from adkf_ift_b200.hypergradient import cauchy_hypergradient, functional_call
    from test_gpu_model import _grad_metrics, build
    g, self, host, tb = build(golden_dir, name, kernel, cuda_device)
    g64 = np.load(os.path.join(golden_dir, "theta_grads_f64.npz"))
    device = cuda_device
    train_task_sample = dkt.torchify(host, device)
    batches = train_task_sample.batches
    batch_labels = train_task_sample.batch_labels
    batch_numeric_labels = train_task_sample.batch_numeric_labels

    # ---- adaptive_dkt_utils.py:361-404, literally -----------------------------------------------------------------
    _ = run_on_batches(self, batches=batches, batch_labels=batch_labels, batch_numeric_labels=batch_numeric_labels, train=True)
    assert len(batches) == 1
    fitted_phi = np.array([float(p.detach().reshape(-1)[0]) for p in self.gp_params()])
    assert np.abs(fitted_phi - g["train_phi_star"]).max() < 2e-3          # the on-device fit lands on SciPy's optimum
    # fixed-phi comparison: continue from the reference's own phi* (its L-BFGS-B stopping point)
    with torch.no_grad():
        for p, v in zip(self.gp_params(), g["train_phi_star"]):
            p.fill_(float(v))
    self.train()
    feature_extractor_params_names = [n for n, _ in self.named_parameters() if not n.startswith("gp_")]
    gp_params_names = [n for n, _ in self.named_parameters() if n.startswith("gp_")]

    def f_inner(params_outer, params_inner):
        feature_extractor_params_dict = {n: p for n, p in zip(feature_extractor_params_names, params_outer)}
        gp_params_dict = {n: p for n, p in zip(gp_params_names, params_inner)}
        self_params_dict = {**feature_extractor_params_dict, **gp_params_dict}
        return functional_call(self, self_params_dict, batches[0], kwargs={"train_loss": True, "is_functional_call": True})

    def f_outer(params_outer, params_inner):
        feature_extractor_params_dict = {n: p for n, p in zip(feature_extractor_params_names, params_outer)}
        gp_params_dict = {n: p for n, p in zip(gp_params_names, params_inner)}
        self_params_dict = {**feature_extractor_params_dict, **gp_params_dict}
        return functional_call(self, self_params_dict, batches[0],
                               kwargs={"train_loss": False, "predictive_val_loss": True, "is_functional_call": True})

    batch_loss = cauchy_hypergradient(f_outer, f_inner, tuple(self.feature_extractor_params()), tuple(self.gp_params()),
                                      self.device, ignore_grad_correction=self.config.ignore_grad_correction)
    # -------------------------------------------------------------------------------------------------------------------
    names = [str(x) for x in g["grad_names"]]
    params = dict(self.named_parameters())
    numels = np.array([params[n].numel() for n in names], dtype=np.float64)
    generic = {n: params[n].grad.detach().clone() for n in names}
    norms = np.array([generic[n].double().norm().item() for n in names])
    dots = np.array([(generic[n].double().cpu() * probe_vector(n, generic[n].shape)).sum().item() for n in names])
    n64, d64 = g64[f"{name}::grad_norms"], g64[f"{name}::grad_dots"]
    base_norm, base_dot = _grad_metrics(g["grad_norms"], g["grad_dots"], n64, d64, numels)
    e_norm, e_dot = _grad_metrics(norms, dots, n64, d64, numels)
    eF = abs(float(batch_loss) - float(g64[f"{name}::F"])) / abs(float(g64[f"{name}::F"]))
    print(f"reference call sequence on this model, {name}: F {eF:.2e} grad-norm {e_norm:.2e} (reference fp32 {base_norm:.2e}) "
          f"probe-dot {e_dot:.2e} (reference fp32 {base_dot:.2e})")
    assert eF < 1e-4
    assert e_norm < max(1e-4, 3 * base_norm) and e_dot < max(1e-4, 3 * base_dot)
    # GP parameters receive dF/dphi (cauchy_hypergradient.py:157-161)
    gp_grads = np.array([float(p.grad.reshape(-1)[0]) for p in self.gp_params()])
    want = g64[f"{name}::dF_dphi"]
    assert np.abs(gp_grads - want).max() < 1e-4 * np.abs(want).max() + 1e-6, (gp_grads, want)

    # ---- (b) the fused batched path at the same phi -----------------------------------------------------------------------
    self.zero_grad(set_to_none=True)
    phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
    res = run_on_task_batch(self, tb, train=True, grad_scale=1.0, phi=phi)
    top = max(float(generic[n].norm()) for n in names)
    worst = 0.0
    for n in names:
        got = params[n].grad if params[n].grad is not None else torch.zeros_like(generic[n])
        worst = max(worst, float((got - generic[n]).norm() / max(float(generic[n].norm()), 1e-3 * top)))
    print(f"    fused train_step path vs generic routine: worst per-tensor rel l2 {worst:.2e}, F {float(res.outer_loss.item())}")
    assert worst < 1e-4
    assert abs(float(res.outer_loss.item()) - float(batch_loss)) < 1e-5 * abs(float(batch_loss))
