"""-m gpu parity tests of BASELINE config 4 (MoleculeNet few-shot ADKF-IFT: GIN encoder + Matern GP) against
tests/golden/c4_moleculenet.npz = outputs of the reference's own encoder.py / adkf_model.py / cauchy_hypergradient.py on the
seeded synthetic graphs, plus kernel-level checks of gin.cu against plain torch.
Tolerances: graph indexing bit-exact; features 2e-4; hypergradient at the reference's phi* 2e-3 on per-tensor norms
(fp32 reference, 5 BatchNorm layers); after each side's own fit 5e-3."""
import os
import zlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def mol_batch(seed, n, device):
    from adkf_ift_b200 import synthetic
    from adkf_ift_b200.moleculenet import MolBatch
    d = synthetic.make_mol_graphs(seed, n)
    return MolBatch(torch.from_numpy(d["x"]), torch.from_numpy(d["edge_index"]), torch.from_numpy(d["edge_attr"]),
                    torch.from_numpy(d["batch"]), torch.from_numpy(d["y"]), n).to(device)


def fresh_model(device, dropout=0.0):
    from adkf_ift_b200 import synthetic
    from adkf_ift_b200.moleculenet import ADKFModel, default_args
    model = ADKFModel(default_args(dropout=dropout))
    synthetic.deterministic_init_(model, seed=1)
    return model.to(device)


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def test_csr_from_edge_index_bit_exact(cuda_device):
    from adkf_ift_b200.moleculenet import build_gin_index
    b = mol_batch(3, 17, cuda_device)
    gi = build_gin_index(b.x, b.edge_index, b.edge_attr)
    ei = b.edge_index.cpu().numpy()
    V, E = b.x.shape[0], ei.shape[1]
    for key_row, rowptr, col in ((1, gi.rowptr_t, gi.col_src), (0, gi.rowptr_s, gi.col_tgt)):
        order = np.lexsort((np.arange(E), ei[key_row]))                 # by key, then by edge id
        want_ptr = np.concatenate([[0], np.cumsum(np.bincount(ei[key_row], minlength=V))])
        assert np.array_equal(rowptr.cpu().numpy(), want_ptr.astype(np.int32))
        assert np.array_equal(col.cpu().numpy()[:E], ei[1 - key_row][order].astype(np.int32))
    assert np.array_equal(gi.col_eid_t.cpu().numpy()[:E], np.lexsort((np.arange(E), ei[1])).astype(np.int32))


def test_gin_kernels_vs_torch(cuda_device):
    """aggregate (fwd + all gradients), segmented BatchNorm (fwd/bwd, running estimates), mean pool -- against torch ops in fp64."""
    from adkf_ift_b200 import moleculenet as mn
    torch.manual_seed(0)
    b = mol_batch(5, 12, cuda_device)
    V, D = b.x.shape[0], 300
    gi = mn.build_gin_index(b.x, b.edge_index, b.edge_attr)
    h = torch.randn(V, D, device=cuda_device, requires_grad=True)
    E1 = torch.randn(6, D, device=cuda_device, requires_grad=True)
    E2 = torch.randn(3, D, device=cuda_device, requires_grad=True)
    go = torch.randn(V, D, device=cuda_device)
    out = mn._GinAggregate.apply(h, E1, E2, b.edge_attr, gi)
    out.backward(go)
    hd, E1d, E2d = (t.detach().double().cpu().requires_grad_(True) for t in (h, E1, E2))
    ei, ea = b.edge_index.cpu(), b.edge_attr.cpu()
    msg = hd[ei[0]] + E1d[ea[:, 0]] + E2d[ea[:, 1]]
    ref = torch.zeros(V, D, dtype=torch.float64).index_add_(0, ei[1], msg) + hd + E1d[4] + E2d[0]
    ref.backward(go.double().cpu())
    assert rel(out.detach().cpu(), ref.detach()) < 1e-6
    for got, want, name in ((h.grad, hd.grad, "dh"), (E1.grad, E1d.grad, "dE1"), (E2.grad, E2d.grad, "dE2")):
        assert rel(got.cpu(), want) < 5e-6, name

    # segmented BatchNorm: 3 segments == three separate torch BatchNorm1d calls
    seg = torch.tensor([0, V // 3, V // 2, V], dtype=torch.int32, device=cuda_device)
    x = torch.randn(V, D, device=cuda_device, requires_grad=True)
    gamma = (1 + 0.1 * torch.randn(D, device=cuda_device)).requires_grad_(True)
    beta = (0.1 * torch.randn(D, device=cuda_device)).requires_grad_(True)
    stats = []
    y = mn._SegBatchNorm.apply(x, gamma, beta, seg, 1, 1.0, 0, stats)
    y.backward(go)
    bn = torch.nn.BatchNorm1d(D).double()
    with torch.no_grad():
        bn.weight.copy_(gamma.detach().double().cpu()); bn.bias.copy_(beta.detach().double().cpu())
    xd = x.detach().double().cpu().requires_grad_(True)
    s = seg.cpu().tolist()
    yr = torch.cat([torch.relu(bn(xd[s[i]:s[i + 1]])) for i in range(3)])
    yr.backward(go.double().cpu())
    assert rel(y.detach().cpu(), yr.detach()) < 2e-6
    assert rel(x.grad.cpu(), xd.grad) < 2e-5
    assert rel(gamma.grad.cpu(), bn.weight.grad) < 2e-5 and rel(beta.grad.cpu(), bn.bias.grad) < 2e-5
    mine = torch.nn.BatchNorm1d(D).to(cuda_device)
    mn.GNN._update_running(mine, *stats[0])
    assert rel(mine.running_mean.cpu(), bn.running_mean) < 1e-5 and rel(mine.running_var.cpu(), bn.running_var) < 1e-5
    assert int(mine.num_batches_tracked) == 3
    # dropout: the backward applies the very mask of the forward (d/dx of sum(y) vanishes exactly where y was dropped)
    x2 = torch.randn(V, D, device=cuda_device, requires_grad=True)
    y2 = mn._SegBatchNorm.apply(x2, gamma.detach(), beta.detach(), seg, 0, 0.5, 1234, [])
    kept = (y2 != 0).float().mean().item()
    assert 0.45 < kept < 0.55
    y2b = mn._SegBatchNorm.apply(x2, gamma.detach(), beta.detach(), seg, 0, 0.5, 1234, [])
    assert torch.equal(y2, y2b)

    # mean pool
    gp = torch.zeros(13, dtype=torch.int32, device=cuda_device)
    gp[1:] = torch.cumsum(torch.bincount(b.batch, minlength=12), 0).to(torch.int32)
    x3 = torch.randn(V, D, device=cuda_device, requires_grad=True)
    p = mn._SegMean.apply(x3, gp, 12)
    gg = torch.randn(12, D, device=cuda_device)
    p.backward(gg)
    x3d = x3.detach().double().cpu().requires_grad_(True)
    cnt = torch.bincount(b.batch.cpu(), minlength=12).double()
    pr = torch.zeros(12, D, dtype=torch.float64).index_add_(0, b.batch.cpu(), x3d) / cnt[:, None]
    pr.backward(gg.double().cpu())
    assert rel(p.detach().cpu(), pr.detach()) < 1e-6 and rel(x3.grad.cpu(), x3d.grad) < 1e-6


def _merged_train_batch(g, device):
    from adkf_ift_b200.moleculenet import merge_mol_batches
    T, n_s, n_q = (int(v) for v in g["meta"])
    parts = []
    for t in range(T):
        parts += [mol_batch(100 + 2 * t, n_s, "cpu"), mol_batch(101 + 2 * t, n_q, "cpu")]
    return merge_mol_batches(parts, device, [n_s] * T, [n_q] * T), T, n_s, n_q


def test_c4_features_and_hypergradient_vs_reference(cuda_device, golden_dir):
    from adkf_ift_b200.moleculenet import run_on_mol_task_batch
    g = np.load(os.path.join(golden_dir, "c4_moleculenet.npz"))
    mb, T, n_s, n_q = _merged_train_batch(g, cuda_device)
    model = fresh_model(cuda_device)
    names = [n for n, _ in model.named_parameters() if not n.startswith("gp_")]
    assert names == [str(x) for x in g["grad_names"]]
    # (1) train-mode features of the merged pass == per-call features of the reference (segment BatchNorm statistics)
    model.train()
    with torch.no_grad():
        Z, _ = model.mol_encoder.encode_merged(mb)
    Z = Z.cpu().numpy()
    for t in range(T):
        zs, zq = Z[t * (n_s + n_q):t * (n_s + n_q) + n_s], Z[t * (n_s + n_q) + n_s:(t + 1) * (n_s + n_q)]
        assert rel(zs, g[f"t{t}_z_support"]) < 2e-4 and rel(zq, g[f"t{t}_z_query"]) < 2e-4
    # (2) hypergradient at the reference's phi*
    phi_ref = torch.tensor(np.stack([g[f"t{t}_phi_star"] for t in range(T)]), dtype=torch.float32, device=cuda_device)
    model = fresh_model(cuda_device)
    model.zero_grad(set_to_none=True)
    res = run_on_mol_task_batch(model, mb, train=True, phi=phi_ref)
    F = res.outer_loss.cpu().numpy()
    for t in range(T):
        assert abs(F[t] - g[f"t{t}_F"]) < 5e-4 * abs(g[f"t{t}_F"]), (t, F[t], g[f"t{t}_F"])
    grads = {n: p.grad for n, p in model.named_parameters() if not n.startswith("gp_")}
    norms = np.array([grads[n].double().norm().item() for n in names])
    dots = np.array([(grads[n].double().cpu() * probe_vector(n, grads[n].shape)).sum().item() for n in names])
    worst = np.abs(norms - g["grad_norms"]) / np.maximum(g["grad_norms"], 1e-3 * g["grad_norms"].max())
    print("fixed-phi: worst per-tensor norm error", worst.max(), names[int(worst.argmax())])
    assert worst.max() < 2e-3
    assert np.abs(dots - g["grad_dots"]).max() < 2e-3 * np.abs(g["grad_dots"]).max()
    for key in ("mol_encoder.gnn.batch_norms.4.bias", "mol_encoder.gnn.gnns.0.edge_embedding1.weight",
                "mol_encoder.gnn.x_embedding2.weight"):
        # batch_norms.4.bias only shifts all features of a segment: the Matern GP is translation invariant, so its gradient
        # is rounding noise (~1e-8) on both sides -> compared on an absolute floor
        a, b = grads[key].cpu().numpy(), g["grad::" + key]
        assert np.abs(a - b).max() < 2e-3 * max(np.abs(b).max(), 1e-3 * g["grad_norms"].max()), key
    # (3) own inner fit
    model = fresh_model(cuda_device)
    model.zero_grad(set_to_none=True)
    res = run_on_mol_task_batch(model, mb, train=True)
    phi = res.phi.cpu().numpy()
    assert int(res.status.max().item()) & 6 == 0
    for t in range(T):
        assert np.abs(phi[t] - g[f"t{t}_phi_star"]).max() < 3e-3, (t, phi[t], g[f"t{t}_phi_star"])
    grads = {n: p.grad for n, p in model.named_parameters() if not n.startswith("gp_")}
    norms = np.array([grads[n].double().norm().item() for n in names])
    worst = np.abs(norms - g["grad_norms"]) / np.maximum(g["grad_norms"], 1e-3 * g["grad_norms"].max())
    print("own fit: worst per-tensor norm error", worst.max())
    assert worst.max() < 5e-3


def test_c4_test_step_vs_reference(cuda_device, golden_dir):
    """test_step of the reference: GP fitted on TRAIN-mode support features (which also moves the BatchNorm running
    estimates once), predictions from EVAL-mode features (adkfift_trainer.py:279-286)."""
    from adkf_ift_b200.moleculenet import merge_mol_batches, run_on_mol_task_batch
    g = np.load(os.path.join(golden_dir, "c4_moleculenet.npz"))
    n_s = int(g["meta"][1])
    s, q = mol_batch(900, n_s, "cpu"), mol_batch(901, 48, "cpu")
    model = fresh_model(cuda_device)
    model.train()
    sup_only = merge_mol_batches([s], cuda_device, [n_s], [0])
    fit = run_on_mol_task_batch(model, sup_only, train=False)
    assert rel(model.mol_encoder.gnn.batch_norms[0].running_mean.cpu().numpy(), g["test_running_mean_l0"]) < 1e-4
    assert rel(model.mol_encoder.gnn.batch_norms[4].running_var.cpu().numpy(), g["test_running_var_l4"]) < 1e-4
    assert np.abs(fit.phi.cpu().numpy()[0] - g["test_phi_star"]).max() < 3e-3
    model.eval()
    full = merge_mol_batches([s, q], cuda_device, [n_s], [48])
    res = run_on_mol_task_batch(model, full, train=False, phi=fit.phi)
    assert rel(res.preds.cpu().numpy()[n_s:], g["test_preds"]) < 2e-3
    # the reference's own single-task surface gives the same numbers
    model2 = fresh_model(cuda_device)
    model2.train()
    sd = s.to(cuda_device)
    assert model2(s_data=sd, q_data=None, s_label=sd.y, train_loss=True) is None
    model2.mll.fit()
    model2.eval()
    preds, labels = model2.forward_query_loader(sd, [q], train_loss=None, s_label=sd.y)
    assert rel(preds.cpu().numpy(), g["test_preds"]) < 2e-3 and torch.equal(labels.cpu(), q.y)


def test_c4_trainer_step_moves_weights(cuda_device, golden_dir):
    from adkf_ift_b200.moleculenet import ADKF_Meta_Trainer, default_args
    g = np.load(os.path.join(golden_dir, "c4_moleculenet.npz"))
    mb, T, _, _ = _merged_train_batch(g, cuda_device)
    model = fresh_model(cuda_device, dropout=0.5)
    trainer = ADKF_Meta_Trainer(default_args(dropout=0.5), model)
    before = [p.detach().clone() for p in model.feature_extractor_params()]
    l0 = trainer.train_step(mb).loss_per_sample.mean().item()
    for _ in range(3):
        l1 = trainer.train_step(mb).loss_per_sample.mean().item()
    assert np.isfinite(l0) and np.isfinite(l1)
    assert any(not torch.equal(a, b) for a, b in zip(before, model.feature_extractor_params()))
