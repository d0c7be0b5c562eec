"""-m gpu tests of the trainer / evaluation surface around the hot path (fs_mol/utils/adaptive_dkt_utils.py:119-211, 261-303,
329-458) on a small synthetic dataset: batched evaluation == the reference's per-sample protocol, train_loop runs, validates
and checkpoints, GNN-only checkpoints load with the reference's two learning-rate groups."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def small_model(device, **over):
    from adkf_ift_b200 import synthetic
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    kw = dict(graph_feature_extractor_config=cli_default_config(2), batch_size=96, tasks_per_batch=4, support_set_size=16,
              query_set_size=32, num_train_steps=3, validate_every_num_steps=2, validation_support_set_sizes=(16,),
              validation_query_set_size=32, validation_num_samples=1, learning_rate=1e-4, clip_value=1.0)
    kw.update(over)
    model = ADKTModelTrainer(ADKTModelTrainerConfig(**kw))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    return model.to(device)


@pytest.mark.parametrize("numeric", [False, True])
def test_batched_evaluation_matches_per_sample_protocol(cuda_device, tmp_path, numeric):
    from adkf_ift_b200 import dkt, eval_utils as E, synthetic
    from adkf_ift_b200.adaptive_dkt_utils import evaluate_adkt_model, run_on_batches
    model = small_model(cuda_device, use_numeric_labels=numeric)
    ds = synthetic.SyntheticFSMolDataset(2, 1, 3, molecules_per_task=72, numeric_scale=50.0 if numeric else 1.0)
    batched = evaluate_adkt_model(model, ds, support_sizes=[16, 24], num_samples=2, seed=1, batch_size=96, query_size=None,
                                  save_dir=str(tmp_path), tasks_per_launch=5)
    batcher = dkt.get_dkt_batcher(max_num_graphs=96)

    def per_sample(sample, tmp, seed):
        dev = dkt.torchify(dkt.task_sample_to_dkt_task_sample(sample, batcher, numeric), cuda_device)
        return run_on_batches(model, dev.batches, dev.batch_labels, dev.batch_numeric_labels, train=False)
    single = E.eval_model(per_sample, ds, [16, 24], num_samples=2, seed=1, filter_numeric_labels=numeric)
    assert sorted(batched) == sorted(single) and len(batched) == 3
    keys = ("mse", "mae", "r2") if numeric else ("acc", "roc_auc", "avg_precision", "f1")
    for name in batched:
        assert len(batched[name]) == len(single[name]) == 4
        for a, b in zip(batched[name], single[name]):
            assert (a.seed, a.num_train, a.num_test) == (b.seed, b.num_train, b.num_test)
            for k in keys:
                assert abs(getattr(a, k) - getattr(b, k)) <= 2e-3 * max(1.0, abs(getattr(b, k))), (name, k, getattr(a, k), getattr(b, k))
        assert os.path.exists(tmp_path / f"{name}_eval_results.csv")
    # a query set larger than one batch goes through run_on_batches (several query batches per task)
    multi = evaluate_adkt_model(model, ds, support_sizes=[16], num_samples=1, seed=1, batch_size=40, query_size=None)
    assert all(len(v) == 1 and v[0].num_test == 56 for v in multi.values())


def test_train_loop_and_checkpoints(cuda_device, tmp_path):
    from adkf_ift_b200 import synthetic
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    model = small_model(cuda_device)
    ds = synthetic.SyntheticFSMolDataset(5, 2, 1, molecules_per_task=64)
    before = torch.cat([p.detach().reshape(-1).clone() for p in model.feature_extractor_params()])
    losses = model.train_loop(str(tmp_path), ds, cuda_device)
    after = torch.cat([p.detach().reshape(-1) for p in model.feature_extractor_params()])
    assert len(losses) == 3 and np.isfinite(losses).all()
    assert float((after - before).abs().max()) > 0.0
    for f in ("best_validation.pt", "fully_trained.pt"):
        assert os.path.exists(tmp_path / f)
    again = ADKTModelTrainer.build_from_model_file(str(tmp_path / "fully_trained.pt"), device=cuda_device)
    got = torch.cat([p.detach().reshape(-1) for p in again.feature_extractor_params()])
    assert torch.equal(got, after)


def test_load_model_gnn_weights(cuda_device, tmp_path):
    model = small_model(cuda_device)
    donor = small_model(cuda_device)
    with torch.no_grad():
        for p in donor.graph_feature_extractor.parameters():
            p.add_(0.25)
    # a GNNMultitask checkpoint names the extractor's tensors without the prefix
    torch.save({"model_state_dict": {k: v.clone() for k, v in donor.graph_feature_extractor.state_dict().items()}}, tmp_path / "gnn.pt")
    fc_before = model.fc[0].weight.detach().clone()
    model.load_model_gnn_weights(str(tmp_path / "gnn.pt"), device=cuda_device)
    for (n, a), (_, b) in zip(model.graph_feature_extractor.state_dict().items(), donor.graph_feature_extractor.state_dict().items()):
        assert torch.equal(a, b), n
    assert torch.equal(model.fc[0].weight, fc_before)
    lrs = [g["lr"] for g in model.optimizer.param_groups]
    assert len(lrs) == 2 and lrs[1] == 0.0 and abs(lrs[0] - 1e-4) < 1e-12          # warm-up: the loaded group starts at 0
    assert sum(p.numel() for p in model.optimizer.param_groups[1]["params"]) == sum(p.numel() for p in model.graph_feature_extractor.parameters())


def test_device_prefetcher_matches_direct_staging(cuda_device):
    """task_batch.DevicePrefetcher (side-stream H2D + CSR build, double buffered) yields the same device batches, in order,
    as task_batch.to_device, and a model step on them gives bit-identical results."""
    from adkf_ift_b200 import dkt, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer, run_on_task_batch
    from adkf_ift_b200.feature_extractor import cli_default_config
    model = ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(2)))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(cuda_device)
    hbs = []
    for j in range(3):
        hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(300 + 10 * j + t, 16, 20 + j),
                                                    dkt.get_dkt_batcher(max_num_graphs=320), False).batches[0] for t in range(2)]
        hbs.append(task_batch.collate_tasks(hosts, False))
    got = list(task_batch.DevicePrefetcher(iter(hbs), cuda_device))
    assert len(got) == 3
    for hb, tb in zip(hbs, got):
        ref = task_batch.to_device(hb, cuda_device)
        assert torch.equal(tb.node_features, ref.node_features) and torch.equal(tb.fingerprints, ref.fingerprints)
        assert torch.equal(tb.graph_index.rowptr, ref.graph_index.rowptr)
        assert torch.equal(tb.graph_index.col_src, ref.graph_index.col_src)
        a = run_on_task_batch(model, tb, train=False)
        b = run_on_task_batch(model, ref, train=False)
        assert torch.equal(a.mean, b.mean) and torch.equal(a.phi, b.phi)


def test_train_loop_pipeline_identical_losses(cuda_device, tmp_path):
    """train_loop fed by worker processes (featurisation + collation in the background, pinned staging, async
    H2D, loss read one step late) gives EXACTLY the losses and parameters of the inline path over the same shard generators."""
    from adkf_ift_b200 import synthetic
    ds = synthetic.SyntheticFSMolDataset(12, 1, 1, molecules_per_task=64)
    runs = []
    for workers in (0, 2):
        model = small_model(cuda_device, num_train_steps=5, validate_every_num_steps=100)
        losses = model.train_loop(str(tmp_path / f"w{workers}"), ds, cuda_device, num_workers=workers, emulate_workers=2)
        runs.append((losses, torch.cat([p.detach().reshape(-1).clone() for p in model.feature_extractor_params()])))
    assert len(runs[0][0]) == 5 and np.isfinite(runs[0][0]).all()
    assert runs[0][0] == runs[1][0], (runs[0][0], runs[1][0])
    assert torch.equal(runs[0][1], runs[1][1])


def test_failed_task_is_dropped_and_mean_renormalised(cuda_device, monkeypatch):
    """A task whose covariance is not positive definite even after the jitter ladder (forced here through its GP parameters:
    outputscale 1e8, lengthscale 1e9 -> a rank-one K + 1e-4 I in fp32) is flagged, contributes a zero gradient and a NaN
    outer loss, and the meta-batch mean is taken over the remaining tasks: the step equals the step on the batch WITHOUT
    that task (the reference raises in this situation: gpytorch NotPSDError)."""
    from adkf_ift_b200 import adaptive_dkt_utils as U
    from adkf_ift_b200 import dkt, ops, synthetic, task_batch
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(500 + t, 16, 24), dkt.get_dkt_batcher(max_num_graphs=96),
                                                False).batches[0] for t in range(3)]
    real_fit = U._fit_or_fixed

    def sabotaged(g, phi, max_iter, gtol):
        fit = real_fit(g, phi, max_iter, gtol)
        if fit.phi.shape[0] == 3:
            fit.phi[2] = torch.tensor([-40.0, 1e8, 1e9], device=fit.phi.device)   # softplus(x) = x: outputscale 1e8, lengthscale 1e9
        return fit

    monkeypatch.setattr(U, "_fit_or_fixed", sabotaged)
    results = []
    for drop in (False, True):
        model = small_model(cuda_device)
        use = hosts if drop else hosts[:2]
        tb = task_batch.to_device(task_batch.collate_tasks(use, False), cuda_device)
        res = model.train_step(tb, global_num_tasks=len(use))
        torch.cuda.synchronize()
        # the (renormalised, clipped) meta-gradient the optimiser saw -- parameters after ONE Adam step are lr * sign-like and
        # would amplify rounding differences of near-zero gradient elements
        results.append((res, model._flat.grad.detach().clone()))
    res_ok, p_ok = results[0]
    res_drop, p_drop = results[1]
    st = res_drop.status.cpu().numpy()
    print("status of the three tasks", st)
    assert st[2] & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED) and not (st[:2] & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED)).any()
    assert torch.isnan(res_drop.outer_loss[2]) and torch.isfinite(res_drop.outer_loss[:2]).all()
    # same meta-gradient as the 2-task batch (the extractor forward of the extra graphs does not touch the others)
    assert torch.isfinite(p_drop).all()
    rel = float((p_drop - p_ok).norm() / p_ok.norm())
    print("dropped-task step vs 2-task step: relative l2 difference of the meta-gradient", rel)
    assert rel < 1e-5


def test_optimizer_state_round_trip(cuda_device, tmp_path):
    """Adam moments survive save_model / load_model_weights although the live optimiser steps ONE flat tensor: the checkpoint
    holds per-parameter state (the reference's layout), a fresh trainer continues bit-identically."""
    from adkf_ift_b200 import dkt, synthetic, task_batch
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(600 + t, 16, 24), dkt.get_dkt_batcher(max_num_graphs=96),
                                                False).batches[0] for t in range(2)]
    hb = task_batch.collate_tasks(hosts, False)
    a = small_model(cuda_device)
    for _ in range(2):
        a.train_step(task_batch.to_device(hb, cuda_device))
    a.save_model(str(tmp_path / "ckpt.pt"), optimizer=a.optimizer, epoch=2)
    sd = torch.load(str(tmp_path / "ckpt.pt"), weights_only=False)["optimizer_state_dict"]
    n_params = len(a.feature_extractor_params())
    assert len(sd["state"]) == n_params and sd["param_groups"][0]["params"] == list(range(n_params))
    assert tuple(sd["state"][0]["exp_avg"].shape) == tuple(a.feature_extractor_params()[0].shape)
    b = small_model(cuda_device)
    b._ensure_flat()
    b.load_model_weights(str(tmp_path / "ckpt.pt"), device=cuda_device)
    a.train_step(task_batch.to_device(hb, cuda_device))
    b.train_step(task_batch.to_device(hb, cuda_device))
    pa = torch.cat([p.detach().reshape(-1) for p in a.feature_extractor_params()])
    pb = torch.cat([p.detach().reshape(-1) for p in b.feature_extractor_params()])
    assert torch.equal(pa, pb)
    # and into the reference's own optimiser class (per-parameter Adam)
    ref_opt = torch.optim.Adam(small_model(cuda_device).feature_extractor_params(), 1e-4)
    ref_opt.load_state_dict(sd)


def test_weight_images_follow_the_flat_buffer_optimizer(cuda_device):
    """Regression test: ADKTModelTrainer re-seats every parameter on a view of one flat buffer and Adam updates that buffer in place,
    which does NOT move the parameters' own version counters.  The cached pre-split weight images (ops.weight_images) must still
    be rebuilt after every optimizer step: after two training steps the extractor's features computed through the images equal
    the features computed with the in-kernel split of the CURRENT weights, bit for bit."""
    from adkf_ift_b200 import dkt, ops, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(2), learning_rate=1e-2)
    model = ADKTModelTrainer(cfg)
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(cuda_device)
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(300 + i, 16, 24), dkt.get_dkt_batcher(max_num_graphs=320),
                                                False).batches[0] for i in range(2)]
    tb = task_batch.to_device(task_batch.collate_tasks(hosts, False), cuda_device)
    old_rows, old_on = ops._PRE_MIN_ROWS, ops._PRESPLIT
    ops._PRE_MIN_ROWS = 256                      # small test graphs: let the image path run (it is reserved for >= 32k rows otherwise)
    try:
        assert ops._presplit_mode()
        w = model.graph_feature_extractor.readout._weighted_mean_pooler._transformation_mlp._layers[0].weight
        before = w.detach().clone()
        for _ in range(2):
            model.train_step(tb)
        assert not torch.equal(w.detach(), before)                       # the optimizer moved the weight ...
        with torch.no_grad():
            z_img = model.embed(tb)
            assert getattr(w, "_adkf_images", None), "the image path did not run"
            hi = ops.weight_images(w, False)[0]
            ops._PRESPLIT = False
            z_plain = model.embed(tb)
        assert torch.equal(z_img, z_plain)                                # ... and the images followed it
        assert float((hi - w.detach()).abs().max()) <= float(w.detach().abs().max()) * 2 ** -10     # hi image = tf32(current weight)
    finally:
        ops._PRE_MIN_ROWS, ops._PRESPLIT = old_rows, old_on


def test_training_trajectory_same_on_the_one_cta_gemm_without_round2_paths(cuda_device):
    """Regression guard over everything the GEMM gained in round 2 (CTA pairs, 16-bit correction operands, TMA-store epilogue, pre-split
    weight images rebuilt after every optimizer step, weight-gradient orientation): four training steps from the same initial model
    on the default path and with adkf_gemm_configure(impl=1) -- every product on the round-1 one-CTA kernel (in-kernel split, tf32
    corrections, LDS/STG epilogue, no images) -- give the same loss trajectory.  (Parameters are not compared entry by entry: Adam's
    first updates are lr * sign(g), so an entry whose gradient is ~0 may legitimately move the other way.)"""
    from adkf_ift_b200 import _cabi, dkt, ops, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    lib = _cabi.load()
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(500 + i, 16, 24), dkt.get_dkt_batcher(max_num_graphs=320),
                                                False).batches[0] for i in range(3)]
    old_rows = ops._PRE_MIN_ROWS
    ops._PRE_MIN_ROWS = 256                       # let the image path run on the small test graphs

    def run(impl):
        prev = lib.adkf_gemm_configure(impl, -1)
        try:
            cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(2), learning_rate=1e-3)
            model = ADKTModelTrainer(cfg)
            synthetic.deterministic_init_(model, seed=0, alpha=0.5)
            model.to(cuda_device)
            tb = task_batch.to_device(task_batch.collate_tasks(hosts, False), cuda_device)
            losses = [float(model.train_step(tb).loss_per_sample.mean()) for _ in range(4)]
            used_images = any(getattr(p, "_adkf_images", None) for p in model.parameters())
            return losses, used_images
        finally:
            lib.adkf_gemm_configure(prev & 0xff, prev >> 8)

    try:
        new, img_new = run(2)
        old, img_old = run(1)
    finally:
        ops._PRE_MIN_ROWS = old_rows
    print("loss trajectories: default", new, "| one-CTA round-1 path", old)
    assert img_new and not img_old
    assert new[-1] != new[0]                                        # the model moved
    scale = max(abs(x) for x in old)                               # (the loss crosses zero on the way down: judge on the trajectory's scale)
    for a, b in zip(new, old):
        assert abs(a - b) <= 2e-4 * scale, (new, old)
