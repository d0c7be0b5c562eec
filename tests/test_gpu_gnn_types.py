"""-m gpu parity of the non-default GNN types (`--gnn_type MultiAggr`, `Plain`, `MultiHeadAttention`;
fs_mol/modules/gnn.py:409-447, 152-165, 197-240, 268-387) against fixtures produced by the reference's own GraphFeatureExtractor
(tests/golden/make_golden_gnn_types.py): features <= 1e-4 relative, per-tensor gradient norms and probe dots of a fixed
probe loss <= 1e-3 (fp32 against fp32, 87 tensors)."""
import os
import zlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def probe_vector(name, shape):
    g = torch.Generator().manual_seed(zlib.crc32(("probe:" + name).encode()) & 0x7FFFFFFF)
    return torch.randn(shape, generator=g, dtype=torch.float64)


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


@pytest.mark.parametrize("gnn_type", ["MultiAggr", "Plain", "MultiHeadAttention"])
def test_gnn_type_features_and_gradients(cuda_device, golden_dir, gnn_type):
    from adkf_ift_b200 import dkt, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    g = np.load(os.path.join(golden_dir, "gnn_types.npz"))
    task_id, n_s, n_q, layers = (int(x) for x in g["meta"])
    model = ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(layers, gnn_type)))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(cuda_device)
    host = dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(task_id, n_s, n_q), dkt.get_dkt_batcher(max_num_graphs=320),
                                              False)
    tb = task_batch.to_device(task_batch.collate_tasks([host.batches[0]], False), cuda_device)
    fe = model.graph_feature_extractor
    fe.zero_grad(set_to_none=True)
    z = fe(tb)[:n_s]                                   # support graphs come first in the task's rows
    want = g[f"{gnn_type}::features"]
    e = rel(z.detach().cpu().numpy(), want)
    print(gnn_type, "feature rel err", e)
    assert e < 1e-4
    probe = probe_vector("features", want.shape).to(torch.float32).to(cuda_device)
    (z * probe).sum().backward()
    names = [str(x) for x in g[f"{gnn_type}::grad_names"]]
    params = dict(fe.named_parameters())
    norms = np.array([params[n].grad.double().norm().item() for n in names])
    dots = np.array([(params[n].grad.double().cpu() * probe_vector(n, params[n].shape)).sum().item() for n in names])
    ref_n, ref_d = g[f"{gnn_type}::grad_norms"], g[f"{gnn_type}::grad_dots"]
    big = ref_n > 1e-3 * ref_n.max()
    e_norm = (np.abs(norms - ref_n)[big] / ref_n[big]).max()
    e_dot = (np.abs(dots - ref_d) / np.maximum(ref_n, 1e-3 * ref_n.max()) / np.sqrt([params[n].numel() for n in names])).max()
    print(gnn_type, "max rel grad-norm err", e_norm, "max probe-dot err", e_dot)
    assert e_norm < 1e-3 and e_dot < 1e-3
    tower = "0" if gnn_type == "MultiHeadAttention" else "2"
    for n in ("gnn.gnn_blocks.0.alpha", "init_node_proj.weight", f"gnn.gnn_blocks.1.mp_layers.{tower}.message_fns.0._layers.0.weight"):
        got, ref = params[n].grad.cpu().numpy(), g[f"{gnn_type}::grad::{n}"]
        assert rel(got, ref) < 1e-3 or np.abs(got - ref).max() < 1e-5 * ref_n.max(), (n, rel(got, ref))


def test_unknown_gnn_type(cuda_device):
    from adkf_ift_b200.feature_extractor import GraphFeatureExtractor, cli_default_config
    with pytest.raises(ValueError):
        GraphFeatureExtractor(cli_default_config(2, "NoSuchType"))
