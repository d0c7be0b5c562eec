"""Development aid: whole-model hypergradient of one golden case with each tensor-core path switched to its SIMT/cuBLAS
twin in turn, to localise accuracy differences.  usage: python tools/grad_diag.py [case] [kernel]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import test_gpu_model as T  # noqa: E402
from adkf_ift_b200 import ops  # noqa: E402
from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch  # noqa: E402


def cublas_linear(x, weight, bias=None, act=None):
    y = torch.nn.functional.linear(x, weight, bias)
    if ops.ACT_IDS[act] == 1:
        y = torch.relu(y)
    elif ops.ACT_IDS[act] == 2:
        y = torch.nn.functional.leaky_relu(y, 0.01)
    return y


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "s64_rbf_l2"
    kernel = sys.argv[2] if len(sys.argv) > 2 else "rbf"
    dev = torch.device("cuda:0")
    torch.backends.cuda.matmul.allow_tf32 = False
    g, model, host, tb = T.build(os.path.join(ROOT, "tests", "golden"), name, kernel, dev)
    names = [str(x) for x in g["grad_names"]]
    params = dict(model.named_parameters())
    phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
    orig_lin, orig_pna, orig_sq, orig_gemm = ops.linear_tc, ops.pna_message_passing, ops.gp_sqdist, ops.gemm_tf32x3

    def gemm_sel(forms):
        def f(a, trans_a, b, trans_b, bias=None, act=None, out=None):
            hit = (int(trans_a), int(trans_b)) in forms
            if not trans_a and not trans_b and any(isinstance(f_, str) for f_ in forms):
                N_, K_ = b.shape
                cat = "fc" if N_ == 2048 else "proj" if (N_, K_) == (128, 3072) else "boom1" if (N_, K_) == (1024, 128) else \
                    "boom2" if (N_, K_) == (128, 1024) else "readout"
                hit = cat in forms
            if not hit:
                return orig_gemm(a, trans_a, b, trans_b, bias, act, out)
            y = (a.t() if trans_a else a) @ (b if trans_b else b.t())
            if bias is not None:
                y = y + bias
            if ops.ACT_IDS[act] == 1:
                y = y.clamp_min(0)
            elif ops.ACT_IDS[act] == 2:
                y = torch.where(y > 0, y, 0.01 * y)
            return y
        return f

    grads = {}
    TAGS = ("all_tc", "fwd_cublas", "dx_cublas", "dw_cublas", "linear_cublas", "all_simt") if len(sys.argv) <= 3 else tuple(sys.argv[3:])
    for tag in TAGS:
        ops.gemm_tf32x3 = gemm_sel({"fwd_cublas": [(0, 0)], "dx_cublas": [(0, 1)], "dw_cublas": [(1, 1)]}.get(tag, [tag[4:]] if tag.startswith("cat_") else []))
        ops.linear_tc = cublas_linear if tag in ("linear_cublas", "all_simt") else orig_lin
        ops.pna_message_passing = (lambda h, W, b, gi, impl="simt": orig_pna(h, W, b, gi, "simt")) \
            if tag in ("pna_simt", "all_simt") else orig_pna
        ops.gp_sqdist = (lambda Z, layout, qq_block, impl="simt", stat=ops.STAT_SQDIST: orig_sq(Z, layout, qq_block, "simt", stat)) \
            if tag in ("sqdist_simt", "all_simt") else orig_sq
        model.zero_grad(set_to_none=True)
        res = run_on_task_batch(model, tb, train=True, grad_scale=1.0, phi=phi)
        F = float(res.outer_loss.item())
        norms = np.array([params[n].grad.double().norm().item() if params[n].grad is not None else 0.0 for n in names])
        dots = np.array([(params[n].grad.double().cpu() * T.probe_vector(n, params[n].shape)).sum().item()
                         if params[n].grad is not None else 0.0 for n in names])
        ref_n, ref_d = g["grad_norms"], g["grad_dots"]
        big = ref_n > 1e-3 * ref_n.max()
        e_norm = np.abs(norms - ref_n)[big] / ref_n[big]
        e_dot = np.abs(dots - ref_d) / np.maximum(ref_n, 1e-3 * ref_n.max()) / np.sqrt([params[n].numel() for n in names])
        print(f"{tag:14s} F {F:.6f} (ref {float(g['train_F']):.6f}) norm err {e_norm.max():.2e} @ {np.array(names)[big][e_norm.argmax()]}"
              f"  dot err {e_dot.max():.2e} @ {names[int(e_dot.argmax())]}")
        for n in ("graph_feature_extractor.gnn.gnn_blocks.0.alpha", "graph_feature_extractor.init_node_proj.weight", "fc.2.weight"):
            got, want = params[n].grad.cpu().numpy(), g["grad::" + n] if ("grad::" + n) in g.files else None
            if want is not None:
                print(f"      {n}: rel {T.rel(got, want):.3e}  abs/maxnorm {np.abs(got - want).max() / ref_n.max():.3e} |g| {np.linalg.norm(want):.3e} maxnorm {ref_n.max():.3e}")
        grads[tag] = {n: params[n].grad.detach().double().cpu() for n in names if params[n].grad is not None}
    if "all_simt" not in grads:
        return
    base = grads["all_simt"]
    for tag in grads:
        worst = max(((grads[tag][n] - base[n]).norm().item() / max(base[n].norm().item(), 1e-30), n) for n in base)
        print(f"{tag:14s} vs all_simt: worst per-tensor rel l2 diff {worst[0]:.3e} @ {worst[1]}")


if __name__ == "__main__":
    main()
