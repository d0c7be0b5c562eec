"""Development aid: where does a forward-GEMM perturbation enter the hypergradient?  Compares Z, dL/dZ and the theta-gradient
between the tensor-core forward and a cuBLAS forward of one layer class, and crosses them."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import test_gpu_model as T  # noqa: E402
from adkf_ift_b200 import gp as gpk, ops  # noqa: E402
from adkf_ift_b200.adaptive_dkt_utils import _fit_or_fixed  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


def main():
    dev = torch.device("cuda:0")
    g, model, host, tb = T.build(os.path.join(ROOT, "tests", "golden"), "s64_rbf_l2", "rbf", dev)
    cfg = model.config
    phi = torch.from_numpy(g["train_phi_star"].astype(np.float32)).view(1, 3)
    orig = ops.gemm_tf32x3

    def boom2_cublas(a, trans_a, b, trans_b, bias=None, act=None, out=None):
        if not trans_a and not trans_b and tuple(b.shape) == (128, 1024):
            y = a @ b.t()
            return y + bias if bias is not None else y
        return orig(a, trans_a, b, trans_b, bias, act, out)

    out = {}
    captured = {}
    blocks = model.graph_feature_extractor.gnn.gnn_blocks

    def mk(i):
        def hook(mod, inp, o):
            o.retain_grad()
            captured[(cur[0], i)] = o
        return hook
    cur = [None]
    for i, blk in enumerate(blocks):
        blk.register_forward_hook(mk(i))
    for tag, fn in (("tc", orig), ("boom2_cublas", boom2_cublas)):
        ops.gemm_tf32x3 = fn
        cur[0] = tag
        model.train()
        model.zero_grad(set_to_none=True)
        Z = model.embed(tb)
        gp = gpk.BatchedExactGP(Z, tb.labels, tb.layout, cfg.gp_kernel, cfg.use_numeric_labels, cfg.use_lengthscale_prior, qq_block=True)
        fit = _fit_or_fixed(gp, phi, 50, 2e-6)
        dZ, aux, status = gp.hypergradient(fit, mode=0)
        out[tag] = dict(Z=Z, dZ=dZ.clone(), aux=aux.clone(), status=int(status.item()), D2=gp.sq.D2.clone())
    a, b = out["tc"], out["boom2_cublas"]
    n = len(g["z_support"])
    zref = torch.from_numpy(np.concatenate([g["z_support"], g["z_query"]]))
    print("status", a["status"], b["status"])
    print("Z  tc vs cublas-boom2", rel(a["Z"], b["Z"]), " tc vs golden", rel(a["Z"].cpu(), zref), " boom2 vs golden", rel(b["Z"].cpu(), zref))
    print("D2 tc vs boom2", rel(a["D2"], b["D2"]))
    print("dZ tc vs boom2", rel(a["dZ"], b["dZ"]))
    dref = torch.from_numpy(np.concatenate([g["fixed_f32_dL_dzs"], g["fixed_f32_dL_dzq"]]))
    d64 = torch.from_numpy(np.concatenate([g["fixed_f64_dL_dzs"], g["fixed_f64_dL_dzq"]]))
    print("dZ tc vs golden f32", rel(a["dZ"].cpu(), dref), " boom2 vs golden f32", rel(b["dZ"].cpu(), dref), " golden f32 vs f64", rel(dref, d64))
    print("aux tc   ", a["aux"][0, :20].cpu().numpy())
    print("aux boom2", b["aux"][0, :20].cpu().numpy())
    post(captured, out, tb)
    # cross: backward of each graph with each dZ
    names = [str(x) for x in g["grad_names"]]
    params = dict(model.named_parameters())
    res = {}
    for zt in ("tc", "boom2_cublas"):
        for dt in ("tc", "boom2_cublas"):
            model.zero_grad(set_to_none=True)
            out[zt]["Z"].backward(out[dt]["dZ"], retain_graph=True)
            res[(zt, dt)] = {k: params[k].grad.detach().double().clone() for k in names if params[k].grad is not None}
    base = res[("boom2_cublas", "boom2_cublas")]
    for key, r in res.items():
        worst = max(((r[k] - base[k]).norm().item() / max(base[k].norm().item(), 1e-30), k) for k in base if base[k].norm() > 1e-3)
        print("graph", key[0], "dZ", key[1], "worst per-tensor l2 diff vs (boom2,boom2)", f"{worst[0]:.3e}", worst[1])




def post(captured, out, tb):
    import torch
    for zt in ("tc", "boom2_cublas"):
        for k, o in captured.items():
            if o.grad is not None:
                o.grad = None
    res = {}
    for zt in ("tc", "boom2_cublas"):
        out[zt]["Z"].backward(out["tc"]["dZ"], retain_graph=True)
        for (t, i), o in captured.items():
            if t == zt:
                res[(zt, i)] = o.grad.detach().double().clone()
    for i in sorted({i for (_, i) in captured}):
        a, b = res[("tc", i)], res[("boom2_cublas", i)]
        d = (a - b)
        pn = d.pow(2).sum(1)
        top = torch.topk(pn, 5)
        print(f"block {i}: d(out) grad rel l2 diff {d.norm() / b.norm():.3e}; top-5 node share of diff energy {float(top.values.sum() / pn.sum()):.3f}; "
              f"nodes {top.indices.tolist()}; fwd rel diff {float((captured[('tc', i)].double() - captured[('boom2_cublas', i)].double()).norm() / captured[('boom2_cublas', i)].double().norm()):.3e}")
        if i == 0:
            v0 = int(top.indices[0])
            ch = torch.topk(d[v0].abs(), 4).indices.tolist()
            print("         node", v0, "channels", ch, "grad tc", a[v0, ch].tolist(), "grad ref", b[v0, ch].tolist())
            ntg = tb.node_to_graph
            gid = int(ntg[v0])
            members = (ntg == gid).nonzero().flatten().tolist()
            print("         graph", gid, "members", members[0], "..", members[-1])
            for zt in ("tc", "boom2_cublas"):
                o = captured[(zt, 0)].detach()
                for c in ch[:2]:
                    col = o[members, c]
                    srt = torch.sort(col, descending=True)
                    print(f"         {zt} ch {c}: top-3 values {[f'{x:.9g}' for x in srt.values[:3].tolist()]} at nodes {[members[j] for j in srt.indices[:3].tolist()]}")
            import itertools
            idx = top.indices.tolist()
            for zt in ("tc", "boom2_cublas"):
                o = captured[(zt, 0)].detach()
                for x, y in itertools.combinations(idx, 2):
                    dd = float((o[x] - o[y]).abs().max())
                    if dd < 1e-3:
                        print(f"         {zt}: rows {x},{y} max |diff| {dd:.3e}  n_equal {int((o[x] == o[y]).sum())}/128  |row| {float(o[x].abs().max()):.3f}")
        cosd = float((d * b).sum() / (d.norm() * b.norm()))
        print(f"         cos(diff, grad) {cosd:+.3f}")


if __name__ == "__main__":
    main()
