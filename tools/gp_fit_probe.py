"""Development probe: duration and per-task Newton iterations of adkf_gp_fit / adkf_gp_ift on the C2 bench batch."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402
from adkf_ift_b200 import gp as gpk, ops, task_batch  # noqa: E402

dev = torch.device("cuda:0")
model = bench.make_model(dev)
tb = task_batch.to_device(task_batch.collate_tasks(bench.host_tasks(1000, 16), False), dev)
with torch.no_grad():
    Z = model.embed(tb)
g = gpk.BatchedExactGP(Z, tb.labels, tb.layout, "matern", False, True, qq_block=True)


def t(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps, r


ms, fit = t(lambda: g.fit())
print("gp_fit ms", ms, "iters", fit.aux[:, ops.AUX_ITERS].cpu().numpy(), "nfev", fit.aux[:, ops.AUX_NFEV].cpu().numpy(), "status",
      fit.status.cpu().numpy())
ms2, _ = t(lambda: g.hypergradient(fit))
print("gp_ift + sqdist_bwd ms", ms2)
ms3, _ = t(lambda: g.fit(max_iter=0))
print("gp_fit(max_iter=0) ms", ms3)
