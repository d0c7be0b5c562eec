"""Every GEMM call of one C2 meta-training step with its shape and device time (CUDA events around each call), aggregated by shape.
usage: python tools/gemm_trace.py [tasks=16]        (ADKF_GEMM_TMA_STORE=0 / ADKF_GEMM_CORR=... for A/B runs)"""
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    tasks = int(sys.argv[1]) if len(sys.argv) > 1 else bench.TASKS_PER_GPU
    dev = torch.device("cuda:0")
    from adkf_ift_b200 import ops, task_batch
    model = bench.make_model(dev)
    tb = task_batch.to_device(task_batch.collate_tasks(bench.host_tasks(1000, tasks), False), dev)
    log = []
    on = [False]

    def wrap(name, fn, describe):
        def inner(*a, **k):
            if not on[0]:
                return fn(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            log.append((describe(*a, **k), e0, e1))
            return out
        setattr(ops, name, inner)

    def d_plain(a, ta, b, tb_, bias=None, act=None, out=None, weight_b=False):
        K, M = (a.shape if ta else a.shape[::-1])
        N = b.shape[1] if tb_ else b.shape[0]
        form = "(T,T) dW" if ta else ("(K,T) dX" if tb_ else "(K,K) fwd")
        return f"{form:10s} M={M:<7d} N={N:<5d} K={K:<7d} {'bias ' if bias is not None else ''}{act or ''}"

    def d_masked(a, b, mask, mask_act, weight_b=False):
        return f"(K,T) dX   M={a.shape[0]:<7d} N={b.shape[1]:<5d} K={a.shape[1]:<7d} masked"

    def d_batched(a, ta, ams, aks, b, tb_, bms, bks, M, N, K, batch, out, ldd, d_stride, bias=None, bias_stride=0, act=None, allow_splitk=False):
        form = "(T,T) dW" if ta else ("(K,T) dX" if tb_ else "(K,K) fwd")
        return f"{form:10s} M={M:<7d} N={N:<5d} K={K:<7d} x{batch} batched"

    wrap("gemm_tf32x3", ops.gemm_tf32x3, d_plain)
    wrap("gemm_tf32x3_masked", ops.gemm_tf32x3_masked, d_masked)
    wrap("gemm_tf32x3_batched", ops.gemm_tf32x3_batched, d_batched)
    for _ in range(2):
        model.train_step(tb)
    torch.cuda.synchronize()
    on[0] = True
    model.train_step(tb)
    torch.cuda.synchronize()
    on[0] = False
    agg = collections.OrderedDict()
    for desc, e0, e1 in log:
        c, t = agg.get(desc, (0, 0.0))
        agg[desc] = (c + 1, t + e0.elapsed_time(e1))
    tot = sum(t for _, t in agg.values())
    print(f"{len(log)} GEMM calls, {tot:.2f} ms (events around each call: includes the split-K reduce and ~5 us of launch gap per call)")
    for desc, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {t:7.3f} ms  x{c:<3d} {t / c:7.3f} each  {desc}")


if __name__ == "__main__":
    main()
