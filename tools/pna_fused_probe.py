"""Times the fused PNA message-passing kernels (pna_fused.cu) alone on the C2 graph batch (16 tasks: V ~ 134k nodes, 2E ~ 296k
messages) and the three-kernel chain they replace, CUDA events, L2 flushed between repetitions by the 0.5 GB outputs themselves.
usage: python tools/pna_fused_probe.py [tasks=16] [reps=10]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def main():
    tasks = int(sys.argv[1]) if len(sys.argv) > 1 else bench.TASKS_PER_GPU
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    dev = torch.device("cuda:0")
    from adkf_ift_b200 import ops, task_batch
    hb = task_batch.collate_tasks(bench.host_tasks(1000, tasks), False)
    tb = task_batch.to_device(hb, dev)
    gi = tb.graph_index
    V, E = gi.num_nodes, gi.num_edges
    gen = torch.Generator(device=dev).manual_seed(0)
    h = torch.randn(V, 128, device=dev, generator=gen)
    W = torch.randn(4, 3, 192, 64, device=dev, generator=gen) * 0.15
    b = torch.randn(4, 3, 192, device=dev, generator=gen) * 0.1
    g = torch.randn(V, 1024, device=dev, generator=gen)
    print(f"V={V} E={E} max_degree={gi.max_degree} fused_ok={gi.fused_ok} types={gi.type_sorted.type_off}")
    for fused in (True, False):
        hc, Wc, bc = (x.clone().requires_grad_(True) for x in (h, W, b))
        out = [None]

        def fwd():
            out[0] = ops.multiaggr_message_passing(hc, Wc, bc, gi, fused=fused)

        def bwd():
            torch.autograd.grad(out[0], (hc, Wc, bc), g, retain_graph=True)

        f_med, f_min = timed(fwd, reps)
        b_med, b_min = timed(bwd, reps)
        flops = 2.0 * E * 4 * 192 * 64
        if fused:
            from adkf_ift_b200 import _cabi
            lib = _cabi.load()
            dbg = torch.zeros(160, 16, dtype=torch.int64, device=dev)
            for name, fn in (("forward", fwd), ("backward", bwd)):
                dbg.zero_()
                lib.adkf_pna_fused_set_debug(ctypes.c_void_p(dbg.data_ptr()))
                fn()
                torch.cuda.synchronize()
                lib.adkf_pna_fused_set_debug(ctypes.c_void_p(0))
                d = dbg[:148].double().mean(0).cpu().numpy() / 1e3
                print(f"  {name} kcycles/CTA: mma wait-A {d[0]:.0f} wait-acc {d[1]:.0f} total {d[2]:.0f} | gather wait-mma {d[3]:.0f} work {d[4]:.0f} "
                      f"total {d[5]:.0f} | epilogue wg0 bar {d[6]:.0f} wait-acc {d[7]:.0f} drain {d[8]:.0f} nodes {d[9]:.0f} total {d[10]:.0f} | "
                      f"wg1 bar {d[11]:.0f} wait-acc {d[12]:.0f} drain {d[13]:.0f} nodes {d[14]:.0f} total {d[15]:.0f}")
        print(f"{'fused' if fused else 'chain'}: forward {f_med:.3f} ms (min {f_min:.3f}; {flops / f_med / 1e9:.1f} TFLOP/s of the message MLP), "
              f"backward {b_med:.3f} ms (min {b_min:.3f})")


if __name__ == "__main__":
    main()
