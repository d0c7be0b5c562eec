"""Per-kernel device-time breakdown of one C2 meta-training step (torch.profiler / CUPTI; development aid --
numbers that are reported come from bench.py's CUDA events and the ncu captures under profiles/)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

import bench  # noqa: E402


def main():
    tasks = int(sys.argv[1]) if len(sys.argv) > 1 else bench.TASKS_PER_GPU
    mode = sys.argv[2] if len(sys.argv) > 2 else "train"
    dev = torch.device("cuda:0")
    from adkf_ift_b200 import task_batch
    from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch
    model = bench.make_model(dev)
    hb = task_batch.collate_tasks(bench.host_tasks(1000, tasks), False)
    tb = task_batch.to_device(hb, dev)

    def step():
        if mode == "train":
            model.train_step(tb)
        else:
            run_on_task_batch(model, tb, train=False)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step()
        torch.cuda.synchronize()
    rows = [(e.key, e.device_time_total, e.count) for e in prof.key_averages() if e.device_time_total > 0
            and e.device_type == torch.autograd.DeviceType.CUDA]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"total device kernel time {tot / 1e3:.2f} ms over {sum(r[2] for r in rows)} launches; nodes {tb.num_nodes} "
          f"edges {tb.graph_index.num_edges} graphs {tb.num_graphs}")
    for k, t, c in rows[:40]:
        print(f"{t / 1e3:10.3f} ms {100 * t / tot:6.2f}% x{c:<5d} {k[:110]}")
    # idle gaps of the device between consecutive kernels (where the host did not keep the stream fed)
    evs = sorted(((e.time_range.start, e.time_range.end, e.name) for e in prof.events()
                  if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start), key=lambda x: x[0])
    if evs:
        span = evs[-1][1] - evs[0][0]
        gaps = []
        end = evs[0][1]
        for i in range(1, len(evs)):
            g = evs[i][0] - end
            if g > 0:
                gaps.append((g, evs[i - 1][2][:50], evs[i][2][:50], (evs[i][0] - evs[0][0]) / 1e3))
            end = max(end, evs[i][1])
        idle = sum(g[0] for g in gaps)
        print(f"device span {span / 1e3:.2f} ms, idle between kernels {idle / 1e3:.2f} ms in {len(gaps)} gaps; "
              f"gaps > 20 us: {sum(g[0] for g in gaps if g[0] > 20) / 1e3:.2f} ms")
        for g in sorted(gaps, key=lambda x: -x[0])[:12]:
            print(f"    {g[0]:8.1f} us at {g[3]:7.2f} ms  after {g[1]}  before {g[2]}")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        step()
    e1.record()
    torch.cuda.synchronize()
    print(f"wall (events) {e0.elapsed_time(e1) / 3:.2f} ms/step, peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")


if __name__ == "__main__":
    main()
