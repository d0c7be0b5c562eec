"""Host-side cost of one meta-training step (cProfile over N steps at a small task count, where the step is launch-bound).
usage: python tools/host_profile.py [tasks=2] [steps=10]"""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    tasks = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    dev = torch.device("cuda:0")
    from adkf_ift_b200 import _cabi, task_batch
    lib = _cabi.load()
    model = bench.make_model(dev)
    tb = task_batch.to_device(task_batch.collate_tasks(bench.host_tasks(1000, tasks), False), dev)
    for _ in range(3):
        model.train_step(tb)
    torch.cuda.synchronize()
    l0 = lib.adkf_launch_count()
    t0 = time.perf_counter()
    for _ in range(steps):
        model.train_step(tb)
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print(f"{tasks} tasks: host issue time {1e3 * t_host / steps:.2f} ms/step, wall {1e3 * t_all / steps:.2f} ms/step, "
          f"own launches/step {(lib.adkf_launch_count() - l0) // steps}")
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(steps):
        model.train_step(tb)
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)


if __name__ == "__main__":
    main()
