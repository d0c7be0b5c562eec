"""One warm-up step + N profiled steps of the C2 workload (what the ncu command lines in profiles/README.md run).
usage: python tools/one_step.py [tasks=16] [train|test] [steps=1]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402


def main():
    tasks = int(sys.argv[1]) if len(sys.argv) > 1 else bench.TASKS_PER_GPU
    mode = sys.argv[2] if len(sys.argv) > 2 else "train"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    dev = torch.device("cuda:0")
    from adkf_ift_b200 import _cabi, task_batch
    from adkf_ift_b200.adaptive_dkt_utils import run_on_task_batch
    lib = _cabi.load()
    model = bench.make_model(dev)
    hb = task_batch.collate_tasks(bench.host_tasks(1000, tasks), False)

    def step():
        tb = task_batch.to_device(hb, dev)
        if mode == "train":
            return model.train_step(tb).loss_per_sample.mean()
        return run_on_task_batch(model, tb, train=False).preds.mean()

    step()
    torch.cuda.synchronize()
    l0 = lib.adkf_launch_count()
    for _ in range(steps):
        out = step()
    torch.cuda.synchronize()
    print(f"ok mode={mode} tasks={tasks} out={float(out):.6f} own_launches_per_step={(lib.adkf_launch_count() - l0) // steps}")


if __name__ == "__main__":
    main()
