"""Development aid: run two C2 training steps (graph index build included, as bench.py's resident step does) with
torch.cuda.set_sync_debug_mode("warn") and print where the host synchronises with the device."""
import os
import sys
import warnings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from adkf_ift_b200 import ops, task_batch  # noqa: E402

dev = torch.device("cuda:0")
model = bench.make_model(dev)
hb = task_batch.collate_tasks(bench.host_tasks(1000, bench.TASKS_PER_GPU), False)
tb = task_batch.to_device(hb, dev, build_index=False)


def step():
    tb.graph_index = ops.build_graph_index(tb.adjacency_lists, tb.num_nodes, tb.node_to_graph, tb.num_graphs)
    return model.train_step(tb)


for _ in range(2):
    step()
torch.cuda.synchronize()
torch.cuda.set_sync_debug_mode("warn")
with warnings.catch_warnings(record=True) as w:
    warnings.simplefilter("always")
    step()
torch.cuda.set_sync_debug_mode("default")
torch.cuda.synchronize()
print(f"{len(w)} synchronising calls in one step")
for x in w:
    print(f"  {x.filename}:{x.lineno}: {str(x.message)[:100]}")
import time
for label, n in (("with index build", 5),):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n):
        step()
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"{label}: host enqueue {1e3 * (t1 - t0) / n:.1f} ms/step, total {1e3 * (t2 - t0) / n:.1f} ms/step")
