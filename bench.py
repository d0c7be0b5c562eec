#!/usr/bin/env python
"""bench.py -- FS-Mol-shaped tasks/sec (GP fit + IFT hypergradient) of the ADKF-IFT meta-training step.

    python bench.py --gpus N --steps K --warmup W            # this repo's B200 path
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU path (oracle) on the host cores

Workload (BASELINE.json configs[1]): ADKTModel meta-training step, 16 tasks per GPU per step, support 64, query 256,
gnn+ecfp+fc features (PNA 10 layers, H=128; FC 2560->2048->2048), Matern-5/2 GP, inner fit + IFT hypergradient
+ one backward through the extractor, NCCL all-reduce of the flat 25.3 M-float meta-gradient, clip 1.0, Adam.
Weak scaling: every rank processes its own 16 tasks per step (global meta-batch 16 N).

One JSON line on rank 0 (contract in the task statement): value = device-resident tasks/s, e2e = the same step
through to_device() from pinned host memory + a host read of the loss, roofline for the dominant own kernel
(timed live with CUDA events), cpu_baseline = oracle on the host cores (N=1 only), clocks, gpu_launches.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

TASKS_PER_GPU = 16
SUPPORT, QUERY = 64, 256
METRIC = "FS-Mol-shaped meta-training tasks/sec (GP fit + IFT hypergradient)"
UNIT = "tasks/s"


def tasks_per_gpu(n_gpus, scaling="weak"):
    """weak: 16 tasks per GPU per step (global meta-batch 16 N); strong: BASELINE config 2 as written -- a global meta-batch
    of 16 tasks split over the GPUs (2 per GPU at N = 8)."""
    if scaling == "strong":
        if TASKS_PER_GPU % n_gpus:
            raise SystemExit(f"--scaling strong needs a number of GPUs that divides {TASKS_PER_GPU}")
        return TASKS_PER_GPU // n_gpus
    return TASKS_PER_GPU


def workload_config(n_gpus, scaling="weak"):
    per = tasks_per_gpu(n_gpus, scaling)
    return {"workload": f"C2: ADKTModel meta-training step, {per} tasks/GPU/step, support 64, query 256, gnn+ecfp+fc, "
                        "matern, IFT hypergradient, clip 1.0, Adam",
            "tasks_per_step": per * n_gpus, "support": SUPPORT, "query": QUERY,
            "parallelism": f"task-sharded dp{n_gpus}, one NCCL all-reduce of the 101 MB meta-gradient per step",
            "l2": "inputs+weights (~160 MB) and ~20 GB of per-step activations exceed the 126 MB L2; no explicit flush"}


def make_model(device, num_layers=10):
    from adkf_ift_b200 import synthetic
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(num_layers), gp_kernel="matern",
                                 tasks_per_batch=TASKS_PER_GPU, support_set_size=SUPPORT, query_set_size=QUERY,
                                 learning_rate=1e-4, clip_value=1.0)
    model = ADKTModelTrainer(cfg)
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    return model.to(device)


def host_tasks(first_task, num_tasks):
    from adkf_ift_b200 import dkt, synthetic
    out = []
    for t in range(num_tasks):
        s = synthetic.make_task_sample(first_task + t, SUPPORT, QUERY)
        out.append(dkt.task_sample_to_dkt_task_sample(s, dkt.get_dkt_batcher(max_num_graphs=SUPPORT + QUERY), False).batches[0])
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "200"], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def timed_region(fn, steps, world, device):
    """barrier + sync, K steps between CUDA events on the current stream, barrier + sync; max over ranks (ms)."""
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return float(ms.item())


def _time_ms(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def dominant_kernel_roofline(model, tb, peaks, tf32_peak=None):
    """The dominant kernel of the step is the tcgen05 3xTF32 GEMM (gemm_tf32x3_kernel, ~60% of the device time over its
    three operand forms, profiles/r1c_*): timed here alone, with CUDA events, on its most expensive launch of the step --
    the readout hidden layer over all V nodes of the 16-task batch (M = V, N = 768, K = 1408, bias + ReLU fused).
    achieved = ALGORITHMIC flops 2 M N K (fp32-equivalent; the kernel issues 2 TF32-equivalents per product, reported beside it)
    over the measured duration; peak = TF32 dense = half of the measured bf16 peak of MEASURED_PEAKS.json.
    A second entry gives the HBM-bound PNA aggregation kernel of the step (reads the message pre-activations pre [E,768],
    writes the unscaled aggregate [V,1024] + 2 x 256 aux bytes per node; the degree scalers are folded into the projection)."""
    from adkf_ift_b200 import ops
    gi = tb.graph_index
    V, E = gi.num_nodes, gi.num_edges
    dev = tb.node_features.device
    M, N, K = V, 768, 1408
    x, w, b = torch.randn(M, K, device=dev), torch.randn(N, K, device=dev) / K ** 0.5, torch.randn(N, device=dev)
    # exactly the call the step issues for this layer (weight_b: the B operand is a weight -> its pre-split images, built once per
    # weight version, feed the MMA; the first untimed call of _time_ms builds them)
    ms = _time_ms(lambda: ops.gemm_tf32x3(x, False, w, False, b, "relu", weight_b=True))
    flops = 2.0 * M * N * K
    achieved = flops / (ms * 1e-3) / 1e12
    bf16 = peaks.get("bf16_tflops")
    if tf32_peak is not None:
        peak, src = tf32_peak, ("measured in this run: torch.matmul allow_tf32 8192^3, best of 10, CUDA events (the MEASURED_PEAKS.json "
                                f"recipe; bf16 there: {bf16} TFLOP/s)")
    elif bf16 is not None:
        peak, src = bf16 / 2.0, "MEASURED_PEAKS.json bf16_tflops / 2 (TF32 runs at half the bf16 rate)"
    else:
        peak, src = 2250.0 * 0.73 / 2.0, "fallback (B200_PROFILING.md nominal bf16 x 0.73) / 2"
    roof = {"kernel": "gemm_pair_kernel<K,K; scaled-fp16 corrections; pre-split weight images> (cta_group::2; readout hidden layer, M=V N=768 "
                      "K=1408, bias+ReLU, TMA-store epilogue)", "bound": "tensor",
            "achieved": round(achieved, 1), "peak": round(peak, 1), "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
            "traffic": 1.148e9, "traffic_source": "ncu --set full, profiles/r2w_ncu_full_summary.md (dram read+write per launch of this shape)",
            "peak_source": src, "launch_ms": round(ms, 4), "algorithmic_flops": flops,
            # tensor work issued per product: one tf32 MMA (hi.hi) + one kind::f16 chain over [lo | hi] x [hi | lo] at twice the tf32
            # rate = 2 tf32-equivalents (3 with tf32 corrections): the scheme's ceiling is frac 0.5
            "issued_tf32_tflops": round(2 * achieved, 1), "frac_issued": round(2 * achieved / peak, 4),
            "algorithmic_bytes": 4.0 * (M * K + N * K + M * N), "M": M, "N": N, "K": K}
    # secondary: the HBM-bound aggregation kernel of one GNN block
    del x, w, b
    blk = model.graph_feature_extractor.gnn.gnn_blocks[0]
    with torch.no_grad():
        W, bb = blk.stacked_message_weights()
        h = torch.randn(V, 128, device=dev)
        from adkf_ift_b200 import _cabi
        lib = _cabi.load()
        ts = gi.type_sorted
        pre = torch.randn(max(E, 1), 768, device=dev)
        out = torch.empty(V, 1024, device=dev)
        cnt = torch.empty(V, 4, 64, dtype=torch.uint8, device=dev)
        arg = torch.empty(V, 4, 64, dtype=torch.uint8, device=dev)
        ms2 = _time_ms(lambda: ops.check(lib.adkf_mp_aggregate_fwd(ops.dptr(pre), ops.dptr(ts.tptr), V, ops.MP_MULTIAGGR, ops.dptr(out),
                                                                   ops.dptr(cnt), ops.dptr(arg), ops.stream_ptr()),
                                         "adkf_mp_aggregate_fwd"))
    alg2 = 4.0 * 768 * E + 4.0 * 1024 * V + 512.0 * V + 4.0 * (3 * V + 1)
    hbm = peaks.get("hbm_gbs") or 6650.0
    roof["hbm_kernel"] = {"kernel": "pna_aggregate_fwd_kernel<false> (unscaled aggregate [V,1024])", "bound": "hbm", "achieved": round(alg2 / (ms2 * 1e-3) / 1e9, 1),
                          "peak": hbm, "unit": "GB/s", "frac": round(alg2 / (ms2 * 1e-3) / 1e9 / hbm, 4), "launch_ms": round(ms2, 4),
                          "algorithmic_bytes": alg2, "nodes": V, "directed_edges": E}
    return roof


def meta_test_throughput(device, cpu_tasks=1):
    """BASELINE config 3 (reported beside the headline, not the bench metric): 256 regression meta-test tasks
    (--use-numeric-labels), support 128, query 192, ONE batched pass = extractor forward + GP fit + prediction; 8 distinct
    synthetic tasks tiled 32x (positions are independent, tests/test_gpu_model.py::test_c3_*).  The CPU figure is the oracle
    port of run_on_batches (reference algorithm, SciPy L-BFGS-B) on `cpu_tasks` of the same tasks."""
    from adkf_ift_b200 import dkt, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer, run_on_task_batch
    from adkf_ift_b200.feature_extractor import cli_default_config
    n_s, n_q, distinct, copies = 128, 192, 8, 32
    cfg = ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(10), use_numeric_labels=True)
    model = ADKTModelTrainer(cfg)
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(device)
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(300 + i, n_s, n_q),
                                                dkt.get_dkt_batcher(max_num_graphs=n_s + n_q), True) for i in range(distinct)]
    hb = task_batch.collate_tasks([hosts[i % distinct].batches[0] for i in range(distinct * copies)], True)
    for _ in range(2):
        res = run_on_task_batch(model, task_batch.to_device(hb, device), train=False)
    torch.cuda.synchronize()
    reps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        res = run_on_task_batch(model, task_batch.to_device(hb, device), train=False)
        float(res.preds.mean().item())
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    out = {"workload": "C3: 256 regression meta-test tasks, support 128, query 192, one batched pass incl. H2D of the tasks",
           "value": 256 / (ms * 1e-3), "unit": UNIT, "ms_per_pass": ms, "h2d_bytes_per_pass": hb.nbytes()}
    if cpu_tasks > 0:
        torch.set_num_threads(os.cpu_count() or 1)
        try:
            from oracle import ref_configs
            run = ref_configs.c3_setup()
            kind = "reference"
        except Exception:
            from oracle import adkf_oracle as orc
            P = {n: p.detach().cpu() for n, p in model.named_parameters() if not n.startswith("gp_")}
            ocfg = orc.OracleConfig(num_layers=10, gp_kernel="matern", use_numeric_labels=True)

            def run(task_id):
                batch = dkt.torchify(hosts[task_id % distinct], torch.device("cpu")).batches[0]
                t0 = time.perf_counter()
                orc.meta_test_task(P, batch, ocfg)
                return time.perf_counter() - t0
            kind = "port"
        run(300)                                           # warms the thread pools
        ts = sorted(run(301 + i) for i in range(cpu_tasks))
        med = ts[len(ts) // 2]
        out["cpu_baseline"] = {"value": 1.0 / med, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                               "sample": f"median of {cpu_tasks} tasks of the same workload (run_on_batches per task: reinit + "
                                         "fit_gpytorch_scipy + prediction), sequential"}
    del model
    torch.cuda.empty_cache()
    return out


def _cuda_ms(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def other_configs(device, cpu=True):
    """BASELINE configs 1, 4 and 5 (reported beside the headline; BASELINE.md section 3 promised tasks/s for all five), each
    through this package's public surface with host inputs, next to the reference's own files on the host cores."""
    from adkf_ift_b200 import bo_utils, dkt, ops, synthetic
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer, run_on_batches
    from adkf_ift_b200.feature_extractor import cli_default_config
    from adkf_ift_b200.gp_utils import fit_gpytorch_scipy
    out = {}
    torch.set_num_threads(os.cpu_count() or 1)
    ref = None
    if cpu:
        try:
            from oracle import ref_configs as ref
        except Exception:
            ref = None

    # ---- C1: classification meta-test, ONE task per call through the reference's call sequence (run_on_batches) -----------
    model = ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(10)))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    model.to(device)
    hosts = [dkt.task_sample_to_dkt_task_sample(synthetic.make_task_sample(700 + i, 16, 32), dkt.get_dkt_batcher(max_num_graphs=320), False)
             for i in range(8)]
    k = [0]

    def c1():
        dev = dkt.torchify(hosts[k[0] % 8], device)
        k[0] += 1
        run_on_batches(model, dev.batches, dev.batch_labels, dev.batch_numeric_labels, train=False)   # ends with a D2H of the predictions

    ms = _cuda_ms(c1, 16)
    out["c1_meta_test_single_task"] = {"workload": "C1: one classification task (support 16, query 32) per run_on_batches call, H2D + D2H included",
                                       "value": 1e3 / ms, "unit": UNIT, "ms_per_task": ms}
    if ref is not None:
        run = ref.c1_setup()
        run(700)
        ts = sorted(run(701 + i) for i in range(3))
        out["c1_meta_test_single_task"]["cpu_baseline"] = {"value": 1.0 / ts[1], "unit": UNIT, "cores": torch.get_num_threads(), "kind": "reference",
                                                           "sample": "median of 3 tasks through the reference's own run_on_batches"}
    del model
    # ---- C4: MoleculeNet ADKF-IFT meta-training step, 9 tasks x (20 support + 32 query), GIN d=300 ------------------------------------
    from adkf_ift_b200.moleculenet import ADKF_Meta_Trainer, ADKFModel, MolBatch, default_args, merge_mol_batches
    args = default_args()
    mmodel = ADKFModel(args)
    synthetic.deterministic_init_(mmodel, seed=1)
    mmodel.to(device)
    trainer = ADKF_Meta_Trainer(args, mmodel)
    T, n_s, n_q = 9, 20, 32

    def mol(seed, n):
        d = synthetic.make_mol_graphs(seed, n)
        return MolBatch(torch.from_numpy(d["x"]), torch.from_numpy(d["edge_index"]), torch.from_numpy(d["edge_attr"]),
                        torch.from_numpy(d["batch"]), torch.from_numpy(d["y"]), n)
    parts = []
    for t in range(T):
        parts += [mol(100 + 2 * t, n_s), mol(101 + 2 * t, n_q)]

    def c4():
        mb = merge_mol_batches(parts, device, [n_s] * T, [n_q] * T)          # host graphs -> device, every step
        res = trainer.train_step(mb)
        float(res.loss_per_sample.mean().item())

    ms = _cuda_ms(c4, 10)
    out["c4_moleculenet_train_step"] = {"workload": "C4: ADKF_Meta_Trainer.train_step, 9 Tox21-like tasks x (20 support + 32 query), GIN 5 x 300, "
                                                    "AdamW, clip 1.0, H2D + loss read included", "value": T * 1e3 / ms, "unit": UNIT, "ms_per_step": ms}
    if ref is not None:
        run = ref.c4_setup(n_s, n_q)
        run(0)
        ts = sorted(run(1 + i) for i in range(3))
        out["c4_moleculenet_train_step"]["cpu_baseline"] = {"value": 1.0 / ts[1], "unit": UNIT, "cores": torch.get_num_threads(), "kind": "reference",
                                                            "sample": "median of 3 tasks: per-task body of the reference's train_step (fit + cauchy_hypergradient), "
                                                                      "torch_geometric restated (oracle/torch_geometric_lite.py)"}
    del trainer, mmodel
    # ---- C5: bayes_opt on a dockstring-shaped pool: fit on 64, joint NLL over the 2248 others (blocked Cholesky), EI over the pool ---
    x_all, y_all = synthetic.bo_pool(seed=5, pool=2312, dim=2048)
    X, Y = torch.from_numpy(x_all).to(device), torch.from_numpy(y_all).to(device)
    rng = np.random.RandomState(0)
    tr = np.sort(rng.choice(np.arange(2312), size=64, replace=False))
    te = np.array([i for i in range(2312) if i not in set(tr.tolist())])
    trd, ted = torch.from_numpy(tr).to(device), torch.from_numpy(te).to(device)
    y_tr = (Y[trd] - Y[trd].mean()) / Y[trd].std()
    y_te = (Y[ted] - Y[trd].mean()) / Y[trd].std()
    state = {}

    def c5_fit_nll():
        likelihood, gp, mll = bo_utils.create_gp(X[trd], y_tr, "matern", device, 0.01, True)
        gp.train(); likelihood.train()
        fit_gpytorch_scipy(mll)
        gp.eval(); likelihood.eval()
        state["m"] = (likelihood, gp)
        float((-likelihood(gp(X[ted])).log_prob(y_te) / len(te)).item())

    def c5_ei():
        likelihood, gp = state["m"]
        float(bo_utils.expected_improvement_pool(gp, likelihood, X, float(y_tr.min().item()), tr.tolist()).max().item())

    ms_fit = _cuda_ms(c5_fit_nll, 5)
    ms_ei = _cuda_ms(c5_ei, 10)
    # the blocked joint Cholesky alone (m^3/3 flops, trailing updates on the tensor cores)
    N = 2312
    layout = ops.TaskLayout.build([64], [N - 64], device)
    perm = torch.cat([trd, ted])
    sq = ops.gp_sqdist(X[perm].contiguous(), layout, qq_block=True)
    yy = torch.cat([y_tr, y_te]).contiguous()
    phi = state["m"][1]._phi()[0].contiguous()
    ms_chol = _cuda_ms(lambda: ops.gp_joint_nll_blocked(sq.D2, N, 64, yy, "matern", phi), 10)
    out["c5_bayes_opt"] = {"workload": "C5: dockstring-shaped pool 2312 x 2048: create_gp + fit on 64 + joint predictive NLL over 2248 molecules; "
                                       "expected improvement over the whole pool", "fit_plus_joint_nll_ms": ms_fit, "ei_pool_ms": ms_ei,
                           "value": 1e3 / (ms_fit + ms_ei), "unit": "pool evaluations/s (fit + NLL + EI sweep)",
                           "joint_nll_blocked": {"ms": ms_chol, "flops": N ** 3 / 3.0, "tflops": N ** 3 / 3.0 / (ms_chol * 1e-3) / 1e12,
                                                 "bound": "panel factorisations on the FP32 pipe, trailing updates on the tensor pipe"}}
    if ref is not None:
        run = ref.c5_setup()
        run(0, 50)
        a = [run(1 + i, 300) for i in range(3)]
        fit_s = sorted(x[0] for x in a)[1]
        ei_s = sorted(x[1] / x[2] for x in a)[1] * 2312
        out["c5_bayes_opt"]["cpu_baseline"] = {"value": 1.0 / (fit_s + ei_s), "unit": "pool evaluations/s (fit + NLL + EI sweep)", "cores": torch.get_num_threads(),
                                               "kind": "reference", "fit_plus_joint_nll_s": fit_s, "ei_pool_s": ei_s,
                                               "sample": "median of 3: the reference's create_gp + fit + joint log_prob; its one-point-at-a-time EI timed on 300 "
                                                         "pool molecules and scaled to 2312"}
    return out


# ----------------------------------------------------------------------------------------------------------
# CPU reference arm
#   kind "reference": the reference's OWN files (ADKTModelTrainer, run_on_batches, functional_call, cauchy_hypergradient),
#                     unmodified, from /root/reference or its staged copy oracle/_ref (oracle/build_ref.py), with the
#                     third-party libraries they import restated by oracle/shims.py where not installed;
#   kind "port"     : oracle/adkf_oracle.py (the same algorithm restated) when no copy of the reference is available.
# ----------------------------------------------------------------------------------------------------------
def reference_setup():
    """-> (run_task(task_id) -> (seconds, F), kind, description)."""
    torch.set_num_threads(os.cpu_count() or 1)
    from adkf_ift_b200 import synthetic
    try:
        from oracle import ref_runner as rr
        report = rr.install()
    except Exception:                                       # no reference copy on this machine: the restated algorithm
        return oracle_setup(), "port", "oracle/adkf_oracle.py (reference algorithm restated on torch CPU + SciPy L-BFGS-B)"
    model = rr.build_reference_model(rr.reference_config(num_layers=10, gp_kernel="matern"))

    def run_task(task_id):
        s = synthetic.make_task_sample(task_id, SUPPORT, QUERY)
        _, dev = rr.reference_dkt_sample(s, False, max_num_graphs=SUPPORT + QUERY)
        t0 = time.perf_counter()
        out = rr.reference_meta_train_task(model, dev.batches[0])        # adaptive_dkt_utils.py:361-404 for one task
        return time.perf_counter() - t0, float(out["F"])

    third = ", ".join(f"{k}: {v.split(' ')[0]}" for k, v in sorted(report.items()))
    return run_task, "reference", ("the reference's own fs_mol files, unmodified (run_on_batches + fit_gpytorch_scipy + functional_call + "
                                   f"cauchy_hypergradient per task); third-party libraries: {third}")


def oracle_setup():
    from adkf_ift_b200 import dkt, synthetic
    from adkf_ift_b200.adaptive_dkt import ADKTModelTrainerConfig
    from adkf_ift_b200.adaptive_dkt_utils import ADKTModelTrainer
    from adkf_ift_b200.feature_extractor import cli_default_config
    from oracle import adkf_oracle as orc
    torch.set_num_threads(os.cpu_count() or 1)
    model = ADKTModelTrainer(ADKTModelTrainerConfig(graph_feature_extractor_config=cli_default_config(10)))
    synthetic.deterministic_init_(model, seed=0, alpha=0.5)
    P = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters() if not n.startswith("gp_")}
    names = list(P.keys())
    ocfg = orc.OracleConfig(num_layers=10, gp_kernel="matern")

    def run_task(task_id):
        s = synthetic.make_task_sample(task_id, SUPPORT, QUERY)
        host = dkt.task_sample_to_dkt_task_sample(s, dkt.get_dkt_batcher(max_num_graphs=SUPPORT + QUERY), False)
        batch = dkt.torchify(host, torch.device("cpu")).batches[0]
        t0 = time.perf_counter()
        out = orc.meta_train_task(P, names, batch, ocfg)
        return time.perf_counter() - t0, float(out["F"])

    return run_task


def run_reference(args):
    """One "step" = a bounded sample of the C2 step: ONE of its 16 tasks through the reference's per-task body (the reference
    runs the tasks of a meta-batch sequentially, so tasks/s does not depend on how many are timed)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    run_task, kind, desc = reference_setup()
    for i in range(args.warmup):
        run_task(9000 + i)
    t = 0.0
    for i in range(args.steps):
        dt, _ = run_task(9100 + i)
        t += dt
    value = args.steps / t
    cores = torch.get_num_threads()
    sample = (f"{args.steps} steps x 1 task (support {SUPPORT}, query {QUERY}) of the C2 meta-training step, sequential as in "
              f"train_loop; {desc}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus, args.scaling),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


# ----------------------------------------------------------------------------------------------------------
def measure_tf32_peak(device):
    """Dense TF32 tensor-core peak of THIS GPU, the way MEASURED_PEAKS.json measures bf16: torch.matmul (cuBLAS, allow_tf32)
    on 8192^3, best of 10 with CUDA events (burst)."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        a, b = torch.randn(n, n, device=device), torch.randn(n, n, device=device)
        for _ in range(3):
            a @ b
        torch.cuda.synchronize()
        best = 1e30
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def run_b200(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=device)
    from adkf_ift_b200 import _cabi, ops, synthetic, task_batch
    from adkf_ift_b200.adaptive_dkt_utils import _LossReadback
    from adkf_ift_b200.eval_utils import DataFold
    from adkf_ift_b200.host_pipeline import HostBatchPipeline, StreamSpec
    lib = _cabi.load()
    model = make_model(device)
    per_gpu = tasks_per_gpu(world, args.scaling)
    global_tasks = per_gpu * world

    # ---- end-to-end leg set up FIRST (its worker processes are spawned before the timed regions and stay idle-blocked on
    # their bounded queues while the resident leg runs): the user-facing path from FSMolTask objects in host memory --
    # stratified support/query draw, DKT batching, multi-task collation in `workers` background processes, pinned staging,
    # async H2D + CSR build one step ahead, train_step, loss / status read one step late.
    pool_tasks = 48                                             # 3 step-batches per pass over the pool; every pass re-draws
    ds = synthetic.SyntheticFSMolDataset(pool_tasks, 0, 0, molecules_per_task=SUPPORT + QUERY,
                                         first_task_id=20000 + 1000 * rank).materialize(DataFold.TRAIN)
    # featurising one task costs ~8 ms of one core: 16 tasks per step need 3-4 workers to stay ahead of a ~90 ms step, 2 tasks
    # per step (strong scaling at N = 8) need one -- more would only compete with the launch threads for the cores
    workers = args.workers if args.workers is not None else max(1, min(4, (os.cpu_count() or 4) // max(world, 1) - 1, (per_gpu + 3) // 4))
    spec = StreamSpec(dataset=ds, data_fold=int(DataFold.TRAIN), support_size=SUPPORT, query_size=QUERY,
                      max_num_graphs=SUPPORT + QUERY, use_numeric_labels=False, tasks_per_step=per_gpu, num_shards=1, repeat=True)
    pipe = HostBatchPipeline(spec, num_workers=workers, rank=0, world=1)       # every rank owns its task pool

    # ---- resident leg: 2 pre-collated batches already in HBM (the CSR build stays inside the step) ---------------------
    n_distinct = 2
    hbs = [task_batch.collate_tasks(host_tasks(1000 * (rank + 1) + 100 * j, per_gpu), False) for j in range(n_distinct)]
    staged = [task_batch.to_device(hb, device, build_index=False) for hb in hbs]
    statuses = []

    def step_resident(i):
        tb = staged[i % n_distinct]
        tb.graph_index = ops.build_graph_index(tb.adjacency_lists, tb.num_nodes, tb.node_to_graph, tb.num_graphs)
        res = model.train_step(tb, global_num_tasks=global_tasks)
        statuses.append(res.status)

    prefetch = task_batch.DevicePrefetcher(pipe, device)
    readback = _LossReadback(device)
    records, h2d_bytes = [], []

    def step_e2e(i):
        tb = next(prefetch)
        res = model.train_step(tb, global_num_tasks=global_tasks)
        records.extend(readback.push(res))                     # D2H of step i-1's loss + status (12 bytes), one step late
        if i == args.steps - 1:
            records.extend(readback.flush())                   # ... and the last step's own, inside the timed region

    for i in range(args.warmup):
        step_resident(i)
    torch.cuda.synchronize()
    statuses.clear()
    l0 = lib.adkf_launch_count()
    with ClockSampler(local) as clk:
        ms = timed_region(step_resident, args.steps, world, device)
    launches = lib.adkf_launch_count() - l0
    bad = int(sum(int(((st & (ops.ST_NOT_PD | ops.ST_HESS_DAMPED)) != 0).sum().item()) for st in statuses))
    assert bad == 0, f"{bad} task(s) of the timed steps failed (status NOT_PD / HESS_DAMPED): the number would be invalid"
    for i in range(max(3, args.warmup)):
        step_e2e(-1)
    records.extend(readback.flush())
    torch.cuda.synchronize()
    records.clear()
    with ClockSampler(local) as clk2:
        ms_e2e = timed_region(step_e2e, args.steps, world, device)
    pipe.close()
    assert len(records) == args.steps and all(r[1] == 0 for r in records), f"e2e: dropped tasks in {records}"
    total_tasks = global_tasks * args.steps
    value = total_tasks / (ms * 1e-3)
    e2e = total_tasks / (ms_e2e * 1e-3)
    if rank != 0:
        if world > 1:
            torch.distributed.destroy_process_group()
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(world, args.scaling),
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": hbs[0].nbytes(), "d2h_bytes_per_step": 12,
                "ms_per_step": ms_e2e / args.steps, "host_workers_per_rank": workers, "host_cores": os.cpu_count(),
                "pipeline": "ADKTModelTrainer's own input path: FSMolTask objects in host memory -> stratified support/query draw "
                            "+ DKT batching + multi-task collation in background worker processes (every step a NEW draw: no two "
                            "batches alike) -> pinned ring -> async H2D + CSR build one step ahead -> train_step -> loss and "
                            "per-task status read by the host one step late; all of it inside the timed region"},
        "gpu_launches": int(launches), "clocks": clk.summary(), "clocks_e2e": clk2.summary(),
        "last_loss_per_sample": records[-1][0],
    }
    if world == 1:
        staged[0].graph_index = ops.build_graph_index(staged[0].adjacency_lists, staged[0].num_nodes, staged[0].node_to_graph,
                                                      staged[0].num_graphs)
        line["roofline"] = dominant_kernel_roofline(model, staged[0], peaks, measure_tf32_peak(device))
    if world == 1 and not args.no_cpu_baseline:
        run_task, kind, desc = reference_setup()
        run_task(9000)  # warm-up (thread pools, allocator)
        n_cpu = args.cpu_tasks
        t = sum(run_task(9100 + i)[0] for i in range(n_cpu))
        line["cpu_baseline"] = {"value": n_cpu / t, "unit": UNIT, "cores": torch.get_num_threads(), "kind": kind,
                                "sample": f"{n_cpu} tasks (support {SUPPORT}, query {QUERY}) of the same C2 step, sequential; {desc}"}
    if world == 1 and not args.no_meta_test:
        del model, staged
        torch.cuda.empty_cache()
        line["meta_test"] = meta_test_throughput(device, 0 if args.no_cpu_baseline else 3)
        if not args.no_other_configs:
            line["other_configs"] = other_configs(device, cpu=not args.no_cpu_baseline)
    print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-tasks", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-meta-test", action="store_true", help="skip the C3 meta-test throughput reported beside the headline")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the C1 / C4 / C5 figures reported beside the headline")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 16 tasks per GPU per step; strong: a global meta-batch of 16 tasks split over the GPUs")
    ap.add_argument("--workers", type=int, default=None, help="host featurisation processes per rank of the e2e leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
