#!/usr/bin/env python
"""Benchmark of the GCN hot path: reaction graphs/s on synthetic screening batches.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--mode infer|train] [--batch B]
    python bench.py --impl reference ...      # the CPU restatement of the reference's PyG path

With no --mode the line's headline is BASELINE.json configs[1] (screening inference) and it carries three
sub-records measured in the same run: `train` (configs[2]: data-parallel training, batch 4096 per GPU,
FusedTrainer.step with the NCCL all-reduce of the gradient slab, `allreduce_us` timed on its own),
`stress` (configs[4]: H = 256, 4 conv layers, the layer-wise kernels with their `aggregation_roofline`) and
`ensemble` (16 models over the screening batch in one pass vs 16 single-model forwards).

A step = one pass of the hot path over one synthetic batch delivered as PyG would deliver it
(x fp32 [N,25], edge_index int64 [2,E], ptr / edge offsets int64 [B+1]).  Inference: hcg_gcn_forward_coo (the conv
kernel builds the normalised neighbour lists from the edge list itself; 2 x GCNConv + LeakyReLU, max||mean pooling,
readout MLP) - HCG_BENCH_CSR=1 selects the earlier hcg_collate (dst-sorted CSR, deg^-1/2, graph offsets) +
hcg_gcn_forward pair.  --mode train: hcg_collate + forward + hcg_gcn_backward, the NCCL all-reduce of the flat
gradient (N > 1) and the Adam step.  Default workload = BASELINE.json configs[1]: screening inference, ~80-atom
graphs, batch 16384 per GPU, 62 steps ~ 1 M graphs.

`value` times the device-resident path with CUDA events (barrier + synchronize on both sides,
max over ranks); `e2e` times the public nn.Module call with pinned HOST buffers, host->device
copies and the device->host read of the predictions inside the timed region.  Prints ONE JSON
line on rank 0.
"""
from __future__ import annotations

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

METRIC = "reaction_graphs_per_sec"
UNIT = "graphs/s"
F_IN, HID, N_CONV, N_RO, N_CLS = 25, 64, 2, 2, 1
POOL = 4  # distinct synthetic batches cycled through (4 x ~205 MB of inputs >> 126 MB L2)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=62)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"])
    ap.add_argument("--batch", type=int, default=None, help="graphs per GPU per step (16384 infer, 4096 train)")
    ap.add_argument("--hidden", type=int, default=HID)
    ap.add_argument("--convs", type=int, default=N_CONV)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="skip the `train` and `stress` sub-records of the default line")
    ap.add_argument("--sub-steps", type=int, default=20, help="timed steps of the `train` sub-record (`stress`: a quarter)")
    a = ap.parse_args()
    if a.batch is None:
        a.batch = 16384 if a.mode == "infer" else 4096
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    return a


def workload_name(a):
    kind = "screening inference" if a.mode == "infer" else "data-parallel training (fwd+bwd+allreduce+Adam)"
    return (f"synthetic reaction graphs n~U[40,120] e=3n F={F_IN} H={a.hidden} L={a.convs}; {kind}; "
            f"batch {a.batch}/GPU")


def config_of(a):
    """Identical in both arms (`--impl b200` / `--impl reference`): the workload, nothing measured."""
    return {"workload": workload_name(a), "mode": a.mode, "batch_per_gpu": a.batch, "hidden": a.hidden,
            "n_convolutions": a.convs, "n_node_features": F_IN, "graphs_per_gpu_in_timed_region": a.batch * a.steps,
            "l2_policy": f"inputs larger than L2: {POOL} distinct batches cycled", "parallelism": f"dp{a.gpus} (graphs sharded)"}


# ----------------------------------------------------------------------------------------------
# clocks sampling (nvidia-smi in a side process during the timed region)
# ----------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle (plain-torch restatement of the reference's PyG path) on the host cores
# ----------------------------------------------------------------------------------------------
def host_cores() -> int:
    """Host threads actually usable here (cgroup / affinity aware), not what OMP_NUM_THREADS says:
    torchrun exports OMP_NUM_THREADS=1 to its workers."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except (AttributeError, OSError):
        return os.cpu_count() or 1


def cpu_model() -> str:
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def _oracle_step_fn(hidden, convs, mode, b):
    from oracle import ref_gcn
    sd = ref_gcn.init_state_dict(F_IN, hidden, convs, N_RO, N_CLS)
    if mode == "train":
        sd = {k: v.requires_grad_(True) for k, v in sd.items()}
        opt = torch.optim.Adam(list(sd.values()), lr=0.01, eps=1e-9)

        def one():
            opt.zero_grad()
            out, _ = ref_gcn.forward(sd, b.x, b.edge_index, b.batch, b.num_graphs)
            ref_gcn.rmse_loss(out, b.y).backward()
            opt.step()
    else:
        def one():
            with torch.no_grad():
                ref_gcn.forward(sd, b.x, b.edge_index, b.batch, b.num_graphs)
    return one


def _time_fn(fn, budget_s, min_iters=2, max_iters=50):
    fn()
    t0 = time.perf_counter(); fn(); dt1 = time.perf_counter() - t0
    iters = int(max(min_iters, min(max_iters, budget_s / max(dt1, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t0) / iters, iters


def cpu_oracle_rate(hidden, convs, sample_graphs: int, budget_s: float, mode: str):
    """graphs/s of oracle/ref_gcn on `sample_graphs` synthetic graphs, all host threads."""
    from hcatgnet_b200.synth import synth_batch
    cores = host_cores()
    torch.set_num_threads(cores)
    b = synth_batch(sample_graphs, F_IN, seed=20232023)
    dt, iters = _time_fn(_oracle_step_fn(hidden, convs, mode, b), budget_s)
    return sample_graphs / dt, cores, (f"{iters} x {sample_graphs} graphs of the same generator (collated batch in RAM), "
                                       f"torch {torch.__version__}, {cores} threads, {cpu_model()}")


def python_collate_rate(num_graphs: int):
    """graphs/s of the Python collate PyG's DataLoader does per batch (list of per-graph records -> batch):
    timed on `num_graphs` graphs split out of one synthetic batch."""
    from hcatgnet_b200.data import Batch
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(num_graphs, F_IN, seed=20232023)
    ptr = b.ptr.tolist()
    # per-graph records (edge lists are grouped by graph in the generator's output)
    gid = torch.bucketize(b.edge_index[0], b.ptr[1:], right=True)
    counts = torch.bincount(gid, minlength=num_graphs).tolist()
    graphs, e0 = [], 0
    for g in range(num_graphs):
        n0, n1, e1 = ptr[g], ptr[g + 1], e0 + counts[g]
        graphs.append({"x": b.x[n0:n1], "edge_index": b.edge_index[:, e0:e1] - n0, "y": b.y[g:g + 1], "idx": g})
        e0 = e1
    t0 = time.perf_counter()
    Batch.from_data_list(graphs)
    return num_graphs / (time.perf_counter() - t0)


def run_reference(a):
    """The reference's CPU path for this workload: the oracle port on all host threads; every step is one
    full `a.batch`-graph batch of the same generator (the same `config` as the B200 arm)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from hcatgnet_b200.synth import synth_batch
    cores = host_cores()
    torch.set_num_threads(cores)
    pool = [synth_batch(a.batch, F_IN, seed=20232023 + i) for i in range(2)]
    fns = [_oracle_step_fn(a.hidden, a.convs, a.mode, b) for b in pool]
    for i in range(a.warmup):
        fns[i % 2]()
    t0 = time.perf_counter()
    for i in range(a.steps):
        fns[i % 2]()
    dt = time.perf_counter() - t0
    v = a.batch * a.steps / dt
    # also one thread, and the Python collate PyG performs per batch
    torch.set_num_threads(1)
    dt1, it1 = _time_fn(_oracle_step_fn(a.hidden, a.convs, a.mode, synth_batch(1024, F_IN, seed=20232023)), 6.0, 1, 8)
    torch.set_num_threads(cores)
    collate = python_collate_rate(2048)
    sample_txt = (f"each step = one full {a.batch}-graph batch, oracle/ref_gcn.py (plain-torch restatement of the PyG path; "
                  f"torch_geometric not installable), {cores} threads, {cpu_model()}, torch {torch.__version__}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": dt / a.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_of(a),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_txt,
                         "one_thread": {"value": 1024 / dt1, "unit": UNIT, "sample": f"{it1} x 1024 graphs, 1 thread"},
                         "python_collate": {"value": collate, "unit": UNIT,
                                            "sample": "Batch.from_data_list over 2048 per-graph records (what PyG's DataLoader does per batch), 1 thread"},
                         "cpu_model": cpu_model()},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------------------
# algorithmic bytes per launch of each kernel (SURVEY.md §8d formulas)
# ----------------------------------------------------------------------------------------------
def algorithmic_bytes(name: str, N: int, E: int, B: int, F: int, H: int):
    """Bytes one launch of `name` must move (each tensor once)."""
    csr = 4 * (N + 1) + 4 * E
    table = {
        # whole conv stack + pooling of a batch: x, CSR, deg^-1/2 in; embedding + per-graph descriptor out/in
        "k_tc_fwd": 4 * N * F + csr + 4 * N + 8 * B * H + 64 * B,   # (+ one 64-byte work-list record per tile, <= B tiles)
        # the same from the raw edge list: x, int64 (src, dst) per edge, work-list record per tile in; embedding out
        "k_tc_fwd_coo": 4 * N * F + 16 * E + 8 * B * H + 64 * B,
        "k_fused_fwd": 4 * N * F + csr + 4 * N + 8 * B * H + 4 * (B + 1),
        # graph-resident backward: saved h1, h2, x, transposed CSR, dis, emb + its gradient in; per-CTA partials out
        "k_fused_bwd": 8 * N * H + 4 * N * F + csr + 4 * N + 16 * B * H + 296 * 4 * (H * H + H * F + 2 * H),
        "k_tc_bwd": 8 * N * H + 4 * N * F + csr + 4 * N + 16 * B * H + 64 * B + 148 * 4 * (H * H + H * F + 2 * H),
        "k_agg": 4 * N * H + csr + 4 * N + 4 * N * H,        # read T once, CSR, dis, write h
        "k_agg_wide": 4 * N * H + csr + 4 * N + 4 * N * H,   # same contract as k_agg
        "k_agg_narrow": 4 * N * H + csr + 4 * N + 4 * N * H,
        "k_agg_wide_pool": 4 * N * H + csr + 4 * N + 4 * (B + 1) + 8 * B * H,   # inference: rows are pooled, not stored
        "k_pool": 4 * N * H + 4 * (B + 1) + 8 * B * H,
        "k_count": 16 * E,                                     # + integer atomics into cnt (L2-resident)
        "k_fill": 16 * E + 4 * (N + 1) + 8 * E,
        "k_rowsort": 4 * (N + 1) + 8 * E + 8 * E,
    }
    return table.get(name)


def algorithmic_flops(name: str, N: int, F: int, H: int, L: int):
    """FLOPs one launch of a tensor-bound kernel must perform (fp32-equivalent: 2 M K Nc), averaged over
    the launches of a step."""
    if name == "k_tc_gemm":  # forward: layer 0 has K = F, the others K = H (inference: one launch per layer)
        return (2.0 * N * F * H + (L - 1) * 2.0 * N * H * H) / L
    return None


def measured_traffic(name: str, mode: str, batch: int, hidden: int, convs: int):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full`
    capture of this kernel on this workload (profiles/ncu_traffic.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        return None
    rec = t.get(f"{name}|{mode}|batch{batch}|H{hidden}|L{convs}")
    return None if rec is None else rec["dram_bytes"]


# ----------------------------------------------------------------------------------------------
def measure_ensemble(cx, batch: int, n_models: int, steps: int):
    """`n_models` networks over the screening batch in ONE pass (hcg_gcn_forward_coo_ensemble)
    against the same models as `n_models` single-model edge-list forwards; device-resident, CUDA events, 4 batches cycled.
    `value` counts graph x model evaluations per second over all ranks."""
    import hcatgnet_b200 as pkg
    from hcatgnet_b200 import _lib, ops
    from hcatgnet_b200.synth import synth_batch
    dev = cx.dev
    pool = [synth_batch(batch, F_IN, seed=20232023 + 1000 * cx.rank + i).to(dev) for i in range(POOL)]
    models = []
    for k in range(n_models):
        torch.manual_seed(100 + k)
        models.append(pkg.make_network("GCN", pkg.default_options(embedding_dim=HID, n_convolutions=N_CONV), F_IN).to(dev).eval())
    dims = models[0]._dims()
    params = [[p.detach() for p in m._flat_params()] for m in models]
    w = ops.EnsembleWeights(dims, params, dev)

    def one_pass(i):
        b = pool[i % POOL]
        return ops.forward_edge_list_ensemble(b.x, b.edge_index, b.ptr, b.edge_ptr, w, validate=False)

    def per_model(i):
        b = pool[i % POOL]
        for ps in params:
            ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, ps, validate=False)

    for i in range(3):
        one_pass(i); per_model(i)
    l0 = _lib.launch_count()
    ms1 = cx.timed(one_pass, steps)
    launches = int(_lib.launch_count() - l0)
    msm = cx.timed(per_model, steps)
    del pool
    return {"models": n_models, "batch_per_gpu": batch, "steps": steps,
            "value": n_models * batch * steps * cx.world / (ms1 * 1e-3), "unit": "graph-model evaluations/s",
            "ms_per_step": ms1 / steps, "ms_per_step_per_model_launches": msm / steps,
            "per_model_cost_vs_single_forward": (ms1 / steps) / (msm / steps), "gpu_launches": launches,
            "api": "hcg_gcn_forward_coo_ensemble (one pass) vs n_models x hcg_gcn_forward_coo"}


# ----------------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of one bench run."""

    def __init__(self):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.peaks = {}
        try:
            self.peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except (OSError, ValueError):
            pass

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps, whole=False):
        """Device time of `steps` calls (CUDA events on the launch stream), barrier + synchronize on both
        sides, max over ranks."""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        ev0.record()
        if whole:
            fn(steps)
        else:
            for i in range(steps):
                fn(i)
        ev1.record()
        self.barrier()
        ms = ev0.elapsed_time(ev1)
        if self.dist is not None:
            t = torch.tensor([ms], device=self.dev)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms


def measure(cx: Ctx, mode: str, batch: int, hidden: int, convs: int, steps: int, warmup: int, host=None,
            want_e2e=True, want_packed=True, want_profile=True, want_clocks=True):
    """One configuration: device-resident value, end-to-end value(s), per-kernel device time and rooflines.
    Returns (record dict, host batches) — the pinned host batches can be handed to another configuration."""
    import hcatgnet_b200 as pkg
    from hcatgnet_b200 import _lib, ops
    from hcatgnet_b200.synth import synth_batch

    dev, rank, world, dist = cx.dev, cx.rank, cx.world, cx.dist
    opt = pkg.default_options(embedding_dim=hidden, n_convolutions=convs)
    model = pkg.make_network("GCN", opt, F_IN).to(dev)
    model.eval() if mode == "infer" else model.train()
    dims = model._dims()
    # training: the device-resident step (collate, forward, loss seed, backward, all-reduce, Adam)
    trainer = pkg.FusedTrainer(model) if mode == "train" else None
    params = [p.detach() for p in model._flat_params()]

    # synthetic pool: pinned host copies (e2e) and device-resident copies (value)
    if host is None:
        host = [synth_batch(batch, F_IN, seed=20232023 + 1000 * rank + i, pin=True) for i in range(POOL)]
    devb = []
    for b in host:
        d = pkg.Batch(b.x.to(dev), b.edge_index.to(dev), b.batch.to(dev), b.ptr.to(dev), b.y.to(dev), None, b.num_graphs,
                      edges_grouped=True)
        # what a PyG Batch knows on the host (ptr, _slice_dict): offsets of every graph and the size of the largest one
        d.max_graph_nodes, d.max_graph_edges, d.edge_ptr = b.max_graph_nodes, b.max_graph_edges, b.edge_ptr.to(dev)
        d.max_in_degree = b.max_in_degree
        devb.append(d)
    # inference straight from the edge list (hcg_gcn_forward_coo: the conv kernel builds the neighbour lists) when the
    # batches' host-side size facts allow it; HCG_BENCH_CSR=1 forces hcg_collate + hcg_gcn_forward (the round-1/2 path)
    edge_list = mode == "infer" and os.environ.get("HCG_BENCH_CSR", "0") != "1" and all(ops.edge_list_hints_ok(d) for d in devb)
    Ns = [b.num_nodes for b in host]
    Es = [int(b.edge_index.shape[1]) for b in host]
    N, E = int(np.mean(Ns)), int(np.mean(Es))

    def step_device(i):
        b = devb[i % POOL]
        if edge_list:
            res = ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, params, validate=False)
            if res is not None:
                return res[0]
        if mode == "infer":
            plan = ops.collate(b.edge_index, b.batch, b.num_nodes, b.num_graphs, validate=False, grouped=True,
                               max_graph_nodes=b.max_graph_nodes, ptr=b.ptr, edge_ptr=b.edge_ptr,
                               max_graph_edges=b.max_graph_edges)
            with torch.no_grad():
                out, _ = ops.gcn_apply(b.x, plan, dims, params)
            return out
        # a fresh batch object every step: the collate runs inside the step (no topology cache hit)
        fresh = pkg.Batch(b.x, b.edge_index, b.batch, b.ptr, b.y, None, b.num_graphs, edges_grouped=True)
        fresh.max_graph_nodes, fresh.max_graph_edges, fresh.edge_ptr = b.max_graph_nodes, b.max_graph_edges, b.edge_ptr
        fresh.max_in_degree = b.max_in_degree
        return trainer.step(fresh)

    def drain_losses(stream_of_losses):
        prev = None
        for loss in stream_of_losses:
            l_h = torch.empty((), dtype=loss.dtype, pin_memory=True)
            l_h.copy_(loss, non_blocking=True)
            ev = torch.cuda.Event(); ev.record()
            if prev is not None:       # read the previous step's loss: one step of slack keeps the queue full
                prev[1].synchronize(); float(prev[0])
            prev = (l_h, ev)
        if prev is not None:
            prev[1].synchronize(); float(prev[0])

    def run_e2e(n):
        """The public streaming API on HOST batches: every step's inputs cross PCIe from pinned memory
        (staged one batch ahead on a copy stream) and every step's result is read back to the host."""
        stream = (host[i % POOL] for i in range(n))
        if mode == "infer":
            for _out, _, _ in pkg.predict_stream(model, stream):
                pass
        else:
            drain_losses(trainer.step(d) for d in pkg.DevicePrefetcher(stream, device=dev))

    rec = {"mode": mode, "path": "edge_list (hcg_gcn_forward_coo)" if edge_list else "collate + forward", "batch_per_gpu": batch, "hidden": hidden, "n_convolutions": convs, "steps": steps, "warmup": warmup,
           "mean_nodes_per_batch": N, "mean_edges_per_batch": E}

    # ---- device-resident number --------------------------------------------------------------
    for i in range(warmup):
        step_device(i)
    sampler = ClockSampler(cx.local)
    if rank == 0 and want_clocks:
        sampler.start()
    l0 = _lib.launch_count()
    ms = cx.timed(step_device, steps)
    rec["gpu_launches"] = int(_lib.launch_count() - l0)
    if rank == 0 and want_clocks:
        rec["clocks"] = sampler.stop()
    graphs_total = batch * steps * world
    rec["value"], rec["unit"], rec["ms_per_step"] = graphs_total / (ms * 1e-3), UNIT, ms / steps

    # ---- NCCL all-reduce of the gradient slab, timed on its own (events around the collective) ---
    if mode == "train":
        rec["allreduce_us"] = None
        if world > 1:
            trainer.timing = []
            for i in range(steps):
                step_device(i)
            torch.cuda.synchronize()
            us = [e0.elapsed_time(e1) * 1e3 for e0, e1 in trainer.timing]
            trainer.timing = None
            # median: a rank that enters this pass late (rank 0 has just read the clocks) makes the others wait once
            t = torch.tensor([float(np.median(us))], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            rec["allreduce_us"] = float(t.item())
            rec["allreduce_bytes"] = int(trainer.grad.numel() * 4)

    # ---- end to end through the nn.Module API with host buffers ------------------------------
    d2h = batch * N_CLS * 4 if mode == "infer" else 4
    if want_e2e:
        run_e2e(warmup)
        t0 = time.perf_counter()
        ms_e2e = cx.timed(run_e2e, steps, whole=True)
        wall_e2e = (time.perf_counter() - t0) * 1e3
        # what DevicePrefetcher copies per batch: x, edge_index, ptr (the batch vector is rebuilt on the device), y
        h2d = int(np.mean([h.x.numel() * 4 + h.edge_index.numel() * 8 + (h.ptr.numel() * 8 if h.ptr is not None else h.batch.numel() * 8)
                           + h.y.numel() * 4 for h in host]))
        rec["e2e"] = {"value": graphs_total / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "ms_per_step": ms_e2e / steps, "wall_ms_per_step": wall_e2e / steps,
                      "api": "hcatgnet_b200.predict_stream (H2D of batch i+1 overlaps the kernels of batch i)"
                      if mode == "infer" else "hcatgnet_b200.DevicePrefetcher + FusedTrainer.step"}

    # ---- the same, with the batches in the packed wire format ----------
    # Node features with the block structure of the reference featuriser (24 one-hot / binary columns -> one bit
    # each, temperature -> f32), endpoints local to their graph in u8: ~11x fewer bytes over PCIe, consumed as they
    # are by hcg_gcn_forward_packed (the conv kernel expands a tile's rows in shared memory).  Same graphs as `host`.
    if want_packed and mode == "infer":
        packed = []
        for i in range(POOL):
            pb = synth_batch(batch, F_IN, seed=20232023 + 1000 * rank + i, features="featuriser")
            packed.append(pkg.PackedDataset.from_arrays(pb.x.numpy(), pb.edge_index.numpy(), pb.ptr.numpy(), pb.y.numpy(),
                                                        pinned=True))

        def run_packed(n):
            for _out, _, _ in pkg.predict_stream(model, (packed[i % POOL] for i in range(n))):
                pass

        run_packed(warmup)
        ms_p = cx.timed(run_packed, steps, whole=True)
        rec["e2e_packed"] = {"value": graphs_total / (ms_p * 1e-3), "unit": UNIT, "ms_per_step": ms_p / steps,
                             "h2d_bytes_per_step": int(np.mean([p.nbytes for p in packed])), "d2h_bytes_per_step": d2h,
                             "api": "hcatgnet_b200.predict_stream over PackedDataset chunks (one image copy per step, "
                                    "hcg_gcn_forward_packed straight from the image: no unpack, no collate)",
                             "x": "featuriser-shaped: 24 one-hot/binary columns as one bit each + 1 f32 column (lossless)"}
        del packed
    elif want_packed:
        # training from a device-resident packed store: shuffled mini-batches are assembled on the GPU
        # (hcg_unpack_batch); per step only the id list crosses PCIe and the loss comes back
        pbs = [synth_batch(batch, F_IN, seed=20232023 + 1000 * rank + i, features="featuriser") for i in range(POOL)]
        xs = np.concatenate([p.x.numpy() for p in pbs]); ys = np.concatenate([p.y.numpy() for p in pbs])
        offs = np.cumsum([0] + [p.num_nodes for p in pbs])
        eis = np.concatenate([p.edge_index.numpy() + offs[i] for i, p in enumerate(pbs)], axis=1)
        ptrs = np.concatenate([pbs[0].ptr.numpy()] + [p.ptr.numpy()[1:] + offs[i] for i, p in enumerate(pbs) if i > 0])
        store = pkg.DeviceGraphStore(pkg.PackedDataset.from_arrays(xs, eis, ptrs, ys), dev)
        del pbs, xs, eis, ptrs
        rng = np.random.default_rng(7 + rank)

        def run_store(n):
            drain_losses(trainer.step(store.batch(rng.choice(store.ds.G, size=batch, replace=False), with_batch_vector=True))
                         for _ in range(n))

        run_store(warmup)
        ms_p = cx.timed(run_store, steps, whole=True)
        rec["e2e_packed"] = {"value": graphs_total / (ms_p * 1e-3), "unit": UNIT, "ms_per_step": ms_p / steps,
                             "h2d_bytes_per_step": 8 * batch, "d2h_bytes_per_step": 4,
                             "api": "hcatgnet_b200.DeviceGraphStore.batch(shuffled ids) + FusedTrainer.step: the packed data set "
                                    f"({store.ds.G} graphs, {store.ds.nbytes} bytes) is resident in HBM",
                             "x": "featuriser-shaped: 24 one-hot/binary columns as u8 + 1 f32 column (lossless)"}
        del store

    # ---- per-kernel device time (separate pass, same inputs) + roofline of the dominant kernel -
    rec["roofline"], rec["kernels"], rec["aggregation_roofline"] = None, None, None
    if want_profile:
        if rank != 0:
            if world > 1 and mode == "train":   # the training step holds a collective: every rank walks the profile pass
                for i in range(steps):
                    step_device(i)
            torch.cuda.synchronize()
        else:
            peaks = cx.peaks
            peak = float(peaks.get("hbm_gbs", 6650.0))
            peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
            torch.cuda.synchronize()
            _lib.profile_enable(True)
            for i in range(steps):
                step_device(i)
            torch.cuda.synchronize()
            prof = _lib.profile_read()
            _lib.profile_enable(False)
            tot = sum(v[0] for v in prof.values()) or 1.0
            kernels = {k: {"ms_per_launch": v[0] / v[1], "launches_per_step": v[1] / steps, "share": v[0] / tot}
                       for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
            rec["kernels"] = kernels
            top = next(iter(kernels))
            tname = top.split("<")[0]
            ab = algorithmic_bytes(tname, N, E, batch, F_IN, hidden)
            af = algorithmic_flops(tname, N, F_IN, hidden, convs) if mode == "infer" else None
            sec = kernels[top]["ms_per_launch"] * 1e-3
            if af is not None:
                # dense contraction on the tensor cores: achieved fp32-equivalent TFLOP/s against the measured dense
                # bf16 peak (sustained figure: the kernel is timed inside a long step).  The split-TF32 scheme
                # issues 3 TF32 MMAs per product and TF32 runs at half the bf16 rate, so frac <= 1/6 by construction.
                tpeak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1500.0)))
                ach = af / sec / 1e12
                rec["roofline"] = {"bound": "tensor", "kernel": top, "achieved": ach, "peak": tpeak, "unit": "TFLOP/s",
                                   "frac": ach / tpeak, "traffic": measured_traffic(tname, mode, batch, hidden, convs),
                                   "algorithmic_flops_per_launch": af, "ms_per_launch": kernels[top]["ms_per_launch"],
                                   "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (cuBLAS bf16)"
                                   if "bf16_tflops_sustained" in peaks else "fallback 1500 TFLOP/s",
                                   "share_of_step": kernels[top]["share"],
                                   "note": "fp32-equivalent FLOPs; 3 TF32 MMAs per product at half the bf16 rate: frac <= 0.167"}
            elif ab is not None:
                ach = ab / sec / 1e9
                rec["roofline"] = {"bound": "hbm", "kernel": top, "achieved": ach, "peak": peak, "unit": "GB/s",
                                   "frac": ach / peak, "traffic": measured_traffic(tname, mode, batch, hidden, convs),
                                   "algorithmic_bytes_per_launch": ab, "ms_per_launch": kernels[top]["ms_per_launch"],
                                   "peak_source": peak_src, "share_of_step": kernels[top]["share"]}
            # the fused GCN aggregation kernels of the layer-wise path against the HBM roofline (north_star's target
            # metric), whenever they run in this configuration (wide hidden layers)
            if mode == "infer":
                agg = {}
                for kname, v in kernels.items():
                    key = kname.split("<")[0]
                    if key not in ("k_agg_wide", "k_agg_wide_pool", "k_agg_narrow", "k_agg"):
                        continue
                    ab2 = algorithmic_bytes(key, N, E, batch, F_IN, hidden)
                    ach = ab2 / (v["ms_per_launch"] * 1e-3) / 1e9
                    agg[key] = {"kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                                "traffic": measured_traffic(key, mode, batch, hidden, convs),
                                "algorithmic_bytes_per_launch": ab2, "ms_per_launch": v["ms_per_launch"],
                                "launches_per_step": v["launches_per_step"]}
                rec["aggregation_roofline"] = agg or None
    del devb, model, trainer
    torch.cuda.empty_cache()
    return rec, host


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product has no CPU path (use --impl reference for the CPU arm)")
    cx = Ctx()
    from hcatgnet_b200 import _lib
    _lib.load()

    head, host = measure(cx, a.mode, a.batch, a.hidden, a.convs, a.steps, a.warmup, want_profile=not a.no_profile)

    # The other two configurations BASELINE.json names, as sub-records of the same line (the driver passes no flags):
    #   train  = configs[2]: data-parallel training, batch 4096 per GPU, NCCL all-reduce of the gradient slab
    #   stress = configs[4]: H = 256, 4 conv layers, screening inference, batch 16384 per GPU (the layer-wise kernels)
    train = stress = None
    default_cfg = a.mode == "infer" and a.hidden == HID and a.convs == N_CONV
    if default_cfg and not a.no_sub:
        sub = max(3, int(a.sub_steps))
        stress, _ = measure(cx, "infer", a.batch, 256, 4, max(3, sub // 4), 3, host=host, want_e2e=False, want_packed=False,
                            want_profile=not a.no_profile, want_clocks=False)
        del host
        train, _ = measure(cx, "train", 4096, HID, N_CONV, sub, max(3, a.warmup), want_packed=False,
                           want_profile=not a.no_profile, want_clocks=False)

    ensemble = None
    if default_cfg and not a.no_sub:
        ensemble = measure_ensemble(cx, a.batch, 16, max(3, int(a.sub_steps) // 4))

    cpu_baseline = None
    if cx.rank == 0 and cx.world == 1 and not a.no_cpu_baseline:
        v, cores, sample = cpu_oracle_rate(a.hidden, a.convs, 2048, 12.0, a.mode)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}

    if cx.rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": cx.world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_of(a),
            "workload_stats": {"mean_nodes_per_batch": head["mean_nodes_per_batch"], "mean_edges_per_batch": head["mean_edges_per_batch"]},
            "clocks": head.get("clocks"), "e2e": head.get("e2e"), "e2e_packed": head.get("e2e_packed"),
            "gpu_launches": head["gpu_launches"], "roofline": head["roofline"],
            "aggregation_roofline": head["aggregation_roofline"], "kernels": head["kernels"], "cpu_baseline": cpu_baseline,
            "train": train, "stress": stress, "ensemble": ensemble,
        }
        if a.mode == "train":
            line["allreduce_us"] = head.get("allreduce_us")
        print(json.dumps(line))
    if cx.dist is not None:
        cx.dist.barrier()
        cx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
