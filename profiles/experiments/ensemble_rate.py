"""One-pass ensemble forward vs M single-model edge-list forwards on the screening batch (16384 graphs):
device time per batch (CUDA events, 4 distinct batches cycled), per-model cost relative to a stand-alone forward."""
import json, sys
import torch
import hcatgnet_b200 as pkg
from hcatgnet_b200 import ops, _lib
from hcatgnet_b200.synth import synth_batch

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
pool = [synth_batch(B, 25, seed=20232023 + i).to(dev) for i in range(4)]
res = {"batch": B, "rows": []}

def timed(fn, n=12, warm=3):
    for i in range(warm): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for M in (1, 2, 4, 8, 16, 32, 90):
    models = []
    for k in range(M):
        torch.manual_seed(k)
        models.append(pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval())
    dims = models[0]._dims()
    params = [[p.detach() for p in m._flat_params()] for m in models]
    w = ops.EnsembleWeights(dims, params, dev)
    def one_pass(i):
        b = pool[i % 4]
        ops.forward_edge_list_ensemble(b.x, b.edge_index, b.ptr, b.edge_ptr, w, validate=False)
    def per_model(i):
        b = pool[i % 4]
        for ps in params:
            ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, ps, validate=False, )
    n = 12 if M <= 16 else 4
    t1 = timed(one_pass, n)
    tm = timed(per_model, n)
    row = {"models": M, "one_pass_ms": t1, "per_model_launches_ms": tm, "ratio": t1 / tm,
           "graph_model_evals_per_s": M * B / (t1 * 1e-3)}
    print(row, flush=True)
    res["rows"].append(row)
json.dump(res, open("gpurun_out/ensemble_rate.json", "w"), indent=1)
