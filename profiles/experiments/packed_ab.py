"""Why is predict_stream over packed chunks slow after a device-resident edge-list loop?  Replays bench.py's sequence."""
import time, sys, os
import torch
import hcatgnet_b200 as pkg
from hcatgnet_b200 import _lib, ops
from hcatgnet_b200.synth import synth_batch

dev = torch.device("cuda:0")
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
dims = model._dims(); params = [p.detach() for p in model._flat_params()]
host = [synth_batch(16384, 25, seed=20232023 + i, pin=True) for i in range(4)]
packed = []
for i in range(4):
    pb = synth_batch(16384, 25, seed=20232023 + i, features="featuriser")
    packed.append(pkg.PackedDataset.from_arrays(pb.x.numpy(), pb.edge_index.numpy(), pb.ptr.numpy(), pb.y.numpy(), pinned=True))

def run_packed(n):
    for _ in pkg.predict_stream(model, (packed[i % 4] for i in range(n))):
        pass
def run_host(n):
    for _ in pkg.predict_stream(model, (host[i % 4] for i in range(n))):
        pass
def t(label, fn, n=20):
    fn(5); torch.cuda.synchronize()
    t0 = time.perf_counter(); fn(n); torch.cuda.synchronize(); dt = time.perf_counter() - t0
    print(label, "graphs/s %.3e  ms/step %.3f" % (n * 16384 / dt, dt / n * 1e3), "mem", torch.cuda.memory_reserved() >> 20, "MiB", flush=True)

t("packed (fresh)", run_packed)
devb = [h.__class__(h.x.to(dev), h.edge_index.to(dev), h.batch.to(dev), h.ptr.to(dev), h.y.to(dev), None, h.num_graphs, edges_grouped=True) for h in host]
eptr = [h.edge_ptr.to(dev) for h in host]
def run_dev(n):
    for i in range(n):
        b = devb[i % 4]
        ops.forward_edge_list(b.x, b.edge_index, b.ptr, eptr[i % 4], dims, params, validate=False)
t("device edge-list loop", run_dev)
t("packed (after device loop)", run_packed)
t("host e2e", run_host)
t("packed (after host e2e)", run_packed)
_lib.profile_enable(True)
run_packed(8); torch.cuda.synchronize()
print(_lib.profile_read())
