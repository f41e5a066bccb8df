"""Reference-sized regime (668 graphs, batch 40 - BASELINE configs[0]): wall time per training / evaluation epoch through
FusedTrainer over a device-resident store versus host loaders.  Synthetic graphs of RhCAA size; prints one JSON line."""
import json
import random
import sys
import time

import torch

sys.path.insert(0, ".")
import hcatgnet_b200 as pkg
from hcatgnet_b200.synth import synth_batch

b = synth_batch(668, 25, seed=3, n_min=60, n_max=127, features="featuriser")
ds = pkg.PackedDataset.from_arrays(b.x.numpy(), b.edge_index.numpy(), b.ptr.numpy(), b.y.numpy())
ds.fold[:] = torch.arange(668).numpy() % 10
store = pkg.DeviceGraphStore(ds, "cuda")
opt = pkg.default_options()
tr_l, va_l, te_l = next(iter(pkg.create_loaders(store, opt)))
host_graphs = []
hb = ds.unpack_host()
for g in range(ds.G):
    lo, hi = int(hb.ptr[g]), int(hb.ptr[g + 1])
    sel = (hb.edge_index[0] >= lo) & (hb.edge_index[0] < hi)
    host_graphs.append({"x": hb.x[lo:hi], "edge_index": hb.edge_index[:, sel] - lo, "y": hb.y[g:g + 1], "idx": g, "fold": g % 10})
htr, hva, hte = next(iter(pkg.create_loaders(host_graphs, opt)))
res = {}
for name, (trl, val) in {"store": (tr_l, va_l), "host": (htr, hva)}.items():
    torch.manual_seed(0)
    model = pkg.make_network("GCN", opt, 25).cuda()
    t = pkg.FusedTrainer(model)
    random.seed(1)
    for _ in range(3):
        t.train_network(trl); t.eval_network(val)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        t.train_network(trl)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    for _ in range(20):
        t.eval_network(val)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    res[name] = {"train_epoch_ms": (t1 - t0) / 20 * 1e3, "train_steps_per_epoch": len(trl),
                 "eval_epoch_ms": (t2 - t1) / 20 * 1e3, "eval_steps_per_epoch": len(val)}
print(json.dumps(res))
