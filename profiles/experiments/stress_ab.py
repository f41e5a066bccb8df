"""A/B of the layer-wise (stress configuration) forward between library builds on one box."""
import os, torch
import hcatgnet_b200 as pkg
from hcatgnet_b200 import ops, _lib
from hcatgnet_b200.synth import synth_batch
dev = torch.device("cuda", 0)
model = pkg.make_network("GCN", pkg.default_options(embedding_dim=256, n_convolutions=4), 25).to(dev).eval()
dims = model._dims(); params = [p.detach() for p in model._flat_params()]
pool = [synth_batch(16384, 25, seed=20232023 + i).to(dev) for i in range(2)]
plans = [ops.collate(b.edge_index, b.batch, b.num_nodes, b.num_graphs, validate=False, grouped=True, ptr=b.ptr, edge_ptr=b.edge_ptr) for b in pool]
def fn(i):
    with torch.no_grad():
        ops.gcn_apply(pool[i % 2].x, plans[i % 2], dims, params)
for i in range(3): fn(i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10): fn(i)
e1.record(); torch.cuda.synchronize()
_lib.profile_enable(True)
for i in range(4): fn(i)
torch.cuda.synchronize()
pr = _lib.profile_read()
print(os.environ.get("HCG_LIB_PATH", "default"), "ms/step %.3f" % (e0.elapsed_time(e1) / 10), {k: round(v[0] / v[1] * 1e3) for k, v in pr.items() if "agg" in k})
