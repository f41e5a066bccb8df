"""A/B of the edge-list forward between library builds on ONE box: HCG_LIB_PATH=... python coo_ab.py"""
import os, torch
from hcatgnet_b200 import _lib
for k in ("hcg_gcn_forward_packed", "hcg_forward_packed_supported"):
    if os.environ.get("HCG_AB_OLD"): _lib.SIGNATURES.pop(k, None)
import hcatgnet_b200 as pkg
from hcatgnet_b200 import ops
from hcatgnet_b200.synth import synth_batch
dev = torch.device("cuda", 0)
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
dims = model._dims(); params = [p.detach() for p in model._flat_params()]
pool = [synth_batch(16384, 25, seed=20232023 + i).to(dev) for i in range(4)]
def fn(i):
    b = pool[i % 4]
    ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, params, validate=False)
res = []
for rep in range(3):
    for i in range(5): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40): fn(i)
    e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) / 40)
print(os.environ.get("HCG_LIB_PATH", "default"), ["%.4f" % r for r in res])
_lib.profile_enable(True)
for i in range(8): fn(i)
torch.cuda.synchronize()
print({k: round(v[0] / v[1] * 1e3, 1) for k, v in _lib.profile_read().items()})
