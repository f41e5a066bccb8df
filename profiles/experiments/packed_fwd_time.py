import torch, time
import hcatgnet_b200 as pkg
from hcatgnet_b200 import ops, _lib
from hcatgnet_b200.synth import synth_batch
dev = torch.device("cuda", 0)
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
dims = model._dims(); params = [p.detach() for p in model._flat_params()]
stores = []
for i in range(4):
    pb = synth_batch(16384, 25, seed=20232023 + i, features="featuriser")
    stores.append(pkg.DeviceGraphStore(pkg.PackedDataset.from_arrays(pb.x.numpy(), pb.edge_index.numpy(), pb.ptr.numpy(), pb.y.numpy()), dev))
def timed(fn, n=20):
    for i in range(5): fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
def direct(i):
    s = stores[i % 4]
    ops.forward_packed(s.c_store, 0, s.ds.G, dims, params, dev, validate=False)
def unpack(i):
    s = stores[i % 4]
    b = s.batch(None, with_batch_vector=False)
    ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, params, validate=False)
print("packed direct ms", timed(direct)); print("unpack + edge list ms", timed(unpack))
_lib.profile_enable(True)
for i in range(8): direct(i)
torch.cuda.synchronize(); print(_lib.profile_read())
