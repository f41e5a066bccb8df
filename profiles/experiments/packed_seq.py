import time, torch
import hcatgnet_b200 as pkg
from hcatgnet_b200.synth import synth_batch
dev = torch.device("cuda:0")
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
packed = []
for i in range(4):
    pb = synth_batch(16384, 25, seed=20232023 + i, features="featuriser")
    packed.append(pkg.PackedDataset.from_arrays(pb.x.numpy(), pb.edge_index.numpy(), pb.ptr.numpy(), pb.y.numpy(), pinned=True))
def run(n):
    for _ in pkg.predict_stream(model, (packed[i % 4] for i in range(n))):
        pass
for n in (5, 20, 20, 20, 60):
    torch.cuda.synchronize(); t0 = time.perf_counter(); run(n); torch.cuda.synchronize()
    print(n, "steps: ms/step %.3f" % ((time.perf_counter() - t0) / n * 1e3), "reserved MiB", torch.cuda.memory_reserved() >> 20, flush=True)
