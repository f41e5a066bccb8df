"""Where does the host time of predict_stream over packed chunks go?  cProfile over 60 steps."""
import cProfile, pstats, time, io
import torch
import hcatgnet_b200 as pkg
from hcatgnet_b200.synth import synth_batch

dev = torch.device("cuda:0")
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev).eval()
packed = []
for i in range(4):
    pb = synth_batch(16384, 25, seed=20232023 + i, features="featuriser")
    packed.append(pkg.PackedDataset.from_arrays(pb.x.numpy(), pb.edge_index.numpy(), pb.ptr.numpy(), pb.y.numpy(), pinned=True))
print("bytes per chunk", packed[0].nbytes)

def run(n):
    for _ in pkg.predict_stream(model, (packed[i % 4] for i in range(n))):
        pass

run(10); torch.cuda.synchronize()
t0 = time.perf_counter(); run(60); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print("ms/step %.3f" % (dt / 60 * 1e3))
pr = cProfile.Profile(); pr.enable(); run(60); torch.cuda.synchronize(); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue()[:6000])
