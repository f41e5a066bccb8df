"""A/B of the training step between library builds on ONE box: HCG_LIB_PATH=... python train_ab.py"""
import os, torch
import hcatgnet_b200 as pkg
from hcatgnet_b200 import _lib
from hcatgnet_b200.synth import synth_batch
dev = torch.device("cuda", 0)
model = pkg.make_network("GCN", pkg.default_options(), 25).to(dev)
trainer = pkg.FusedTrainer(model)
pool = [synth_batch(4096, 25, seed=20232023 + i).to(dev) for i in range(4)]
res = []
for rep in range(3):
    for i in range(5): trainer.step(pool[i % 4])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(40): trainer.step(pool[i % 4])
    e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) / 40)
_lib.profile_enable(True)
for i in range(8): trainer.step(pool[i % 4])
torch.cuda.synchronize()
pr = _lib.profile_read()
print(os.environ.get("HCG_LIB_PATH", "default"), ["%.4f" % r for r in res], {k: round(v[0] / v[1] * 1e3) for k, v in pr.items() if "tc_" in k or "readout" in k or "collate" in k})
