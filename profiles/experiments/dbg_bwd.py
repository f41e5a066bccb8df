import sys, torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import hcatgnet_b200 as pkg
from hcatgnet_b200 import _lib, ops
from hcatgnet_b200.synth import synth_batch
from oracle import ref_gcn
import tests.test_gpu_parity as T
hcg = ops
for (F, nmin, nmax, B) in [(32, 1, 9, 700), (32, 2, 9, 700), (25, 1, 9, 700), (32, 5, 9, 100), (32, 1, 9, 20)]:
    b = synth_batch(B, n_features=F, seed=11, n_min=nmin, n_max=nmax)
    sd = ref_gcn.init_state_dict(F, 64, 2, 2, 1, seed=5)
    args = (sd, b.x, b.edge_index, b.batch, b.num_graphs, b.y)
    _, g_ref, _ = T._oracle_grads(*args)
    for path in (2, 3):
        prev = _lib.config_set(_lib.HCG_CFG_BACKWARD_PATH, path)
        _, g, _ = T._cuda_grads(hcg, *args)
        _lib.config_set(_lib.HCG_CFG_BACKWARD_PATH, prev)
        errs = {k: float((g[k].double().cpu() - g_ref[k]).abs().max() / max(float(g_ref[k].abs().max()), 1e-30)) for k in g_ref if k.startswith("conv")}
        print((F, nmin, nmax, B), "path", path, {k: f"{v:.2e}" for k, v in errs.items()})
