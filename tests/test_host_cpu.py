"""CPU-only checks: the C-ABI library loads and exports every symbol include/hcg.h declares,
the host-side mirror keeps the reference's module surface, and there is no CPU fallback."""
import ctypes
import io
import os
import re

import numpy as np
import pytest
import torch

from oracle import ref_gcn
from tests import golden_io

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    src = open(os.path.join(ROOT, "include", "hcg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hcg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from hcatgnet_b200 import _lib
    names = _declared_functions()
    assert "hcg_collate" in names and "hcg_gcn_forward" in names and "hcg_gcn_backward" in names
    _lib.load()  # rebuilds first when the sources are newer and nvcc exists
    lib = ctypes.CDLL(_lib.lib_path())
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/hcg.h but not exported by libhcg.so"
    assert sorted(_lib.SIGNATURES) == names, "ctypes table and header disagree"
    assert b"sm_100a" in _lib.load().hcg_version()


def test_struct_layouts_match_header_sizes():
    from hcatgnet_b200 import _lib
    lib = _lib.load()
    for name, cls in (("HcgGraph", _lib.HcgGraph), ("HcgParams", _lib.HcgParams),
                      ("HcgGrads", _lib.HcgGrads), ("HcgSaved", _lib.HcgSaved)):
        fn = getattr(lib, "hcg_sizeof_" + name.lower()[3:])
        fn.restype = ctypes.c_size_t
        assert fn() == ctypes.sizeof(cls), name


def test_argument_errors_without_gpu():
    from hcatgnet_b200 import _lib
    lib = _lib.load()
    p = _lib.HcgParams()
    p.F, p.H, p.L, p.R, p.C = 25, 64, 99, 2, 1
    assert lib.hcg_forward_workspace_bytes(10, 10, 1, ctypes.byref(p)) == 0  # L out of range
    rc = lib.hcg_collate(None, -1, None, 0, 0, 1.0, None, 1.0, 0, None, None, None, None, None, None, None, None,
                         None, None, None, None, None, 0, None)
    assert rc == 1 and b"negative" in lib.hcg_last_error()
    assert lib.hcg_collate_workspace_bytes(1000, 3000, 10, 1) > lib.hcg_collate_workspace_bytes(1000, 3000, 10, 0) > 0


def test_module_surface_and_state_dict_layout():
    import hcatgnet_b200 as pkg
    opt = pkg.default_options()
    m = pkg.make_network("GCN", opt, 25)
    assert m.name == "GCN" and isinstance(m.loss, torch.nn.MSELoss)
    assert isinstance(m.optimizer, torch.optim.Adam) and m.optimizer.defaults["eps"] == 1e-9
    assert isinstance(m.scheduler, torch.optim.lr_scheduler.ReduceLROnPlateau)
    assert list(m.state_dict().keys()) == ref_gcn.state_dict_keys(2, 2)
    sd, _ = golden_io.load_model("learning", 1, 2)
    m.load_state_dict(sd, strict=True)
    assert sum(p.numel() for p in m.parameters()) == 14145
    m4 = pkg.GCN(pkg.default_options(n_convolutions=4, embedding_dim=256, readout_layers=3), 32)
    assert list(m4.state_dict().keys()) == ref_gcn.state_dict_keys(4, 3)
    assert m4.state_dict()["readout.1.0.weight"].shape == (128, 256)
    with pytest.raises(ValueError):
        pkg.make_network("GAT", opt, 25)
    with pytest.raises(ValueError):
        pkg.GCN(pkg.default_options(problem_type="ranking"), 25)
    with pytest.raises(NotImplementedError):
        pkg.GCN(pkg.default_options(optimizer="lbfgs"), 25)
    with pytest.raises(NotImplementedError):
        pkg.GCN(pkg.default_options(scheduler="cosine"), 25)
    buf = io.BytesIO()
    torch.save(m, buf)  # reference utils/utils_model.py:143 pickles the whole module
    buf.seek(0)
    assert torch.equal(torch.load(buf, weights_only=False).conv1.bias, m.conv1.bias)


def test_constructor_reseeds_and_init_is_reproducible():
    import hcatgnet_b200 as pkg
    a = pkg.GCN(pkg.default_options(), 25)
    torch.rand(3)
    b = pkg.GCN(pkg.default_options(), 25)
    for (k, va), vb in zip(a.state_dict().items(), b.state_dict().values()):
        assert torch.equal(va, vb), k
    w = a.conv1.lin.weight
    assert w.abs().max() <= (6.0 / (25 + 64)) ** 0.5 and torch.count_nonzero(a.conv1.bias) == 0


def test_batch_container_matches_oracle_collate():
    import hcatgnet_b200 as pkg
    graphs = golden_io.load_graphs("biAryl")[:17]
    b = pkg.Batch.from_data_list(graphs)
    r = ref_gcn.collate(graphs)
    for k in ("x", "edge_index", "batch", "ptr", "y", "idx"):
        assert torch.equal(getattr(b, k), r[k]), k
    assert b.num_graphs == 17 and b.to("cpu") is b
    loader = pkg.DataLoader(graphs, batch_size=5, shuffle=True)
    assert len(loader) == 4 and len(loader.dataset) == 17
    assert sorted(int(i) for bb in loader for i in bb.idx) == sorted(g["idx"] for g in graphs)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU behaviour")
def test_no_cpu_fallback():
    import hcatgnet_b200 as pkg
    graphs = golden_io.load_graphs("rhcaa_learning")[:2]
    m = pkg.GCN(pkg.default_options(), 25)
    with pytest.raises(RuntimeError, match="CUDA device"):
        m(pkg.Batch.from_data_list(graphs))


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "hcatgnet_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_synthetic_generator_shape():
    from hcatgnet_b200.synth import synth_batch_numpy
    d = synth_batch_numpy(256, seed=1)
    ei, N = d["edge_index"], d["x"].shape[0]
    assert ei.dtype == np.int64 and d["batch"].dtype == np.int64 and d["x"].dtype == np.float32
    assert abs(ei.shape[1] / N - 3.0) < 0.01 and (ei[0] != ei[1]).all()
    assert np.array_equal(ei[:, 0::2], ei[::-1, 1::2])               # (i,j),(j,i) adjacent pairs
    assert len(np.unique(ei[0] * N + ei[1])) == ei.shape[1]          # no duplicate bonds
    assert (d["batch"][ei[0]] == d["batch"][ei[1]]).all()            # no bond crosses graphs
    n = np.diff(d["ptr"])
    assert n.min() >= 40 and n.max() <= 120
    d2 = synth_batch_numpy(256, seed=1)
    assert np.array_equal(d2["edge_index"], ei) and np.array_equal(d2["x"], d["x"])
