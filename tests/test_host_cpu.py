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
                      ("HcgGrads", _lib.HcgGrads), ("HcgSaved", _lib.HcgSaved), ("HcgEnsemble", _lib.HcgEnsemble)):
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


# ------------------------------------------------------------------------------------------
# packed graph image: format round trips and the host decoder
# ------------------------------------------------------------------------------------------
def _real_graphs(count=40):
    from tests import golden_io
    g = golden_io.graphs_by_idx("rhcaa_learning")
    return [g[k] for k in sorted(g)[:count]]


def _same_batch(a, b):
    assert torch.equal(a.x, b.x) and torch.equal(a.edge_index, b.edge_index) and torch.equal(a.batch, b.batch)
    assert torch.equal(a.ptr, b.ptr) and torch.equal(a.y, b.y) and a.num_graphs == b.num_graphs


def test_packed_image_round_trip_real_graphs(tmp_path):
    from hcatgnet_b200.data import Batch
    from hcatgnet_b200.packed import PackedDataset
    graphs = _real_graphs()
    ds = PackedDataset.from_graphs(graphs)
    assert ds.G == len(graphs) and ds.edge_kind == 1
    # the featuriser's one-hot / flag columns take one bit each, lossless: the image is many times smaller than the tensors
    full = Batch.from_data_list(graphs)
    assert ds.Fb + ds.Fu >= ds.F - 2 and ds.Fb >= ds.F - 4
    assert ds.nbytes * 6 < full.x.numel() * 4 + full.edge_index.numel() * 8
    old = PackedDataset.from_arrays(full.x.numpy(), full.edge_index.numpy(), full.ptr.numpy(), full.y.numpy(), full.idx.numpy(),
                                    bits=False)   # the byte-per-value split of round 1
    assert old.Fb == 0 and old.Fu == ds.Fb + ds.Fu and ds.nbytes < old.nbytes
    _same_batch(old.unpack_host(), full)
    _same_batch(ds.unpack_host(), full)
    ids = [7, 3, 3, 39, 0, 12]
    _same_batch(ds.unpack_host(ids), Batch.from_data_list([graphs[i] for i in ids]))
    path = str(tmp_path / "set.hcgp")
    ds.save(path)
    for mmap in (True, False):
        back = PackedDataset.load(path, mmap=mmap)
        assert back.G == ds.G and np.array_equal(np.asarray(back.blob), ds.blob)
        _same_batch(back.unpack_host(ids), ds.unpack_host(ids))
    with pytest.raises(ValueError):
        PackedDataset(np.zeros(256, np.uint8))


def test_packed_image_column_and_edge_kinds():
    from hcatgnet_b200.packed import PackedDataset
    gen = np.random.default_rng(0)
    # graph sizes above 256 force u16 endpoints; column 1 non-integer, column 2 negative, column 3 above 255 -> f32
    sizes = [300, 1, 17]
    graphs = []
    for n in sizes:
        x = gen.integers(0, 3, size=(n, 5)).astype(np.float32)
        x[:, 1] += 0.5
        x[0, 2] = -1.0
        x[0, 3] = 256.0
        e = gen.integers(0, n, size=(2, 2 * n)) if n > 1 else np.zeros((2, 0), np.int64)
        graphs.append({"x": torch.from_numpy(x), "edge_index": torch.from_numpy(e.astype(np.int64)), "y": torch.tensor([float(n)]),
                       "idx": n, "fold": n % 3})
    ds = PackedDataset.from_graphs(graphs)
    assert ds.edge_kind == 2 and ds.Fu == 2 and ds.Ff == 3 and ds.Fb == 0 and ds.max_graph_nodes == 300
    # bit columns: every value 0 or 1; a column holding -0.0 must stay f32 (the sign would be lost); 11 bit columns
    # need two bytes per node
    xb = (gen.random((40, 14)) < 0.4).astype(np.float32)
    xb[:, 12] *= 7.0                       # 0 / 7: a byte column
    xb[3, 13] = -0.0                       # compares equal to 0
    xb[:, 5] = gen.random(40).astype(np.float32)
    gb = [{"x": torch.from_numpy(xb[:25]), "edge_index": torch.tensor([[0, 1], [1, 0]]), "y": torch.tensor([1.0])},
          {"x": torch.from_numpy(xb[25:]), "edge_index": torch.zeros(2, 0, dtype=torch.int64), "y": torch.tensor([2.0])}]
    dsb = PackedDataset.from_graphs(gb)
    assert (dsb.Fb, dsb.Fu, dsb.Ff) == (11, 1, 2) and dsb.xbits.shape == (40, 2)
    back = dsb.unpack_host()
    assert np.array_equal(back.x.numpy().view(np.uint32), xb.view(np.uint32))      # bit patterns, -0.0 included
    assert list(ds.fold) == [0, 1, 2] and list(ds.idx) == sizes
    from hcatgnet_b200.data import Batch
    _same_batch(ds.unpack_host([2, 0, 1, 0]), Batch.from_data_list([graphs[i] for i in (2, 0, 1, 0)]))
    bad = {"x": torch.zeros(3, 5), "edge_index": torch.tensor([[0, 1], [1, 5]]), "y": torch.tensor([0.0])}
    with pytest.raises(ValueError):
        PackedDataset.from_graphs([bad, bad])      # an endpoint outside its own graph


@pytest.mark.skipif(not os.path.isdir("/root/reference/data/datasets/rhcaa_final_test/processed"),
                    reason="reference data set not present")
def test_packed_converter_reads_the_reference_pickles():
    """convert_processed_dir against the golden fixtures (which were made by importing the reference)."""
    from tests import golden_io
    from hcatgnet_b200.packed import convert_processed_dir
    ds = convert_processed_dir("/root/reference/data/datasets/rhcaa_final_test/processed")
    want = golden_io.graphs_by_idx("rhcaa_final_test")
    assert ds.G == len(want)
    got = ds.unpack_host()
    for b in range(ds.G):
        g = want[int(ds.idx[b])]
        lo, hi = int(got.ptr[b]), int(got.ptr[b + 1])
        assert torch.equal(got.x[lo:hi], torch.as_tensor(g["x"]))
        sel = (got.edge_index[0] >= lo) & (got.edge_index[0] < hi)
        assert torch.equal(got.edge_index[:, sel] - lo, torch.as_tensor(g["edge_index"]).to(torch.int64))


def test_create_loaders_nested_cv_partitions():
    """reference call_methods.py:15-47: folds x (folds - 1) triples; test = outer fold, val = inner-th of the
    remaining folds, train = the rest in fold order."""
    import hcatgnet_b200 as pkg
    graphs = _real_graphs(120)
    opt = pkg.default_options()
    triples = list(pkg.create_loaders(graphs, opt))
    assert len(triples) == opt.folds * (opt.folds - 1)
    all_idx = sorted(int(g["idx"]) for g in graphs)
    for n, (tr, va, te) in enumerate(triples):
        outer, inner = divmod(n, opt.folds - 1)
        rest = [f for f in range(opt.folds) if f != outer]
        assert all(int(g["fold"]) == outer for g in te.dataset)
        assert all(int(g["fold"]) == rest[inner] for g in va.dataset)
        want_train = [int(g["idx"]) for f in rest if f != rest[inner] for g in graphs if int(g["fold"]) == f]
        assert [int(g["idx"]) for g in tr.dataset] == want_train and tr.shuffle and not va.shuffle and not te.shuffle
        assert sorted([int(g["idx"]) for l in (tr, va, te) for g in l.dataset]) == all_idx
