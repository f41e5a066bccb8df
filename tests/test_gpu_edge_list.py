"""hcg_gcn_forward_coo (inference straight from Batch.x / Batch.edge_index / Batch.ptr: the conv kernel builds the
normalised neighbour lists itself) against hcg_collate + hcg_gcn_forward on the same batch: identical bits, and against
the CPU oracle.  Reference call replaced: `model(batch)` under no_grad, utils/utils_model.py:83-96 -> model/gcn.py:52-76."""
import numpy as np
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hcatgnet_b200 as p
    return p


def _both(pkg, batch, n_features=25, seed=3, opt_kw=None):
    from hcatgnet_b200 import ops
    opt = pkg.default_options()
    for k, v in (opt_kw or {}).items():
        setattr(opt, k, v)
    torch.manual_seed(seed)
    model = pkg.make_network("GCN", opt, n_features).cuda().eval()
    b = batch.to("cuda")
    params = [p.detach() for p in model._flat_params()]
    res = ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, model._dims(), params, validate=True)
    plan = ops.collate(b.edge_index, b.batch, b.x.shape[0], b.num_graphs, validate=True)
    out_ref, emb_ref, _ = ops.forward_raw(plan, b.x, model._dims(), params, save=False)
    return res, out_ref, emb_ref, model, b


@pytest.mark.parametrize("n_graphs,F", [(1, 25), (7, 25), (300, 25), (4099, 25), (513, 32), (129, 11)])
def test_edge_list_forward_equals_collate_forward_bit_for_bit(pkg, n_graphs, F):
    from hcatgnet_b200.synth import synth_batch
    batch = synth_batch(n_graphs, F, seed=100 + n_graphs, features="uniform" if F != 25 else "featuriser")
    res, out_ref, emb_ref, _, _ = _both(pkg, batch, F)
    assert res is not None
    out, emb = res
    assert torch.equal(emb, emb_ref)
    assert torch.equal(out, out_ref)


def test_shuffled_edges_unsorted_rows_and_isolated_nodes(pkg):
    """Edges of a graph in arbitrary order (the claim order of the shared-memory atomics must not show), nodes without
    edges, a graph without any edge, in-degrees up to the cap of 16."""
    g = torch.Generator().manual_seed(9)
    graphs = []
    for k in range(257):
        n = int(torch.randint(1, 129, (1,), generator=g))
        if k % 17 == 0 or n == 1:
            ei = torch.zeros(2, 0, dtype=torch.int64)
        else:
            # random simple directed graph with in-degree <= 16: every node draws up to 16 distinct sources (mostly few)
            src, dst = [], []
            for d in range(n):
                if k % 5 == 0 and d % 3 == 0:
                    continue                    # isolated on the incoming side
                cap = 16 if d % 7 == 0 else 5
                deg = int(torch.randint(0, min(cap, n - 1) + 1, (1,), generator=g))
                cand = torch.randperm(n, generator=g)
                cand = cand[cand != d][:deg]
                src += cand.tolist(); dst += [d] * len(cand)
            ei = torch.tensor([src, dst], dtype=torch.int64).reshape(2, -1)
            ei = ei[:, torch.randperm(ei.shape[1], generator=g)][:, :512]
        graphs.append({"x": torch.randn(n, 25, generator=g), "edge_index": ei, "y": torch.zeros(1), "idx": k})
    batch = pkg.Batch.from_data_list(graphs)
    assert 8 < batch.max_in_degree <= 16
    res, out_ref, emb_ref, _, _ = _both(pkg, batch)
    assert res is not None
    # A graph WITHOUT edges never reaches the tensor-core kernel on the collate path (its work list wants at least one
    # CSR slot: the FMA graph-resident kernel serves it, fp32 FMAs instead of the 3-term TF32 split), so those rows agree
    # to rounding only; every other graph must agree bit for bit.
    no_edges = (batch.edge_ptr[1:] == batch.edge_ptr[:-1]).to("cuda")
    differ = (res[1] != emb_ref).any(1) | (res[0] != out_ref).any(1)
    assert not bool((differ & ~no_edges).any()), differ.nonzero().flatten().tolist()
    assert torch.allclose(res[1], emb_ref, rtol=1e-5, atol=1e-6) and torch.allclose(res[0], out_ref, rtol=1e-5, atol=1e-5)


def test_against_the_cpu_oracle(pkg):
    """Through the oracle's restatement of the reference network (oracle/ref_gcn.py), shuffled edge order included."""
    from hcatgnet_b200.synth import synth_batch
    from oracle import ref_gcn
    batch = synth_batch(64, 25, seed=77)
    res, _, _, model, _ = _both(pkg, batch)
    sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    out, emb = ref_gcn.forward(sd, batch.x.cpu(), batch.edge_index.cpu(), batch.batch.cpu(), 64)   # (.to() moved the batch in place)
    # fp32 everywhere; the kernel's 3-term TF32 split and the summation order differ from torch's CPU kernels
    assert torch.allclose(res[1].cpu(), emb, rtol=1e-4, atol=1e-5)
    assert torch.allclose(res[0].cpu(), out, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("kind", ["self_loop", "degree", "too_many_nodes", "cross_edge", "too_many_edges"])
def test_refused_batches_are_reported_and_the_module_falls_back(pkg, kind):
    from hcatgnet_b200 import ops
    g = torch.Generator().manual_seed(4)
    n = 140 if kind == "too_many_nodes" else 20
    ring = torch.arange(n)
    ei = torch.stack([torch.cat([ring, (ring + 1) % n]), torch.cat([(ring + 1) % n, ring])])
    if kind == "self_loop":
        ei = torch.cat([ei, torch.tensor([[3], [3]])], 1)
    elif kind == "degree":
        ei = torch.cat([ei, torch.stack([torch.arange(2, 19), torch.zeros(17, dtype=torch.int64)])], 1)
    elif kind == "too_many_edges":
        extra = torch.stack([torch.randint(0, n, (600,), generator=g), torch.randint(0, n, (600,), generator=g)])
        ei = torch.cat([ei, extra[:, extra[0] != extra[1]]], 1)
    graphs = [{"x": torch.randn(n, 25, generator=g), "edge_index": ei, "y": torch.zeros(1), "idx": 0},
              {"x": torch.randn(5, 25, generator=g), "edge_index": torch.tensor([[0, 1], [1, 0]]), "y": torch.zeros(1), "idx": 1}]
    batch = pkg.Batch.from_data_list(graphs)
    if kind == "cross_edge":
        batch.edge_index[0, 0] = n + 1      # first edge of graph 0 now starts in graph 1
    b = batch.to("cuda")
    torch.manual_seed(0)
    model = pkg.make_network("GCN", pkg.default_options(), 25).cuda().eval()
    params = [p.detach() for p in model._flat_params()]
    assert ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, model._dims(), params, validate=True) is None
    if kind != "cross_edge":
        # the module: hints say "not this path" or the device refuses -> same answer as the general path
        with torch.no_grad():
            out = model(b)
        plan = ops.collate(b.edge_index, b.batch, b.x.shape[0], 2, validate=True)
        want, _, _ = ops.forward_raw(plan, b.x, model._dims(), params, save=False)
        # (the general path picks its kernels from the batch's size hints, which this direct call does not pass:
        #  fp32 FMA vs 3-term TF32 for the oversize graph - agreement to rounding)
        assert torch.allclose(out, want, rtol=1e-5, atol=1e-6)


def test_module_takes_the_edge_list_path_for_inference_only(pkg):
    """model(batch) under no_grad on a batch that carries its offsets and size facts launches no collate kernel;
    with grad enabled it goes through the plan (backward needs the transposed CSR)."""
    from hcatgnet_b200 import _lib
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(500, 25, seed=5).to("cuda")
    torch.manual_seed(1)
    model = pkg.make_network("GCN", pkg.default_options(), 25).cuda()
    model.eval()
    with torch.no_grad():
        n0 = _lib.launch_count()
        out = model(b)
        launched = _lib.launch_count() - n0
    assert not hasattr(b, "_hcg_plan")
    assert launched == 3, launched            # k_tile_desc, k_tc_fwd, k_readout2
    model.train()
    out2 = model(b)
    assert hasattr(b, "_hcg_plan")
    assert torch.equal(out, out2.detach())


# ---- hcg_gcn_forward_packed: the same forward straight from a device-resident packed image (no unpack) ----------------
def _packed_pair(pkg, ds, first, count, F, seed=5):
    from hcatgnet_b200 import ops
    torch.manual_seed(seed)
    model = pkg.make_network("GCN", pkg.default_options(), F).cuda().eval()
    params = [p.detach() for p in model._flat_params()]
    store = pkg.DeviceGraphStore(ds, "cuda")
    got = ops.forward_packed(store.c_store, first, count, model._dims(), params, torch.device("cuda", 0), validate=True)
    b = store.batch(np.arange(first, first + count))
    want = ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, model._dims(), params, validate=True)
    return got, want, model, store


@pytest.mark.parametrize("n_graphs,F,style", [(1, 25, "featuriser"), (700, 25, "featuriser"), (3001, 25, "uniform"), (257, 32, "mixed"),
                                              (64, 9, "mixed")])
def test_packed_forward_equals_unpack_plus_edge_list_forward(pkg, n_graphs, F, style):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(n_graphs, F, seed=900 + n_graphs, features="featuriser" if style == "featuriser" else "uniform")
    x = b.x.clone()
    if style == "mixed":                      # bit, byte and f32 columns interleaved
        x[:, 0::3] = (x[:, 0::3] < 0.4).float()
        x[:, 1::3] = torch.floor(x[:, 1::3] * 7)
    ds = pkg.PackedDataset.from_arrays(x.numpy(), b.edge_index.numpy(), b.ptr.numpy(), b.y.numpy())
    if style == "mixed":
        assert ds.Fb > 0 and ds.Fu > 0 and ds.Ff > 0
    if style == "uniform" and F == 25:        # 25 f32 columns = 100 packed bytes per node: above the kernel's stage
        got, want, _, _ = _packed_pair(pkg, ds, 0, n_graphs, F)
        assert got is None and want is not None
        return
    for first, count in ((0, n_graphs), (n_graphs // 3, n_graphs - n_graphs // 3)):
        got, want, _, _ = _packed_pair(pkg, ds, first, count, F)
        assert got is not None and want is not None
        assert torch.equal(got[1], want[1]) and torch.equal(got[0], want[0])


def test_predict_stream_runs_packed_chunks_without_unpacking(pkg):
    from hcatgnet_b200 import _lib
    from hcatgnet_b200.synth import synth_batch
    chunks, hosts = [], []
    for k in range(3):
        b = synth_batch(400 + 37 * k, 25, seed=70 + k, features="featuriser")
        chunks.append(pkg.PackedDataset.from_arrays(b.x.numpy(), b.edge_index.numpy(), b.ptr.numpy(), b.y.numpy(), pinned=True))
        hosts.append(b)
    torch.manual_seed(3)
    model = pkg.make_network("GCN", pkg.default_options(), 25).cuda().eval()
    want = [(o.clone(), e.clone()) for o, e, _ in pkg.predict_stream(model, hosts, return_embedding=True)]
    n0 = _lib.launch_count()
    got = [(o.clone(), e.clone()) for o, e, _ in pkg.predict_stream(model, chunks, return_embedding=True)]
    assert _lib.launch_count() - n0 == 3 * len(chunks)          # k_tile_desc, k_tc_fwd_packed, k_readout2: no k_unpack_batch
    for (o, e), (ow, ew) in zip(got, want):
        assert torch.equal(o, ow) and torch.equal(e, ew)
    # the lazy batch unpacks itself when somebody reads it (training, or a model that takes another path)
    store = pkg.DeviceGraphStore(chunks[0], "cuda")
    lazy = store.lazy_batch()
    assert lazy._real is None and lazy.num_graphs == 400
    assert torch.equal(lazy.x.cpu(), hosts[0].x) and torch.equal(lazy.edge_index.cpu(), hosts[0].edge_index)
    out = model(lazy)                                            # grad mode: the collate path, on the unpacked tensors
    assert torch.allclose(out.detach().cpu(), want[0][0], rtol=1e-5, atol=1e-6)


def test_packed_forward_on_edgeless_graphs_single_nodes_and_all_bit_columns(pkg):
    """Graphs without edges, one-node graphs, shuffled edge order; an image whose columns are ALL bit columns (no byte, no
    f32 section) and one with f32 columns only."""
    g = torch.Generator().manual_seed(23)
    for F, kind in ((12, "bits"), (9, "f32")):
        graphs = []
        for k in range(97):
            n = 1 if k % 11 == 0 else int(torch.randint(2, 129, (1,), generator=g))
            if k % 7 == 0 or n == 1:
                ei = torch.zeros(2, 0, dtype=torch.int64)
            else:
                src = torch.randint(0, n, (3 * n,), generator=g)
                dst = (src + 1 + torch.randint(0, n - 1, (3 * n,), generator=g)) % n        # never a self-loop
                key = torch.unique(dst * n + src)[:500]                                      # simple graph
                ei = torch.stack([key % n, key // n])[:, torch.randperm(key.numel(), generator=g)]
                ei = ei[:, torch.bincount(ei[1], minlength=n)[ei[1]] <= 16]                  # in-degree cap of the gather kernel
            x = (torch.rand(n, F, generator=g) < 0.3).float() if kind == "bits" else torch.randn(n, F, generator=g)
            graphs.append({"x": x, "edge_index": ei, "y": torch.zeros(1), "idx": k})
        ds = pkg.PackedDataset.from_graphs(graphs)
        assert (ds.Fb, ds.Fu, ds.Ff) == ((F, 0, 0) if kind == "bits" else (0, 0, F))
        got, want, _, _ = _packed_pair(pkg, ds, 0, len(graphs), F)
        assert got is not None and want is not None
        assert torch.equal(got[1], want[1]) and torch.equal(got[0], want[0])
