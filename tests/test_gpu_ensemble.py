"""One-pass ensemble forward (hcg_gcn_forward_coo_ensemble; SURVEY.md section 8 f2: the reference loops whole
`predict_network` calls over the trained models, scripts_experiments/predict_test.py:47-70): M models over one batch in
one kernel pass must give, PER MODEL, the bits of the single-model edge-list forward (same arithmetic, same order)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hcatgnet_b200 as p
    return p


def _models(pkg, count, F=25, seed=0):
    out = []
    for k in range(count):
        torch.manual_seed(seed + k)
        m = pkg.make_network("GCN", pkg.default_options(), F).cuda().eval()
        with torch.no_grad():          # non-trivial biases (the default init leaves the conv biases at zero)
            for p in m.parameters():
                if p.dim() == 1:
                    p.uniform_(-0.3, 0.3)
        out.append(m)
    return out


def _compare(pkg, models, batch):
    from hcatgnet_b200 import ops
    b = batch.to("cuda")
    dims = models[0]._dims()
    params = [[p.detach() for p in m._flat_params()] for m in models]
    w = ops.EnsembleWeights(dims, params, b.x.device)
    res = ops.forward_edge_list_ensemble(b.x, b.edge_index, b.ptr, b.edge_ptr, w, validate=True)
    assert res is not None
    out, emb = res
    assert out.shape == (len(models), b.num_graphs, dims.C) and emb.shape == (len(models), b.num_graphs, 2 * dims.H)
    for k, ps in enumerate(params):
        one = ops.forward_edge_list(b.x, b.edge_index, b.ptr, b.edge_ptr, dims, ps, validate=True)
        assert one is not None
        assert torch.equal(emb[k], one[1]), k
        assert torch.equal(out[k], one[0]), k


@pytest.mark.parametrize("n_models,n_graphs,F", [(1, 1, 25), (1, 300, 25), (2, 7, 25), (5, 700, 25), (3, 2500, 32), (4, 129, 11),
                                                 (17, 64, 25)])
def test_ensemble_pass_equals_the_single_model_forward_per_model(pkg, n_models, n_graphs, F):
    from hcatgnet_b200.synth import synth_batch
    batch = synth_batch(n_graphs, F, seed=300 + n_graphs, features="uniform" if F != 25 else "featuriser")
    _compare(pkg, _models(pkg, n_models, F, seed=n_graphs), batch)


def test_ensemble_pass_on_shuffled_edges_isolated_nodes_and_edgeless_graphs(pkg):
    g = torch.Generator().manual_seed(19)
    graphs = []
    for k in range(150):
        n = int(torch.randint(1, 129, (1,), generator=g))
        if k % 13 == 0 or n == 1:
            ei = torch.zeros(2, 0, dtype=torch.int64)
        else:
            src, dst = [], []
            for d in range(n):
                if k % 5 == 0 and d % 3 == 0:
                    continue
                deg = int(torch.randint(0, min(16 if d % 7 == 0 else 5, n - 1) + 1, (1,), generator=g))
                cand = torch.randperm(n, generator=g)
                cand = cand[cand != d][:deg]
                src += cand.tolist(); dst += [d] * len(cand)
            ei = torch.tensor([src, dst], dtype=torch.int64).reshape(2, -1)
            ei = ei[:, torch.randperm(ei.shape[1], generator=g)][:, :512]
        graphs.append({"x": torch.randn(n, 25, generator=g), "edge_index": ei, "y": torch.zeros(1), "idx": k})
    _compare(pkg, _models(pkg, 3, seed=50), pkg.Batch.from_data_list(graphs))


def test_ensemble_pass_against_the_cpu_oracle(pkg):
    from hcatgnet_b200 import ops
    from hcatgnet_b200.synth import synth_batch
    from oracle import ref_gcn
    models = _models(pkg, 3, seed=7)
    batch = synth_batch(48, 25, seed=8)
    want = []
    for m in models:
        sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
        want.append(ref_gcn.forward(sd, batch.x, batch.edge_index, batch.batch, 48))
    b = batch.to("cuda")
    w = ops.EnsembleWeights(models[0]._dims(), [[p.detach() for p in m._flat_params()] for m in models], b.x.device)
    out, emb = ops.forward_edge_list_ensemble(b.x, b.edge_index, b.ptr, b.edge_ptr, w)
    for k, (o, e) in enumerate(want):
        # fp32 end to end; split-TF32 products and another summation order than torch's CPU kernels
        assert float((emb[k].cpu() - e).abs().max()) <= 2e-5 * float(e.abs().max()), k
        assert float((out[k].cpu() - o).abs().max()) <= 2e-5 * max(1.0, float(o.abs().max())), k


def test_predict_ensemble_stream_one_pass_equals_per_model_launches(pkg):
    from hcatgnet_b200 import _lib
    from hcatgnet_b200.synth import synth_batch
    models = _models(pkg, 4, seed=11)
    batches = [synth_batch(n, 25, seed=60 + n) for n in (200, 1, 333)]
    runs = {}
    for one_pass in (True, False):
        n0 = _lib.launch_count()
        runs[one_pass] = [(p.clone(), mean.clone(), std.clone(), emb.clone())
                          for p, mean, std, _h, emb in pkg.predict_ensemble_stream(models, iter(batches), one_pass=one_pass,
                                                                                    return_embedding=True)]
        runs[one_pass, "launches"] = _lib.launch_count() - n0
    for a, b in zip(runs[True], runs[False]):
        for u, v in zip(a, b):
            assert torch.equal(u, v)
    # per batch: k_ens_prep, k_tile_desc, k_tc_fwd_ens, k_readout2_ens (the batch vector is never built)
    assert runs[True, "launches"] == 4 * len(batches)
    assert runs[False, "launches"] > runs[True, "launches"]


def test_refused_batch_falls_back(pkg):
    from hcatgnet_b200 import ops
    g = torch.Generator().manual_seed(4)
    n = 140                                                     # above the tile's 128 rows
    ring = torch.arange(n)
    ei = torch.stack([torch.cat([ring, (ring + 1) % n]), torch.cat([(ring + 1) % n, ring])])
    graphs = [{"x": torch.randn(n, 25, generator=g), "edge_index": ei, "y": torch.zeros(1), "idx": 0},
              {"x": torch.randn(5, 25, generator=g), "edge_index": torch.tensor([[0, 1], [1, 0]]), "y": torch.zeros(1), "idx": 1}]
    models = _models(pkg, 2, seed=3)
    b = pkg.Batch.from_data_list(graphs).to("cuda")
    w = ops.EnsembleWeights(models[0]._dims(), [[p.detach() for p in m._flat_params()] for m in models], b.x.device)
    assert ops.forward_edge_list_ensemble(b.x, b.edge_index, b.ptr, b.edge_ptr, w, validate=True) is None
    host = pkg.Batch.from_data_list(graphs)
    (p, mean, std, _h), = list(pkg.predict_ensemble_stream(models, [host]))
    with torch.no_grad():
        for k, m in enumerate(models):
            assert torch.allclose(p[k], m(b).cpu(), rtol=1e-5, atol=1e-6)
    with pytest.raises(ValueError):
        list(pkg.predict_ensemble_stream(models, [pkg.Batch.from_data_list(graphs)], one_pass=True))
