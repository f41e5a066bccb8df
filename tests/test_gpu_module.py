"""GPU tests of the drop-in boundary: GCN / GCN_explain nn.Modules driven the way the reference's
train / eval / predict loops drive them (reference utils/utils_model.py:55-111)."""
import io

import numpy as np
import pytest
import torch

from oracle import ref_gcn
from tests import golden_io

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    assert torch.cuda.is_available()
    import hcatgnet_b200
    return hcatgnet_b200


def _batches(pkg, graphs, bs=40):
    return list(pkg.DataLoader(graphs, batch_size=bs, shuffle=False))


def test_predict_loop_reproduces_reference_csv(pkg):
    """load shipped weights -> model(batch, True) over the data set -> the rows of embeddings.csv."""
    sd, graphs, emb_gold, pred_gold = golden_io.rows_with_graphs("final_test", 1, 2)
    model = pkg.make_network("GCN", pkg.default_options(), 25)
    model.load_state_dict(sd, strict=True)
    # exactly what predict_network does: model and batches forced to the CPU, numpy() on the results
    model.to("cpu"); model.eval()
    preds, embs = [], []
    for batch in _batches(pkg, graphs):
        batch = batch.to("cpu")
        out, emb = model(batch, True)
        preds.append(out.cpu().detach().numpy()); embs.append(emb.detach().numpy())
    pred, emb = np.concatenate(preds).ravel(), np.concatenate(embs)
    assert (np.abs(emb - emb_gold) / np.maximum(np.abs(emb_gold), 1.0)).max() <= 1e-5   # north-star bound vs the reference CSV
    assert np.abs(pred - pred_gold).max() <= 2e-5


def test_train_loop_tracks_oracle(pkg):
    """A few Adam steps of the reference's train_network body, CUDA module vs plain-torch oracle."""
    graphs = golden_io.load_graphs("rhcaa_learning")[:120]
    opt = pkg.default_options()
    model = pkg.make_network("GCN", opt, 25).to("cuda")
    sd0 = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    ref = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    ref_opt = torch.optim.Adam(list(ref.values()), lr=opt.lr, eps=1e-9)
    model.train()
    for epoch in range(2):
        for batch in _batches(pkg, graphs):
            batch = batch.to("cuda")
            model.optimizer.zero_grad()
            out = model(batch)
            loss = torch.sqrt(model.loss(out, torch.unsqueeze(batch.y, dim=1)))
            loss.backward()
            model.optimizer.step()
            b = batch.to("cpu")
            ref_opt.zero_grad()
            o_ref, _ = ref_gcn.forward(ref, b.x, b.edge_index, b.batch, b.num_graphs)
            l_ref = ref_gcn.rmse_loss(o_ref, b.y)
            l_ref.backward()
            ref_opt.step()
            assert abs(loss.item() - l_ref.item()) <= 2e-3 * max(1.0, abs(l_ref.item()))
    for k, v in model.state_dict().items():
        assert torch.allclose(v.cpu(), ref[k].detach(), rtol=0, atol=5e-3), k


def test_plan_is_cached_on_batch_and_eval_needs_no_grad(pkg):
    graphs = golden_io.load_graphs("rhcaa_learning")[:40]
    model = pkg.make_network("GCN", pkg.default_options(), 25).to("cuda").eval()
    batch = pkg.Batch.from_data_list(graphs).to("cuda")
    with torch.no_grad():
        a0 = model(batch)   # inference on a batch that carries its offsets and size facts: straight from the edge list, no plan
    assert not hasattr(batch, "_hcg_plan")
    c = model(batch)  # grad mode on (reference eval_network has no no_grad): collate with the transpose, plan cached
    plan = batch._hcg_plan[1]
    assert plan.has_transpose and torch.equal(a0, c.detach())
    with torch.no_grad():
        a = model(batch)    # a cached plan is reused by later inference calls
        b = model(batch)
    assert batch._hcg_plan[1] is plan and torch.equal(a, b) and torch.equal(a, a0)
    # a batch without the host-side size facts (e.g. built by hand): inference collates, and caches the plan
    bare = pkg.Batch(batch.x, batch.edge_index, batch.batch, None, batch.y, None, batch.num_graphs)
    with torch.no_grad():
        d = model(bare)
    assert hasattr(bare, "_hcg_plan") and torch.equal(d, a0)


def test_explain_signature_and_pickle(pkg):
    sd, graphs, _, pred_gold = golden_io.rows_with_graphs("learning", 1, 2)
    m = pkg.GCN_explain(pkg.default_options(), 25)
    m.load_state_dict(sd)
    m = m.to("cuda")
    g = graphs[0]
    out = m(x=g["x"].cuda(), edge_index=g["edge_index"].cuda(), batch_index=torch.zeros(g["x"].shape[0], dtype=torch.int64).cuda())
    assert abs(out.item() - pred_gold[0]) <= 2e-5
    out2 = m(x=g["x"].cuda(), edge_index=g["edge_index"].cuda())  # batch_index=None: one graph
    assert torch.equal(out, out2)
    # explicit unit weights + improved=True => self-loop fill 2.0: differs from the unweighted path
    out3 = m(x=g["x"].cuda(), edge_index=g["edge_index"].cuda(), edge_weight=torch.ones(g["edge_index"].shape[1]).cuda())
    o_ref, _ = ref_gcn.forward(sd, g["x"], g["edge_index"], torch.zeros(g["x"].shape[0], dtype=torch.int64), 1,
                               edge_weight=torch.ones(g["edge_index"].shape[1]), improved=True)
    assert abs(out3.item() - o_ref.item()) <= 1e-5 * max(1, abs(o_ref.item()))
    # torch.save(model) of the whole module (reference utils/utils_model.py:143)
    buf = io.BytesIO()
    torch.save(m, buf)
    buf.seek(0)
    m2 = torch.load(buf, weights_only=False)
    assert torch.equal(m2(x=g["x"].cuda(), edge_index=g["edge_index"].cuda()), out)


def test_predict_stream_matches_blocking_forward():
    """hcatgnet_b200.predict_stream (host batches, copies overlapped on a side stream) returns, batch
    by batch and in order, exactly what the blocking model(batch) call returns."""
    import hcatgnet_b200 as pkg
    from hcatgnet_b200.synth import synth_batch
    model = pkg.make_network("GCN", pkg.default_options(), 25).to("cuda:0").eval()
    host = [synth_batch(200 + 17 * i, 25, seed=100 + i, pin=True) for i in range(5)]
    got = list(pkg.predict_stream(model, host, return_embedding=True))
    assert len(got) == len(host)
    for (o, e, h), hb in zip(got, host):
        assert h is hb and not o.is_cuda and o.shape == (hb.num_graphs, 1)
        with torch.no_grad():
            o2, e2 = model(pkg.Batch(hb.x.cuda(), hb.edge_index.cuda(), hb.batch.cuda(), None, None, None, hb.num_graphs), True)
        assert torch.equal(o, o2.cpu()) and torch.equal(e, e2.cpu())
    yp, yt, idx, emb = pkg.predict_network_stream(model, host, return_emb=True)
    assert yp.shape == (sum(h.num_graphs for h in host),) and emb.shape[1] == 128 and yt.shape == yp.shape


def test_fused_trainer_tracks_oracle_and_autograd_path(pkg):
    """FusedTrainer.step (collate, forward, rmse seed, backward, Adam kernel; no autograd) against the
    plain-torch oracle + torch.optim.Adam, and against this package's own autograd path."""
    graphs = golden_io.load_graphs("rhcaa_learning")[:120]
    opt = pkg.default_options()
    model = pkg.make_network("GCN", opt, 25).to("cuda")
    twin = pkg.make_network("GCN", opt, 25).to("cuda")
    twin.load_state_dict(model.state_dict())
    sd0 = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    keys = list(sd0.keys())
    ref = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    ref_opt = torch.optim.Adam(list(ref.values()), lr=opt.lr, eps=1e-9)
    trainer = pkg.FusedTrainer(model)
    assert list(model.state_dict().keys()) == keys          # flattening keeps the state_dict layout
    model.train(); twin.train()
    for epoch in range(2):
        for batch in _batches(pkg, graphs):
            batch = batch.to("cuda")
            loss = trainer.step(batch)
            twin.optimizer.zero_grad()
            l_twin = torch.sqrt(twin.loss(twin(batch), torch.unsqueeze(batch.y, dim=1)))
            l_twin.backward()
            twin.optimizer.step()
            b = batch.to("cpu")
            ref_opt.zero_grad()
            o_ref, _ = ref_gcn.forward(ref, b.x, b.edge_index, b.batch, b.num_graphs)
            l_ref = ref_gcn.rmse_loss(o_ref, b.y)
            l_ref.backward()
            ref_opt.step()
            assert abs(loss.item() - l_ref.item()) <= 2e-3 * max(1.0, abs(l_ref.item()))
            assert abs(loss.item() - l_twin.item()) <= 2e-3 * max(1.0, abs(l_twin.item()))
    for k, v in model.state_dict().items():
        assert torch.allclose(v.cpu(), ref[k].detach(), rtol=0, atol=5e-3), k
        assert torch.allclose(v, twin.state_dict()[k], rtol=0, atol=5e-3), k
    # a scheduler acting on the module's optimizer object changes the trainer's learning rate
    model.optimizer.param_groups[0]["lr"] = 0.0
    before = trainer.flat.clone()
    trainer.step(_batches(pkg, graphs)[0].to("cuda"))
    assert torch.equal(before, trainer.flat)
    # state_dict round trip keeps the parameters inside the slab
    model.load_state_dict({k: v.cuda() for k, v in sd0.items()})
    assert torch.allclose(trainer.flat[:64 * 25].view(64, 25).cpu(), sd0["conv1.lin.weight"])


def test_predict_ensemble_stream_matches_per_model_golden(pkg):
    """Three shipped RhCAA models over one stream of host batches: every model's predictions equal the
    reference's own golden CSV rows, mean / std are those of the stack."""
    models, golds = [], []
    graphs = None
    for outer, inner in [(1, 2), (8, 10), (10, 9)]:
        sd, gs, _, pred_gold = golden_io.rows_with_graphs("learning", outer, inner)
        m = pkg.make_network("GCN", pkg.default_options(), 25)
        m.load_state_dict(sd, strict=True)
        models.append(m.to("cuda").eval())
        if graphs is None:
            graphs = gs
        golds.append((gs, pred_gold))
    # the golden rows of each model cover different reactions (its own test folds): screen the union of the
    # first model's graphs and compare model 0 to its golden rows, the others to the blocking forward
    gs0, gold0 = golds[0]
    loader = pkg.DataLoader(gs0, batch_size=37, shuffle=False)
    preds, means, stds = [], [], []
    for p, mean, std, h in pkg.predict_ensemble_stream(models, loader):
        assert p.shape == (3, h.num_graphs, 1)
        preds.append(p.clone()); means.append(mean.clone()); stds.append(std.clone())
    pred = torch.cat(preds, 1)
    assert np.abs(pred[0].numpy().ravel() - gold0).max() <= 2e-5
    with torch.no_grad():
        for k, m in enumerate(models):
            ref = torch.cat([m(b.to("cuda")).cpu() for b in pkg.DataLoader(gs0, batch_size=37, shuffle=False)], 0)
            assert torch.equal(ref, pred[k])      # same kernels, same batches: identical bits
    assert torch.allclose(torch.cat(means, 0), pred.mean(0), atol=1e-6)
    assert torch.allclose(torch.cat(stds, 0), pred.std(0, unbiased=False), atol=1e-6)


def test_trainer_epoch_functions_match_reference_loop(pkg):
    """FusedTrainer.train_network / eval_network / fit against the reference's loop bodies
    (utils/utils_model.py:55-79, train_GNN.py:73-111) run through the module's autograd path."""
    graphs = golden_io.load_graphs("rhcaa_learning")
    tr, va = graphs[:200], graphs[200:280]
    opt = pkg.default_options()
    model = pkg.make_network("GCN", opt, 25).to("cuda")
    twin = pkg.make_network("GCN", opt, 25).to("cuda")
    twin.load_state_dict(model.state_dict())
    trainer = pkg.FusedTrainer(model)
    tl, vl = pkg.DataLoader(tr, batch_size=40), pkg.DataLoader(va, batch_size=40)
    for epoch in range(2):
        got_train = trainer.train_network(tl)
        got_val = trainer.eval_network(vl)
        # the reference's functions, verbatim semantics
        twin.train(); ref_train = 0.0
        for batch in tl:
            batch = batch.to("cuda")
            twin.optimizer.zero_grad()
            loss = torch.sqrt(twin.loss(twin(batch), torch.unsqueeze(batch.y, dim=1)))
            loss.backward(); twin.optimizer.step()
            ref_train += loss.item() * batch.num_graphs
        ref_train /= len(tl.dataset)
        twin.eval(); ref_val = 0.0
        for batch in vl:
            batch = batch.to("cuda")
            ref_val += torch.sqrt(twin.loss(twin(batch), torch.unsqueeze(batch.y, dim=1))).item() * batch.num_graphs
        ref_val /= len(vl.dataset)
        assert abs(got_train - ref_train) <= 2e-3 * max(1.0, abs(ref_train))
        assert abs(got_val - ref_val) <= 2e-3 * max(1.0, abs(ref_val))
    hist = trainer.fit(tl, vl, epochs=6, early_stopping=6)
    assert len(hist["val"]) == 2 and hist["best_state_dict"] is not None      # epochs 0 and 5 are evaluated
    assert list(hist["best_state_dict"].keys()) == list(model.state_dict().keys())


def test_gnn_explainer_tracks_oracle_run(pkg):
    """hcatgnet_b200.explain.gnn_explainer (masks trained through the CUDA forward/backward) against the same
    loop driven by the CPU oracle, from identical initial masks."""
    from hcatgnet_b200 import explain
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    g = graphs[3]
    x, ei = g["x"], g["edge_index"]
    batch = torch.zeros(x.shape[0], dtype=torch.int64)
    from hcatgnet_b200.model.gcn import GCN_explain
    model = GCN_explain(pkg.default_options(), 25)
    model.load_state_dict(sd, strict=True)
    model = model.to("cuda")
    gen = torch.Generator().manual_seed(11)
    nm0, em0 = explain.init_masks(x.shape[0], x.shape[1], ei.shape[1], "cpu", gen)
    got = explain.gnn_explainer(model, x, ei, batch, epochs=12,
                                masks=(nm0.detach().clone().cuda().requires_grad_(True),
                                       em0.detach().clone().cuda().requires_grad_(True)))
    ref_nm, ref_em = explain.run_gnnexplainer(
        lambda h, m: ref_gcn.forward(sd, h, ei, batch, 1, edge_mask=m)[0], x, ei.shape[1], epochs=12,
        masks=(nm0.detach().clone().requires_grad_(True), em0.detach().clone().requires_grad_(True)))
    assert got["node_mask"].shape == x.shape and got["edge_mask"].shape == (ei.shape[1],)
    assert float((got["node_mask"].cpu() - ref_nm).abs().max()) <= 2e-3
    assert float((got["edge_mask"].cpu() - ref_em).abs().max()) <= 2e-3
    # what the reference consumes: masks / max(masks.sum(dim=1))  (utils/other_utils.py:66-68)
    nm = got["node_mask"].cpu()
    assert torch.isfinite(nm / torch.max(nm.sum(dim=1))).all()


def test_prefetcher_rebuilds_batch_vector_from_ptr(pkg):
    from hcatgnet_b200 import ops
    from hcatgnet_b200.synth import synth_batch
    h = synth_batch(300, 25, seed=3, n_min=1, n_max=60, pin=True)
    d = next(iter(pkg.DevicePrefetcher([h])))
    torch.cuda.synchronize()
    assert torch.equal(d.batch.cpu(), h.batch) and torch.equal(d.ptr.cpu(), h.ptr)
    ptr = torch.tensor([0, 3, 3, 7], dtype=torch.int64, device="cuda")      # an empty graph in the middle
    assert ops.batch_from_ptr(ptr, 7).tolist() == [0, 0, 0, 2, 2, 2, 2]


# ------------------------------------------------------------------------------------------
# packed graph image on the device: hcg_unpack_batch against PyG-style collate
# ------------------------------------------------------------------------------------------
def _assert_same_batch(dev_b, host_b):
    assert torch.equal(dev_b.x.cpu(), host_b.x) and torch.equal(dev_b.edge_index.cpu(), host_b.edge_index)
    assert torch.equal(dev_b.batch.cpu(), host_b.batch) and torch.equal(dev_b.ptr.cpu(), host_b.ptr)
    assert torch.equal(dev_b.y.cpu(), host_b.y) and dev_b.num_graphs == host_b.num_graphs
    assert torch.equal(dev_b.idx, host_b.idx)


@pytest.mark.gpu
def test_device_graph_store_assembles_pyg_batches_bit_exact():
    import hcatgnet_b200 as pkg
    sd, graphs, _, pred = golden_io.rows_with_graphs("learning", 1, 2)
    ds = pkg.PackedDataset.from_graphs(graphs)
    store = pkg.DeviceGraphStore(ds, "cuda")
    _assert_same_batch(store.batch(), pkg.Batch.from_data_list(graphs))
    gen = np.random.default_rng(3)
    for ids in ([5], [4, 4, 4], list(gen.permutation(len(graphs))[:37]), list(range(len(graphs) - 1, -1, -1))):
        _assert_same_batch(store.batch(ids), pkg.Batch.from_data_list([graphs[i] for i in ids]))
    assert store.batch([]).num_graphs == 0 and store.batch([]).x.shape == (0, ds.F)
    with pytest.raises(IndexError):
        store.batch([len(graphs)])
    # the assembled batch feeds the model like any other: golden predictions of the shipped run
    model = pkg.make_network("GCN", pkg.default_options(), ds.F)
    model.load_state_dict(sd, strict=True)
    model.cuda().eval()
    with torch.no_grad():
        out = model(store.batch())
    assert np.abs(out.cpu().numpy().ravel() - pred).max() <= 2e-5


@pytest.mark.gpu
def test_device_graph_store_large_and_wire_format():
    """A 3000-graph synthetic image: shuffled selections against the host decoder (u8 endpoints, one f32
    column), u16 endpoints, and the wire path (pinned image -> one copy -> unpack) through predict_stream."""
    import hcatgnet_b200 as pkg
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(3000, n_features=25, seed=77)
    x = torch.floor(b.x * 3)            # integer-valued columns, as the featuriser's one-hot blocks are
    x[:, -1] = b.x[:, -1]               # ... and one real-valued column (temperature)
    ds = pkg.PackedDataset.from_arrays(x.numpy(), b.edge_index.numpy(), b.ptr.numpy(), b.y.numpy(), pinned=True)
    assert ds.Fu == 24 and ds.Ff == 1 and ds.edge_kind == 1
    store = pkg.DeviceGraphStore(ds, "cuda")
    ids = np.random.default_rng(1).permutation(3000)[:1111]
    _assert_same_batch(store.batch(ids), ds.unpack_host(ids))
    big = synth_batch(3, n_features=7, seed=5, n_min=250, n_max=400)
    ds16 = pkg.PackedDataset.from_arrays(big.x.numpy(), big.edge_index.numpy(), big.ptr.numpy(), big.y.numpy())
    assert ds16.edge_kind == 2 and ds16.Fu == 0
    _assert_same_batch(pkg.DeviceGraphStore(ds16, "cuda").batch([2, 0]), ds16.unpack_host([2, 0]))
    # bit columns (one-hot / flag columns of the featuriser: one bit per node and column), mixed with byte and f32 columns
    fb = synth_batch(500, n_features=25, seed=78, features="featuriser")
    xm = fb.x.clone()
    xm[:, 3] *= 5.0                     # 0 / 5: a byte column between the bit columns
    dsb = pkg.PackedDataset.from_arrays(xm.numpy(), fb.edge_index.numpy(), fb.ptr.numpy(), fb.y.numpy())
    assert (dsb.Fb, dsb.Fu, dsb.Ff) == (23, 1, 1)
    sel = np.random.default_rng(2).permutation(500)[:123]
    got = pkg.DeviceGraphStore(dsb, "cuda").batch(sel)
    _assert_same_batch(got, dsb.unpack_host(sel))
    assert torch.equal(got.x.cpu(), torch.cat([xm[fb.ptr[g]:fb.ptr[g + 1]] for g in sel]))
    # wire format: the same predictions as the f32 / i64 host batch
    opt = pkg.default_options()
    torch.manual_seed(0)
    model = pkg.make_network("GCN", opt, 25).cuda().eval()
    host_b = ds.unpack_host()
    want = [o.clone() for o, _, _ in pkg.predict_stream(model, [host_b])]
    got = [o.clone() for o, _, _ in pkg.predict_stream(model, [ds, ds])]
    assert len(got) == 2 and torch.equal(got[0], want[0]) and torch.equal(got[1], want[0])


@pytest.mark.gpu
def test_store_loaders_match_host_loaders_in_nested_cv():
    """create_loaders over a DeviceGraphStore: the same graphs in the same order as the host loaders of the
    reference's generator, an identical training epoch, and cached evaluation batches."""
    import random
    import hcatgnet_b200 as pkg
    graphs = golden_io.load_graphs("rhcaa_learning")[:200]
    opt = pkg.default_options()
    store = pkg.DeviceGraphStore(pkg.PackedDataset.from_graphs(graphs), "cuda")
    host_it, dev_it = pkg.create_loaders(graphs, opt), pkg.create_loaders(store, opt)
    for n in range(12):
        (htr, hva, hte), (dtr, dva, dte) = next(host_it), next(dev_it)
        for hl, dl in ((htr, dtr), (hva, dva), (hte, dte)):
            assert [int(g["idx"]) for g in hl.dataset] == [int(store.ds.idx[i]) for i in dl.dataset]
            assert len(hl) == len(dl) and len(hl.dataset) == len(dl.dataset)
    losses = []
    for loaders in ((htr, hva), (dtr, dva)):
        torch.manual_seed(1)
        model = pkg.make_network("GCN", opt, 25).cuda()
        tr = pkg.FusedTrainer(model)
        random.seed(5)
        l_train = tr.train_network(loaders[0])
        l_val = tr.eval_network(loaders[1])
        l_val2 = tr.eval_network(loaders[1])
        assert l_val == l_val2
        losses.append((l_train, l_val))
    assert losses[0] == losses[1]          # same batches, same kernels, same order: identical numbers
    assert dva._kept is not None and dtr._kept is None
    first = [id(b) for b in dva]
    assert first == [id(b) for b in dva]


def test_explain_model_on_two_different_graphs_back_to_back(pkg):
    """Two molecules of equal (N, E) through ONE GCN_explain, as the reference's explain loop feeds them
    (scripts_experiments/explain_gnn.py:39-57): each call must use its own graph's topology (the plan cache is
    keyed on the tensors' identity, not on their addresses, which the allocator reuses)."""
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    model = pkg.GCN_explain(pkg.default_options(), 25)
    model.load_state_dict(sd, strict=True)
    model = model.to("cuda").eval()
    by_shape = {}
    for g in graphs:
        by_shape.setdefault((g["x"].shape[0], g["edge_index"].shape[1]), []).append(g)
    pairs = [v for v in by_shape.values() if len(v) >= 2 and not torch.equal(v[0]["edge_index"], v[1]["edge_index"])]
    assert pairs, "the fixture holds same-shape molecules with different bonds"
    checked = 0
    for group in pairs[:6]:
        for g in group[:3]:
            # fresh device tensors per call, freed right after: the next call's tensors land at the same addresses
            with torch.no_grad():
                out = model(x=g["x"].cuda(), edge_index=g["edge_index"].cuda(), batch_index=None).cpu()
            ref_sd = {k: v for k, v in sd.items()}
            o_ref, _ = ref_gcn.forward(ref_sd, g["x"], g["edge_index"], torch.zeros(g["x"].shape[0], dtype=torch.int64), 1)
            assert float((out - o_ref).abs().max()) <= 1e-5 * max(1.0, float(o_ref.abs().max()))
            checked += 1
    assert checked >= 4


def test_report_csvs_reproduce_the_reference_files(pkg, tmp_path):
    """embeddings.csv / predictions_test_set.csv written through the CUDA path: same columns as the reference's
    network_report (utils/utils_model.py:199-205) and the same numbers as the shipped files (the golden fixture IS the
    reference's embeddings.csv of this run), row for row."""
    import pandas as pd
    sd, rows = golden_io.load_model("final_test", 1, 2)
    model = pkg.make_network("GCN", pkg.default_options(), 25)
    model.load_state_dict(sd, strict=True)
    model = model.to("cuda").eval()
    learn, final = golden_io.graphs_by_idx("rhcaa_learning"), golden_io.graphs_by_idx("rhcaa_final_test")
    sets = {"training": [], "val": [], "test": []}
    for r in range(len(rows["index"])):
        name = str(rows["set"][r])
        sets[name].append((final if name == "test" else learn)[int(rows["index"][r])])
    loaders = [pkg.DataLoader(sets[k], batch_size=40) for k in ("training", "val", "test")]
    pkg.write_report_csvs(str(tmp_path), model, loaders)
    emb = pd.read_csv(tmp_path / "embeddings.csv", index_col=0)
    assert list(emb.columns) == [str(i) for i in range(128)] + ["ddG_exp", "ddG_pred", "index", "set"]
    assert len(emb) == len(rows["index"]) and list(emb["set"]) == [str(v) for v in rows["set"]]
    assert np.array_equal(emb["index"].to_numpy(), rows["index"])
    assert np.abs(emb[[str(i) for i in range(128)]].to_numpy() - rows["emb"]).max() <= 1e-5 * max(1.0, np.abs(rows["emb"]).max())
    assert np.abs(emb["ddG_pred"].to_numpy() - rows["ddG_pred"]).max() <= 2e-5
    pred = pd.read_csv(tmp_path / "predictions_test_set.csv", index_col=0)
    assert list(pred.columns) == ["real_ddG", "predicted_ddG", "index"] and len(pred) == len(sets["test"])
    ens = pkg.screen_ensemble_to_csv(str(tmp_path / "screen"), [model, model], pkg.DataLoader(sets["test"], batch_size=16))
    assert np.allclose(ens["mean_predicted_ddG"].to_numpy(), pred["predicted_ddG"].to_numpy(), atol=1e-6)
    assert float(ens["std_predicted_ddG"].abs().max()) == 0.0


def test_gradient_wrt_edge_weight_matches_oracle(pkg):
    """GCN_explain.forward(x, edge_index, batch_index, edge_weight) with edge_weight.requires_grad (reference
    model/gcn.py:124-127: the weights reach conv1 only, through its weighted normalisation): d out / d edge_weight and
    d out / d x against the fp64 oracle's autograd, with and without `improved`, including an existing self-loop."""
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    b = ref_gcn.collate(graphs[:7])
    ei = torch.cat([b["edge_index"], torch.tensor([[3], [3]])], dim=1)       # one self-loop edge with its own weight
    g = torch.Generator().manual_seed(4)
    ew0 = torch.rand(ei.shape[1], generator=g) + 0.25
    for improved in (True, False):
        opt = pkg.default_options(improved=improved)
        model = pkg.GCN_explain(opt, 25)
        model.load_state_dict(sd, strict=True)
        model = model.to("cuda").eval()
        ew = ew0.clone().cuda().requires_grad_(True)
        x = b["x"].clone().cuda().requires_grad_(True)
        out = model(x=x, edge_index=ei.cuda(), batch_index=b["batch"].cuda(), edge_weight=ew)
        out.sum().backward()
        sd64 = {k: v.double() for k, v in sd.items()}
        ew_r = ew0.double().requires_grad_(True)
        x_r = b["x"].double().requires_grad_(True)
        o_ref, _ = ref_gcn.forward(sd64, x_r, ei, b["batch"], 7, edge_weight=ew_r, improved=improved)
        o_ref.sum().backward()
        assert float((out.detach().cpu().double() - o_ref.detach()).abs().max()) <= 1e-5 * max(1.0, float(o_ref.abs().max()))
        for got, ref in ((ew.grad, ew_r.grad), (x.grad, x_r.grad)):
            scale = max(float(ref.abs().max()), 1e-12)
            assert float((got.cpu().double() - ref).abs().max()) <= 1e-4 * scale
