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
        a = model(batch)
    plan = batch._hcg_plan[1]
    with torch.no_grad():
        b = model(batch)
    assert batch._hcg_plan[1] is plan and torch.equal(a, b)
    c = model(batch)  # grad mode on (reference eval_network has no no_grad): plan upgraded with transpose
    assert batch._hcg_plan[1].has_transpose and torch.equal(a, c.detach())


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
