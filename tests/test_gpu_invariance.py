"""Property tests (hypothesis) of the symmetries the GCN path has by construction:
the prediction of a reaction graph does not depend on the order of its edge list, on the numbering of its atoms,
or on where the graph sits in the batch.  Edge / atom order change the floating-point summation order, so those
hold to the path's tolerance; the position in the batch changes nothing a graph's arithmetic sees, so that one
holds bit for bit.  Integer outputs of hcg_collate (degrees, row lengths) are order-independent exactly.
"""
import numpy as np
import pytest
import torch
from hypothesis import given, settings, strategies as st

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import hcatgnet_b200 as pkg
    from tests import golden_io
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    model = pkg.make_network("GCN", pkg.default_options(), 25)
    model.load_state_dict(sd, strict=True)
    return pkg, model.to("cuda").eval(), graphs


def _predict(pkg, model, graphs):
    with torch.no_grad():
        out, emb = model(pkg.Batch.from_data_list(graphs).to("cuda"), True)
    return out.cpu(), emb.cpu()


def _close(a, b):
    return float((a - b).abs().max()) <= 1e-5 * max(1.0, float(b.abs().max()))


@settings(max_examples=8, deadline=None)
@given(seed=st.integers(0, 2**31 - 1), count=st.integers(1, 60))
def test_graph_order_is_bit_exact(ctx, seed, count):
    pkg, model, graphs = ctx
    rng = np.random.default_rng(seed)
    pick = [graphs[i] for i in rng.choice(len(graphs), size=count, replace=False)]
    perm = rng.permutation(count)
    o1, e1 = _predict(pkg, model, pick)
    o2, e2 = _predict(pkg, model, [pick[i] for i in perm])
    assert torch.equal(o1[perm], o2) and torch.equal(e1[perm], e2)


@settings(max_examples=8, deadline=None)
@given(seed=st.integers(0, 2**31 - 1))
def test_edge_order_and_atom_numbering(ctx, seed):
    pkg, model, graphs = ctx
    from hcatgnet_b200 import ops
    rng = np.random.default_rng(seed)
    pick = [graphs[i] for i in rng.choice(len(graphs), size=24, replace=False)]
    shuffled, renumbered = [], []
    for g in pick:
        x, ei = g["x"], g["edge_index"]
        pe = torch.from_numpy(rng.permutation(ei.shape[1]))
        shuffled.append({**g, "edge_index": ei[:, pe]})
        pn = torch.from_numpy(rng.permutation(x.shape[0]))          # new number of atom i = pn[i]
        x2 = torch.empty_like(x); x2[pn] = x
        renumbered.append({**g, "x": x2, "edge_index": pn[ei]})
    o0, e0 = _predict(pkg, model, pick)
    for variant in (shuffled, renumbered):
        o, e = _predict(pkg, model, variant)
        assert _close(o, o0) and _close(e, e0)
    # integer side: same degrees (row lengths) whatever the edge order; dis is a function of the degree only
    b0, b1 = pkg.Batch.from_data_list(pick).to("cuda"), pkg.Batch.from_data_list(shuffled).to("cuda")
    p0 = ops.collate(b0.edge_index, b0.batch, b0.num_nodes, b0.num_graphs)
    p1 = ops.collate(b1.edge_index, b1.batch, b1.num_nodes, b1.num_graphs)
    assert torch.equal(p0.rowptr, p1.rowptr) and torch.equal(p0.dis, p1.dis) and torch.equal(p0.graph_ptr, p1.graph_ptr)
    s0 = torch.stack([p0.col[p0.rowptr[i]:p0.rowptr[i + 1]].sort().values.sum() for i in range(0, b0.num_nodes, 7)])
    s1 = torch.stack([p1.col[p1.rowptr[i]:p1.rowptr[i + 1]].sort().values.sum() for i in range(0, b0.num_nodes, 7)])
    assert torch.equal(s0, s1)
