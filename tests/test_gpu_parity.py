"""GPU parity tests proper: the CUDA path (through the C-ABI of include/hcg.h) against the oracle.

Tolerances: integer arrays bit-exact; fp32 outputs
``abs(Δ) <= 1e-5 * max(abs(ref), 1)``; gradients against the fp64 oracle's autograd at
rtol 1e-4 of the tensor's max magnitude.
"""
import numpy as np
import pytest
import torch

from oracle import ref_collate, ref_gcn
from tests import golden_io

pytestmark = pytest.mark.gpu

TOL = 1e-5


def _close(got, ref, tol=TOL):
    got, ref = np.asarray(got, dtype=np.float64), np.asarray(ref, dtype=np.float64)
    err = np.abs(got - ref) / np.maximum(np.abs(ref), 1.0)
    return float(err.max()) if err.size else 0.0


@pytest.fixture(scope="module")
def hcg():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from hcatgnet_b200 import _lib, ops
    _lib.load()
    return ops


def _sd_params(sd, device):
    n_conv, n_ro = ref_gcn.infer_arch(sd)
    ps = [sd["conv1.lin.weight"], sd["conv1.bias"]]
    for i in range(n_conv - 1):
        ps += [sd[f"conv_layers.{i}.lin.weight"], sd[f"conv_layers.{i}.bias"]]
    for k in range(n_ro - 1):
        ps += [sd[f"readout.{k}.0.weight"], sd[f"readout.{k}.0.bias"]]
    ps += [sd[f"readout.{n_ro - 1}.weight"], sd[f"readout.{n_ro - 1}.bias"]]
    return [p.to(device).contiguous() for p in ps], n_conv, n_ro


def _dims(ops, sd):
    n_conv, n_ro = ref_gcn.infer_arch(sd)
    H, F = sd["conv1.lin.weight"].shape
    C = sd[f"readout.{n_ro - 1}.weight"].shape[0]
    return ops.NetDims(F=F, H=H, L=n_conv, R=n_ro, C=C)


# ------------------------------------------------------------------------------------------
# collate: bit-exact integers
# ------------------------------------------------------------------------------------------
def _check_collate(ops, edge_index, batch, N, B):
    plan = ops.collate(edge_index.cuda(), None if batch is None else batch.cuda(), N, B, transpose=True)
    ei = edge_index.numpy()
    rowptr, col, eid, deg = ref_collate.csr_by_dst(ei, N)
    E_kept = int(rowptr[-1])
    assert np.array_equal(plan.rowptr.cpu().numpy(), rowptr)
    assert np.array_equal(plan.col.cpu().numpy()[:E_kept], col)
    assert np.array_equal(plan.eid.cpu().numpy()[:E_kept], eid)
    assert np.array_equal(plan.dis.cpu().numpy()[:N], ref_collate.dis_from_deg(deg, 1.0))
    t_rowptr, t_col, t_eid, _ = ref_collate.csr_by_dst(ei, N, by="src")
    assert np.array_equal(plan.t_rowptr.cpu().numpy(), t_rowptr)
    assert np.array_equal(plan.t_col.cpu().numpy()[:E_kept], t_col)
    assert np.array_equal(plan.t_eid.cpu().numpy()[:E_kept], t_eid)
    if batch is not None:
        assert np.array_equal(plan.graph_ptr.cpu().numpy(), ref_collate.graph_ptr(batch.numpy(), B))
    return plan



def test_collate_grouped_fast_path_same_bits(hcg):
    """HCG_COLLATE_GROUPED (graph-local CSR build) against the general kernels and the oracle: identical
    arrays on grouped batches; ungrouped edge order, self-loops and over-sized graphs are refused by
    the device check and transparently re-collated by the general path."""
    from hcatgnet_b200 import _lib
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(700, 25, seed=4, n_min=1, n_max=200)
    ei, bt = b.edge_index.cuda(), b.batch.cuda()
    fast = hcg.collate(ei, bt, b.num_nodes, b.num_graphs, transpose=True, validate=False, grouped=True)
    gen = hcg.collate(ei, bt, b.num_nodes, b.num_graphs, transpose=True, grouped=False)
    for name in ("rowptr", "col", "eid", "dis", "graph_ptr", "t_rowptr", "t_col", "t_eid"):
        assert torch.equal(getattr(fast, name), getattr(gen, name)), name
    _check_collate(hcg, b.edge_index, b.batch, b.num_nodes, b.num_graphs)        # auto mode, vs the oracle
    # ungrouped: a random permutation of the edge columns
    perm = torch.randperm(b.edge_index.shape[1], generator=torch.Generator().manual_seed(0))
    _check_collate(hcg, b.edge_index[:, perm], b.batch, b.num_nodes, b.num_graphs)
    # one graph above the shared-memory capacity (1024 nodes): refused, then served by the general path
    big = synth_batch(3, 25, seed=5, n_min=1500, n_max=1600)
    _check_collate(hcg, big.edge_index, big.batch, big.num_nodes, big.num_graphs)
    # the refusal itself, seen through the C ABI: status word -> HCG_ERR_NOT_GROUPED
    p2 = hcg.collate(b.edge_index[:, perm].cuda(), bt, b.num_nodes, b.num_graphs, validate=False, grouped=True)
    assert p2.rowptr.numel() == b.num_nodes + 1                                   # memory-safe garbage, not checked further

@pytest.mark.parametrize("dataset,count", [("rhcaa_learning", 40), ("biAryl", 485), ("rhcaa_final_test", 52)])
def test_collate_bit_exact_on_reference_graphs(hcg, dataset, count):
    b = ref_gcn.collate(golden_io.load_graphs(dataset)[:count])
    _check_collate(hcg, b["edge_index"], b["batch"], b["x"].shape[0], b["num_graphs"])


def test_collate_edge_cases(hcg):
    # pre-existing self-loops, duplicate edges, isolated nodes, a single-node graph, one hub row
    ei = torch.tensor([[0, 1, 1, 2, 2, 2, 5, 5, 0, 3, 3, 6, 7, 8, 9, 10, 11, 12, 13],
                       [1, 0, 1, 2, 0, 0, 5, 6, 2, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0]], dtype=torch.int64)
    batch = torch.tensor([0, 0, 0, 0, 1, 2, 2, 2, 2, 2, 2, 2, 2, 2, 2], dtype=torch.int64)
    N, B = 15, 4  # graph 3 is empty (never produced by collate, but must stay memory-safe)
    _check_collate(hcg, ei, batch, N, B)
    # no edges at all
    _check_collate(hcg, torch.zeros(2, 0, dtype=torch.int64), torch.zeros(5, dtype=torch.int64), 5, 1)
    # batch=None means one graph
    plan = hcg.collate(ei.cuda(), None, N, 1)
    assert plan.graph_ptr.cpu().tolist() == [0, N]


def test_collate_large_synthetic_bit_exact(hcg):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(2048, seed=7)
    _check_collate(hcg, b.edge_index, b.batch, b.num_nodes, b.num_graphs)


def test_collate_reports_bad_input(hcg):
    from hcatgnet_b200._lib import HcgError
    ei = torch.tensor([[0, 1], [1, 9]], dtype=torch.int64).cuda()
    with pytest.raises(HcgError) as e:
        hcg.collate(ei, torch.zeros(3, dtype=torch.int64).cuda(), 3, 1)
    assert e.value.code == 5
    with pytest.raises(HcgError) as e:
        hcg.collate(torch.tensor([[0], [1]]).cuda(), torch.tensor([1, 0, 0]).cuda(), 3, 2)
    assert e.value.code == 6


# ------------------------------------------------------------------------------------------
# forward: against the oracle AND directly against the reference's golden vectors
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rset,outer,inner", golden_io.MODELS)
def test_forward_matches_oracle_and_golden(hcg, rset, outer, inner):
    sd, graphs, emb_gold, pred_gold = golden_io.rows_with_graphs(rset, outer, inner)
    params, _, _ = _sd_params(sd, "cuda")
    dims = _dims(hcg, sd)
    outs, embs, outs_ref, embs_ref = [], [], [], []
    for s in range(0, len(graphs), 40):
        b = ref_gcn.collate(graphs[s:s + 40])
        plan = hcg.collate(b["edge_index"].cuda(), b["batch"].cuda(), b["x"].shape[0], b["num_graphs"])
        with torch.no_grad():
            o, e = hcg.gcn_apply(b["x"].cuda(), plan, dims, params)
        o_ref, e_ref = ref_gcn.forward(sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"])
        outs.append(o.cpu()); embs.append(e.cpu()); outs_ref.append(o_ref); embs_ref.append(e_ref)
    out, emb = torch.cat(outs).numpy().ravel(), torch.cat(embs).numpy()
    out_ref, emb_ref = torch.cat(outs_ref).numpy().ravel(), torch.cat(embs_ref).numpy()
    assert _close(emb, emb_ref) <= TOL and _close(out, out_ref) <= TOL
    # the reference's own outputs (embeddings.csv): same bounds the oracle itself is pinned with
    assert _close(emb, emb_gold) <= TOL
    assert np.abs(out - pred_gold).max() <= 2e-5
    match4 = np.mean([f"{a:.4f}" == f"{b:.4f}" for a, b in zip(out, pred_gold)])
    # 4-decimal string identity with the reference CSV: no re-implementation reaches 100 % (values on a
    # rounding boundary flip with any change of summation order; CPU-vs-CPU already differs in 1-2 % of
    # rows).  Observed: 97-98 % FMA kernels, ~97 % tensor-core kernel.
    assert match4 >= 0.95, match4


@pytest.mark.parametrize("H,L,R,F,C", [(64, 2, 2, 25, 1), (64, 1, 1, 32, 1), (256, 4, 2, 25, 1), (48, 3, 3, 7, 3), (30, 2, 2, 5, 2)])
def test_forward_other_architectures(hcg, H, L, R, F, C):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(37, n_features=F, seed=3, n_min=1, n_max=60)
    sd = ref_gcn.init_state_dict(F, H, L, R, C, seed=5)
    for k in sd:  # non-zero conv biases
        if k.endswith("bias"):
            sd[k] = torch.randn_like(sd[k]) * 0.1
    params, _, _ = _sd_params(sd, "cuda")
    plan = hcg.collate(b.edge_index.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs)
    with torch.no_grad():
        o, e = hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
    o_ref, e_ref = ref_gcn.forward(sd, b.x, b.edge_index, b.batch, b.num_graphs)
    assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL


def test_forward_deterministic_run_to_run(hcg):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(512, seed=11)
    sd = ref_gcn.init_state_dict(25)
    params, _, _ = _sd_params(sd, "cuda")
    res = []
    for _ in range(3):
        plan = hcg.collate(b.edge_index.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs)
        with torch.no_grad():
            o, e = hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
        res.append((o.cpu(), e.cpu()))
    for o, e in res[1:]:
        assert torch.equal(o, res[0][0]) and torch.equal(e, res[0][1])


def test_forward_edge_weight_conv1_only(hcg):
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    b = ref_gcn.collate(graphs[:5])
    E = b["edge_index"].shape[1]
    ew = torch.rand(E, generator=torch.Generator().manual_seed(1)) + 0.1
    # add a pre-existing self-loop with its own weight
    ei = torch.cat([b["edge_index"], torch.tensor([[3], [3]])], 1)
    ew = torch.cat([ew, torch.tensor([0.37])])
    params, _, _ = _sd_params(sd, "cuda")
    for improved in (True, False):
        plan = hcg.collate(ei.cuda(), b["batch"].cuda(), b["x"].shape[0], 5, edge_weight=ew.cuda(), improved=improved)
        with torch.no_grad():
            o, e = hcg.gcn_apply(b["x"].cuda(), plan, _dims(hcg, sd), params)
        o_ref, e_ref = ref_gcn.forward(sd, b["x"], ei, b["batch"], 5, edge_weight=ew, improved=improved)
        assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL


# ------------------------------------------------------------------------------------------
# backward: against the fp64 oracle's autograd
# ------------------------------------------------------------------------------------------
def _oracle_grads(sd, x, ei, batch, B, y, edge_weight=None, improved=True, x_grad=False):
    sd64 = {k: v.double().requires_grad_(True) for k, v in sd.items()}
    x64 = x.double().requires_grad_(x_grad)
    out, emb = ref_gcn.forward(sd64, x64, ei, batch, B, edge_weight=None if edge_weight is None else edge_weight.double(),
                               improved=improved)
    loss = ref_gcn.rmse_loss(out, y.double()) + 0.01 * emb.sum()
    loss.backward()
    return loss.item(), {k: v.grad for k, v in sd64.items()}, x64.grad


def _cuda_grads(hcg, sd, x, ei, batch, B, y, edge_weight=None, improved=True, x_grad=False):
    keys = ["conv1.lin.weight", "conv1.bias"]
    n_conv, n_ro = ref_gcn.infer_arch(sd)
    for i in range(n_conv - 1):
        keys += [f"conv_layers.{i}.lin.weight", f"conv_layers.{i}.bias"]
    for k in range(n_ro - 1):
        keys += [f"readout.{k}.0.weight", f"readout.{k}.0.bias"]
    keys += [f"readout.{n_ro - 1}.weight", f"readout.{n_ro - 1}.bias"]
    params = [sd[k].cuda().requires_grad_(True) for k in keys]
    xc = x.cuda().requires_grad_(x_grad)
    plan = hcg.collate(ei.cuda(), batch.cuda(), x.shape[0], B, transpose=True,
                       edge_weight=None if edge_weight is None else edge_weight.cuda(), improved=improved)
    out, emb = hcg.gcn_apply(xc, plan, _dims(hcg, sd), params)
    loss = torch.sqrt(torch.nn.functional.mse_loss(out, y.cuda().unsqueeze(1))) + 0.01 * emb.sum()
    loss.backward()
    return loss.item(), {k: p.grad.cpu() for k, p in zip(keys, params)}, None if not x_grad else xc.grad.cpu()


def _assert_grads(g_cuda, g_ref, rtol=1e-4):
    for k in g_ref:
        ref = g_ref[k].numpy()
        scale = max(np.abs(ref).max(), 1e-6)
        err = np.abs(g_cuda[k].numpy().astype(np.float64) - ref).max() / scale
        assert err <= rtol, (k, err)


@pytest.mark.parametrize("rset,outer,inner,count", [("learning", 1, 2, 40), ("biaryl", 1, 2, 64)])
def test_backward_matches_oracle_autograd(hcg, rset, outer, inner, count):
    sd, graphs, _, _ = golden_io.rows_with_graphs(rset, outer, inner)
    b = ref_gcn.collate(graphs[:count])
    args = (sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"], b["y"])
    l_ref, g_ref, _ = _oracle_grads(*args)
    l_cuda, g_cuda, _ = _cuda_grads(hcg, *args)
    assert abs(l_cuda - l_ref) <= 1e-5 * max(abs(l_ref), 1)
    _assert_grads(g_cuda, g_ref)
    # d/dx: symmetry-equivalent atoms of the real molecules are (near-)ties of the max pooling;
    # which of them carries the max gradient is decided by rounding noise and differs between the
    # fp64 oracle and any fp32 run (weight gradients are unaffected: the tied atoms are
    # equivalent).  The node-level gradient is therefore compared on de-symmetrised inputs;
    # exact ties are covered by test_backward_exact_ties_split_evenly.
    xj = b["x"] + 0.05 * torch.rand(b["x"].shape, generator=torch.Generator().manual_seed(3))
    args = (sd, xj, b["edge_index"], b["batch"], b["num_graphs"], b["y"])
    _, g_ref, gx_ref = _oracle_grads(*args, x_grad=True)
    _, g_cuda, gx_cuda = _cuda_grads(hcg, *args, x_grad=True)
    _assert_grads(g_cuda, g_ref)
    # per-node gradients are single fp32 chains (no averaging over nodes as in weight gradients):
    # 1e-3 of the largest entry
    _assert_grads({"x": gx_cuda}, {"x": gx_ref}, rtol=1e-3)


def test_backward_exact_ties_split_evenly(hcg):
    """Two identical copies of a molecule inside ONE graph (same edge order): every pooled maximum
    is attained twice, bit-identically.  torch's scatter_reduce(amax) backward — what the
    restated reference path runs — shares the gradient evenly among ties; so must the kernel."""
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 1, 2)
    gs = []
    for g in graphs[:6]:
        n = g["x"].shape[0]
        gs.append({"x": torch.cat([g["x"], g["x"]]), "y": g["y"], "idx": g["idx"],
                   "edge_index": torch.cat([g["edge_index"], g["edge_index"] + n], 1)})
    b = ref_gcn.collate(gs)
    args = (sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"], b["y"])
    _, g_ref, gx_ref = _oracle_grads(*args, x_grad=True)
    _, g_cuda, gx_cuda = _cuda_grads(hcg, *args, x_grad=True)
    _assert_grads(g_cuda, g_ref)
    n0 = gs[0]["x"].shape[0] // 2
    assert torch.equal(gx_cuda[:n0], gx_cuda[n0:2 * n0])      # the two copies get identical gradients
    # vs the fp64 oracle the copies are exact ties too, but ties BETWEEN symmetry-equivalent atoms of
    # one copy stay rounding-dependent: compare the per-copy SUM over equivalent candidates instead,
    # i.e. the total gradient mass per graph and feature
    ptr = b["ptr"].tolist()
    for g in range(len(gs)):
        a, r = gx_cuda[ptr[g]:ptr[g + 1]].double().sum(0), gx_ref[ptr[g]:ptr[g + 1]].sum(0)
        assert float((a - r).abs().max()) <= 1e-4 * max(float(r.abs().max()), 1e-6)


@pytest.mark.parametrize("H,L,R,F,C", [(256, 4, 2, 25, 1), (48, 3, 3, 7, 3), (64, 1, 1, 32, 1)])
def test_backward_other_architectures(hcg, H, L, R, F, C):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(33, n_features=F, seed=4, n_min=1, n_max=50)
    sd = ref_gcn.init_state_dict(F, H, L, R, C, seed=6)
    y = b.y if C == 1 else None
    if C != 1:  # rmse_loss broadcasts y[:,None] against [B,C]; fine for a gradient check
        y = b.y
    args = (sd, b.x, b.edge_index, b.batch, b.num_graphs, y)
    _, g_ref, _ = _oracle_grads(*args)
    _, g_cuda, _ = _cuda_grads(hcg, *args)
    _assert_grads(g_cuda, g_ref)


def test_backward_with_edge_weight(hcg):
    sd, graphs, _, _ = golden_io.rows_with_graphs("learning", 8, 10)
    b = ref_gcn.collate(graphs[:7])
    ew = torch.rand(b["edge_index"].shape[1], generator=torch.Generator().manual_seed(2)) + 0.2
    args = (sd, b["x"], b["edge_index"], b["batch"], 7, b["y"])
    _, g_ref, gx_ref = _oracle_grads(*args, edge_weight=ew, x_grad=True)
    _, g_cuda, gx_cuda = _cuda_grads(hcg, *args, edge_weight=ew, x_grad=True)
    _assert_grads(g_cuda, g_ref)
    _assert_grads({"x": gx_cuda}, {"x": gx_ref})


# ------------------------------------------------------------------------------------------
# the two CUDA forward paths (graph-resident kernel / layer-wise kernels) against the oracle
# ------------------------------------------------------------------------------------------
@pytest.fixture
def forward_path(hcg):
    from hcatgnet_b200 import _lib
    prev = []

    def set_path(v):
        prev.append(_lib.config_set(_lib.HCG_CFG_FORWARD_PATH, v))

    yield set_path
    if prev:
        _lib.config_set(_lib.HCG_CFG_FORWARD_PATH, prev[0])


@pytest.mark.parametrize("path", [1, 2, 3])
@pytest.mark.parametrize("rset,outer,inner,hint", [("learning", 1, 2, 0), ("biaryl", 1, 2, 0), ("biaryl", 5, 6, 184)])
def test_forward_paths_match_golden(hcg, forward_path, path, rset, outer, inner, hint):
    """biaryl has graphs of up to 184 atoms: without a hint they exceed the 128-node shared-memory
    capacity and run on global scratch inside the same kernel; with the hint they stay on chip."""
    forward_path(path)
    sd, graphs, emb_gold, pred_gold = golden_io.rows_with_graphs(rset, outer, inner)
    params, _, _ = _sd_params(sd, "cuda")
    dims = _dims(hcg, sd)
    b = ref_gcn.collate(graphs)
    plan = hcg.collate(b["edge_index"].cuda(), b["batch"].cuda(), b["x"].shape[0], b["num_graphs"],
                       max_graph_nodes=hint)
    with torch.no_grad():
        o, e = hcg.gcn_apply(b["x"].cuda(), plan, dims, params)
    o_ref, e_ref = ref_gcn.forward(sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"])
    assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL
    # directly against the reference's own outputs: the north-star bound (1e-5 relative, floor 1)
    assert _close(e.cpu().numpy(), emb_gold) <= TOL
    assert np.abs(o.cpu().numpy().ravel() - pred_gold).max() <= 2e-5


@pytest.mark.parametrize("path", [2, 3])
@pytest.mark.parametrize("L,F", [(1, 25), (2, 32), (3, 25), (4, 40), (2, 64), (2, 3), (2, 8), (3, 30)])
def test_graph_resident_forward_shapes(hcg, forward_path, path, L, F):
    """H = 64 networks of 1..4 conv layers, odd feature widths, tiny (1-node) and edge-less graphs,
    pre-existing self-loops and duplicate edges; saved activations (training forward) included."""
    from hcatgnet_b200.synth import synth_batch
    if path == 3 and (F > 32 or L > 3 or (L == 3 and F > 28)):
        pytest.skip("outside the tensor-core kernel's shared-memory budget (the dispatcher picks the FMA kernel)")
    forward_path(path)
    b = synth_batch(61, n_features=F, seed=9 + L, n_min=1, n_max=140)
    ei = torch.cat([b.edge_index, torch.tensor([[0, 5, 5], [0, 6, 6]])], 1)   # self-loop + duplicates
    sd = ref_gcn.init_state_dict(F, 64, L, 2, 1, seed=8)
    for k in sd:
        if k.endswith("bias"):
            sd[k] = torch.randn_like(sd[k]) * 0.1
    params, _, _ = _sd_params(sd, "cuda")
    dims = _dims(hcg, sd)
    plan = hcg.collate(ei.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs, transpose=True)
    o_ref, e_ref, acts_ref = ref_gcn.forward(sd, b.x, ei, b.batch, b.num_graphs, return_all=True)
    with torch.no_grad():
        o, e = hcg.gcn_apply(b.x.cuda(), plan, dims, params)
    assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL
    out, emb, saved = hcg.forward_raw(plan, b.x.cuda(), dims, params, save=True)
    for a_cuda, a_ref in zip(saved[0], acts_ref):
        assert _close(a_cuda.cpu(), a_ref) <= TOL
    assert torch.equal(emb, e) and torch.equal(out, o)          # same kernel, same bits


def test_cross_graph_edges(hcg, forward_path):
    """An edge between two graphs cannot come out of PyG collate.  hcg_collate_status reports it and
    the dispatcher then uses the layer-wise kernels (result = oracle); without validation the
    graph-resident kernel must stay memory-safe on it (the entry is skipped)."""
    from hcatgnet_b200 import _lib
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(8, seed=2)
    ei = torch.cat([b.edge_index, torch.tensor([[0], [b.num_nodes - 1]])], 1)
    sd = ref_gcn.init_state_dict(25)
    params, _, _ = _sd_params(sd, "cuda")
    plan = hcg.collate(ei.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs)
    assert plan.flags & _lib.HCG_GRAPH_CROSS_EDGES
    with torch.no_grad():
        o, e = hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
    o_ref, e_ref = ref_gcn.forward(sd, b.x, ei, b.batch, b.num_graphs)
    assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL
    forward_path(2)
    with pytest.raises(_lib.HcgError):
        hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
    plan2 = hcg.collate(ei.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs, validate=False)
    with torch.no_grad():
        o2, _ = hcg.gcn_apply(b.x.cuda(), plan2, _dims(hcg, sd), params)
    torch.cuda.synchronize()
    assert torch.isfinite(o2).all()


# ------------------------------------------------------------------------------------------
# the two CUDA backward paths (graph-resident kernel / layer-wise kernels) against the fp64 oracle
# ------------------------------------------------------------------------------------------
@pytest.fixture
def backward_path(hcg):
    from hcatgnet_b200 import _lib
    prev = []

    def set_path(v):
        prev.append(_lib.config_set(_lib.HCG_CFG_BACKWARD_PATH, v))

    yield set_path
    if prev:
        _lib.config_set(_lib.HCG_CFG_BACKWARD_PATH, prev[0])


@pytest.mark.parametrize("path", [1, 2, 3])   # layer-wise / graph-resident (tcgen05 tiles + overflow) / graph-resident mma.sync only
@pytest.mark.parametrize("rset,outer,inner,count", [("learning", 1, 2, 668), ("biaryl", 5, 6, 485), ("final_test", 8, 10, 52)])
def test_backward_paths_match_oracle(hcg, backward_path, path, rset, outer, inner, count):
    """Whole datasets in one batch: more graphs than CTAs (each CTA accumulates several graphs), biaryl
    graphs above the 128-node shared-memory capacity (global scratch path) and isolated atoms."""
    backward_path(path)
    sd, graphs, _, _ = golden_io.rows_with_graphs(rset, outer, inner)
    b = ref_gcn.collate(graphs[:count])
    args = (sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"], b["y"])
    l_ref, g_ref, _ = _oracle_grads(*args)
    l_cuda, g_cuda, _ = _cuda_grads(hcg, *args)
    assert abs(l_cuda - l_ref) <= 1e-5 * max(abs(l_ref), 1)
    _assert_grads(g_cuda, g_ref)


@pytest.mark.parametrize("F,nmin,nmax,B", [(25, 40, 120, 3000), (32, 1, 9, 700), (3, 100, 300, 50), (8, 128, 128, 40), (25, 1, 1, 10)])
def test_graph_resident_backward_shapes(hcg, backward_path, F, nmin, nmax, B):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(B, n_features=F, seed=11, n_min=nmin, n_max=nmax)
    sd = ref_gcn.init_state_dict(F, 64, 2, 2, 1, seed=5)
    args = (sd, b.x, b.edge_index, b.batch, b.num_graphs, b.y)
    _, g_ref, _ = _oracle_grads(*args)
    for path in (2, 3):     # the tcgen05 kernel (+ overflow list), then the mma.sync kernel alone
        backward_path(path)
        _, g_fused, _ = _cuda_grads(hcg, *args)
        _assert_grads(g_fused, g_ref)
        _, g_again, _ = _cuda_grads(hcg, *args)
        for k in g_fused:  # no floating-point atomics: run-to-run identical bits
            assert torch.equal(g_fused[k], g_again[k]), k
    backward_path(1)
    _, g_layer, _ = _cuda_grads(hcg, *args)
    _assert_grads(g_layer, g_ref)


def test_graph_resident_backward_refuses_unsupported(hcg, backward_path):
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(5, n_features=7, seed=1, n_min=3, n_max=9)
    sd = ref_gcn.init_state_dict(7, 48, 3, 2, 1, seed=5)
    backward_path(2)
    with pytest.raises(RuntimeError):
        _cuda_grads(hcg, sd, b.x, b.edge_index, b.batch, b.num_graphs, b.y)


# ------------------------------------------------------------------------------------------
# the tensor-core dense product of the layer-wise path (wide hidden layers) against the oracle
# ------------------------------------------------------------------------------------------
@pytest.fixture
def gemm_path(hcg):
    from hcatgnet_b200 import _lib
    prev = []

    def set_path(v):
        prev.append(_lib.config_set(_lib.HCG_CFG_GEMM_PATH, v))

    yield set_path
    if prev:
        _lib.config_set(_lib.HCG_CFG_GEMM_PATH, prev[0])


@pytest.mark.parametrize("H,L,F,B,nmax", [(256, 4, 25, 33, 50), (128, 2, 32, 700, 60), (192, 3, 7, 64, 40),
                                          (64, 2, 40, 300, 120), (256, 2, 25, 1, 3)])
def test_tensor_core_gemm_forward_backward(hcg, gemm_path, forward_path, backward_path, H, L, F, B, nmax):
    """Layer-wise kernels with every X.W product (forward T = h W^T, backward g_in = gT W with the
    LeakyReLU' mask) forced onto the split-TF32 tcgen05 GEMM: partial row tiles, K = F < 32, several
    K chunks, 1..4 column tiles."""
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(B, n_features=F, seed=21, n_min=1, n_max=nmax)
    sd = ref_gcn.init_state_dict(F, H, L, 2, 1, seed=9)
    forward_path(1)
    backward_path(1)
    gemm_path(2)
    params, _, _ = _sd_params(sd, "cuda")
    plan = hcg.collate(b.edge_index.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs)
    with torch.no_grad():
        o, e = hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
    o_ref, e_ref = ref_gcn.forward(sd, b.x, b.edge_index, b.batch, b.num_graphs)
    assert _close(e.cpu(), e_ref) <= TOL and _close(o.cpu(), o_ref) <= TOL
    args = (sd, b.x, b.edge_index, b.batch, b.num_graphs, b.y)
    _, g_ref, _ = _oracle_grads(*args)
    _, g_tc, _ = _cuda_grads(hcg, *args)
    _assert_grads(g_tc, g_ref)
    gemm_path(1)
    with torch.no_grad():
        o_fma, e_fma = hcg.gcn_apply(b.x.cuda(), plan, _dims(hcg, sd), params)
    assert _close(e.cpu(), e_fma.cpu()) <= TOL and _close(o.cpu(), o_fma.cpu()) <= TOL


# ------------------------------------------------------------------------------------------
# whole-row aggregation of the layer-wise path (wide hidden layers), pooling in its last epilogue
# ------------------------------------------------------------------------------------------
@pytest.fixture
def agg_path(hcg):
    from hcatgnet_b200 import _lib
    prev = []

    def set_path(v):
        prev.append(_lib.config_set(_lib.HCG_CFG_AGG_PATH, v))

    yield set_path
    if prev:
        _lib.config_set(_lib.HCG_CFG_AGG_PATH, prev[0])


@pytest.mark.parametrize("H,L,F,B,nmin,nmax", [(256, 4, 25, 33, 1, 50), (128, 2, 32, 700, 2, 60), (192, 3, 7, 64, 1, 40),
                                               (64, 2, 25, 300, 1, 120), (136, 2, 9, 5, 1, 8), (256, 2, 25, 6, 100, 300),
                                               (50, 2, 9, 40, 1, 30), (32, 3, 25, 100, 1, 80), (6, 2, 4, 9, 1, 12)])
def test_wide_row_aggregation_forward_backward(hcg, agg_path, gemm_path, forward_path, backward_path, H, L, F, B,
                                                     nmin, nmax):
    """k_agg_wide (warp tiles of whole rows, pooled per graph in the last layer's epilogue, transposed CSR in the
    backward) and k_agg_narrow (H <= 64) against k_agg + k_pool (same summation order: identical activations)
    and the oracle; 1..2 column chunks, partial chunks (H = 192, 136), narrow widths 64 / 50 / 32 / 6, graphs
    from 1 to 300 nodes."""
    from hcatgnet_b200.synth import synth_batch
    b = synth_batch(B, n_features=F, seed=33, n_min=nmin, n_max=nmax)
    # seed 6: with seed 4 one readout pre-activation of the 700-graph case lands within 1e-7 of zero, and which
    # side of the LeakyReLU kink it falls on depends on the rounding of the mean pooling (k_pool and the fp64
    # oracle disagree there; profiles/experiments/agg_grad_probe.py) - a property of the data, not of a kernel
    sd = ref_gcn.init_state_dict(F, H, L, 2, 1, seed=6)
    forward_path(1)
    backward_path(1)
    gemm_path(1)
    params, _, _ = _sd_params(sd, "cuda")
    dims = _dims(hcg, sd)
    plan = hcg.collate(b.edge_index.cuda(), b.batch.cuda(), b.num_nodes, b.num_graphs, transpose=True)
    x = b.x.cuda()
    res = {}
    for path in (1, 2):
        agg_path(path)
        with torch.no_grad():
            o, e = hcg.gcn_apply(x, plan, dims, params)
        o_s, e_s, saved = hcg.forward_raw(plan, x, dims, params, save=True)
        res[path] = (o, e, o_s, e_s, [a.clone() for a in saved[0]])
    o_ref, e_ref = ref_gcn.forward(sd, b.x, b.edge_index, b.batch, b.num_graphs)
    assert _close(res[2][1].cpu(), e_ref) <= TOL and _close(res[2][0].cpu(), o_ref) <= TOL
    assert _close(res[2][3].cpu(), e_ref) <= TOL and _close(res[2][2].cpu(), o_ref) <= TOL
    for a1, a2 in zip(res[1][4], res[2][4]):
        assert torch.equal(a1, a2)                       # same weights, same order of the sums
    assert torch.equal(res[1][1][:, :H], res[2][1][:, :H])   # max pooling is order independent
    assert _close(res[2][1].cpu(), res[1][1].cpu()) <= TOL
    args = (sd, b.x, b.edge_index, b.batch, b.num_graphs, b.y)
    _, g_ref, _ = _oracle_grads(*args)
    for path in (2, 1):
        agg_path(path)
        _, g_res, _ = _cuda_grads(hcg, *args)
        _assert_grads(g_res, g_ref)


# ------------------------------------------------------------------------------------------
# soft edge mask on every conv layer (explainers): forward and gradients against the fp64 oracle
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rset,outer,inner,count", [("learning", 1, 2, 5), ("biaryl", 1, 2, 3)])
def test_edge_mask_forward_and_gradients(hcg, rset, outer, inner, count):
    sd, graphs, _, _ = golden_io.rows_with_graphs(rset, outer, inner)
    b = ref_gcn.collate(graphs[:count])
    E = b["edge_index"].shape[1]
    gen = torch.Generator().manual_seed(5)
    mask = torch.rand(E, generator=gen) * 0.9 + 0.05
    xj = b["x"] + 0.05 * torch.rand(b["x"].shape, generator=gen)     # de-symmetrised (see the x-gradient test above)
    # oracle, fp64
    sd64 = {k: v.double() for k, v in sd.items()}
    x64, m64 = xj.double().requires_grad_(True), mask.double().requires_grad_(True)
    o_ref, e_ref = ref_gcn.forward(sd64, x64, b["edge_index"], b["batch"], b["num_graphs"], edge_mask=m64)
    (o_ref.pow(2).sum() + 0.01 * e_ref.sum()).backward()
    # CUDA
    params, _, _ = _sd_params(sd, "cuda")
    plan = hcg.collate(b["edge_index"].cuda(), b["batch"].cuda(), xj.shape[0], b["num_graphs"], transpose=True)
    xc, mc = xj.cuda().requires_grad_(True), mask.cuda().requires_grad_(True)
    o, e = hcg.gcn_apply(xc, plan, _dims(hcg, sd), params, edge_mask=mc)
    assert _close(o.detach().cpu(), o_ref.detach()) <= TOL and _close(e.detach().cpu(), e_ref.detach()) <= TOL
    (o.pow(2).sum() + 0.01 * e.sum()).backward()
    _assert_grads({"mask": mc.grad.cpu(), "x": xc.grad.cpu()}, {"mask": m64.grad, "x": x64.grad}, rtol=1e-3)
    # a mask of ones is the unmasked network
    with torch.no_grad():
        o1, _ = hcg.gcn_apply(xj.cuda(), plan, _dims(hcg, sd), params, edge_mask=torch.ones(E, device="cuda"))
        o0, _ = hcg.gcn_apply(xj.cuda(), plan, _dims(hcg, sd), params)
    assert _close(o1.cpu(), o0.cpu()) <= TOL
