"""TEST INFRASTRUCTURE — CPU restatement (oracle) of the HCat-GNet GCN hot path.

This file restates, in plain PyTorch, the arithmetic the reference delegates to the
un-vendored dependency ``torch_geometric`` (reference requirements.txt:13,
``torch_geometric>=2.4.0``, unpinned; absent from /root/reference and not installable
here).  It follows the reference's own call sites:

* op order of the network ............ reference model/gcn.py:54-76 (GCN.forward)
* tensor-level entry .................. reference model/gcn.py:124-139 (GCN_explain.forward)
* module layout / state_dict keys ..... reference model/gcn.py:11-50, model/networks.py:10-19
* loss / train step ................... reference utils/utils_model.py:55-70
* published algorithm of GCNConv ...... Kipf & Welling 2017: out = D^-1/2 (A+I) D^-1/2 X W + b,
  as implemented by PyG ``gcn_norm`` + ``propagate(aggr='add')``.

PARITY STATUS: **pinned** for the forward pass — checked against the reference's shipped
golden vectors (``embeddings.csv`` rows written by reference utils/utils_model.py:199-205
with weights from ``model_params.pth``) in tests/test_oracle_golden.py.  Backward is
pinned only against this restatement's autograd (the reference ships no gradient vectors).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference arm
may import this module.  The product (``hcatgnet_b200``) never does.
"""
from __future__ import annotations

from collections import OrderedDict

import torch

LEAKY_SLOPE = 0.01  # nn.LeakyReLU() default, reference model/gcn.py:21,63


def gcn_norm(edge_index: torch.Tensor, num_nodes: int, edge_weight=None, improved: bool = False,
             dtype=torch.float32):
    """PyG ``gcn_norm`` for the flow source_to_target (row = src j, col = dst i).

    With ``edge_weight=None`` (what the reference always passes, model/gcn.py:56,58,62) the
    self-loop weight is 1.0 even when ``improved=True`` — SURVEY.md §0 fact 4, verified
    against the golden vectors.  With explicit weights the fill value is 2.0 if improved.
    Pre-existing self-loop edges are removed and one loop per node is appended LAST
    (``add_remaining_self_loops``); with explicit weights an existing loop keeps its weight.
    """
    row, col = edge_index[0], edge_index[1]
    n = num_nodes
    loop_mask = row == col
    keep = ~loop_mask
    ar = torch.arange(n, dtype=row.dtype)
    if edge_weight is None:
        w_loop = torch.ones(n, dtype=dtype)
        w = torch.cat([torch.ones(int(keep.sum()), dtype=dtype), w_loop])
    else:
        fill = 2.0 if improved else 1.0
        w_loop = torch.full((n,), fill, dtype=dtype)
        # an existing self-loop carries its own weight (last one wins, as index assignment does)
        w_loop[row[loop_mask]] = edge_weight[loop_mask].to(dtype)
        w = torch.cat([edge_weight[keep].to(dtype), w_loop])
    row2 = torch.cat([row[keep], ar])
    col2 = torch.cat([col[keep], ar])
    deg = torch.zeros(n, dtype=dtype).index_add_(0, col2, w)
    dis = deg.pow(-0.5)
    dis = dis.masked_fill(dis == float("inf"), 0.0)
    norm = dis[row2] * w * dis[col2]
    return row2, col2, norm, deg, dis


def gcn_conv(h, edge_index, weight, bias, edge_weight=None, improved=False, edge_mask=None):
    """One GCNConv: transform first (no bias), weighted sum to dst in edge order, bias last.

    ``edge_mask`` ([E], already in (0,1)) restates what PyG's Explainer hooks do inside
    ``MessagePassing.propagate`` (explain mode, ``edge_mask_type='object'``): every message is multiplied
    by its edge's mask, the self-loops GCNConv added get mask 1, the normalisation is untouched."""
    n = h.shape[0]
    row, col, norm, _, _ = gcn_norm(edge_index, n, edge_weight, improved, dtype=h.dtype)
    if edge_mask is not None:
        keep = edge_index[0] != edge_index[1]
        norm = norm * torch.cat([edge_mask[keep].to(h.dtype), torch.ones(n, dtype=h.dtype)])
    t = h @ weight.t()
    out = torch.zeros(n, weight.shape[0], dtype=h.dtype).index_add_(0, col, t[row] * norm[:, None])
    return out + bias


def segment_max(h, batch, num_graphs):
    out = torch.full((num_graphs, h.shape[1]), float("-inf"), dtype=h.dtype)
    idx = batch[:, None].expand(-1, h.shape[1])
    return out.scatter_reduce(0, idx, h, reduce="amax", include_self=False)


def segment_mean(h, batch, num_graphs):
    s = torch.zeros(num_graphs, h.shape[1], dtype=h.dtype).index_add_(0, batch, h)
    cnt = torch.zeros(num_graphs, dtype=h.dtype).index_add_(0, batch, torch.ones_like(batch, dtype=h.dtype))
    return s / cnt.clamp(min=1)[:, None]


def state_dict_keys(n_convolutions: int, readout_layers: int):
    """Key order of the reference state_dict."""
    keys = ["conv1.bias", "conv1.lin.weight"]
    for i in range(n_convolutions - 1):
        keys += [f"conv_layers.{i}.bias", f"conv_layers.{i}.lin.weight"]
    for k in range(readout_layers - 1):
        keys += [f"readout.{k}.0.weight", f"readout.{k}.0.bias"]
    keys += [f"readout.{readout_layers - 1}.weight", f"readout.{readout_layers - 1}.bias"]
    return keys


def infer_arch(sd) -> tuple[int, int]:
    n_conv = 1 + len({k.split(".")[1] for k in sd if k.startswith("conv_layers.")})
    n_ro = len({k.split(".")[1] for k in sd if k.startswith("readout.")})
    return n_conv, n_ro


def forward(sd, x, edge_index, batch, num_graphs=None, edge_weight=None, improved=True,
            return_all=False, edge_mask=None):
    """Whole network on a state_dict-like mapping of tensors (reference model/gcn.py:54-76).

    Returns ``(out[B,C], emb[B,2H])`` (and the per-layer activations when ``return_all``).
    ``edge_weight`` reaches conv1 only, exactly as reference model/gcn.py:127 vs :131.
    """
    n_conv, n_ro = infer_arch(sd)
    if num_graphs is None:
        num_graphs = int(batch.max()) + 1 if batch.numel() else 0
    acts = []
    h = torch.nn.functional.leaky_relu(
        gcn_conv(x, edge_index, sd["conv1.lin.weight"], sd["conv1.bias"], edge_weight, improved, edge_mask), LEAKY_SLOPE)
    acts.append(h)
    for i in range(n_conv - 1):
        h = torch.nn.functional.leaky_relu(
            gcn_conv(h, edge_index, sd[f"conv_layers.{i}.lin.weight"], sd[f"conv_layers.{i}.bias"],
                     edge_mask=edge_mask), LEAKY_SLOPE)
        acts.append(h)
    emb = torch.cat([segment_max(h, batch, num_graphs), segment_mean(h, batch, num_graphs)], dim=1)
    z = emb
    for k in range(n_ro - 1):
        z = torch.nn.functional.leaky_relu(z @ sd[f"readout.{k}.0.weight"].t() + sd[f"readout.{k}.0.bias"], LEAKY_SLOPE)
    out = z @ sd[f"readout.{n_ro - 1}.weight"].t() + sd[f"readout.{n_ro - 1}.bias"]
    if return_all:
        return out, emb, acts
    return out, emb


def rmse_loss(out, y):
    """sqrt(MSELoss(out, y[:,None])) — reference utils/utils_model.py:64."""
    return torch.sqrt(torch.nn.functional.mse_loss(out, y.unsqueeze(1)))


def init_state_dict(n_node_features, embedding_dim=64, n_convolutions=2, readout_layers=2,
                    n_classes=1, seed=20232023, dtype=torch.float32):
    """Random weights with the reference's initialisers: glorot-uniform
    conv weights, zero conv bias, torch-default readout Linear."""
    g = torch.Generator().manual_seed(seed)
    sd = OrderedDict()

    def glorot(o, i):
        a = (6.0 / (i + o)) ** 0.5
        return (torch.rand(o, i, generator=g, dtype=dtype) * 2 - 1) * a

    def lin_default(o, i):
        bound = 1.0 / i ** 0.5
        w = (torch.rand(o, i, generator=g, dtype=dtype) * 2 - 1) * bound
        b = (torch.rand(o, generator=g, dtype=dtype) * 2 - 1) * bound
        return w, b

    sd["conv1.bias"] = torch.zeros(embedding_dim, dtype=dtype)
    sd["conv1.lin.weight"] = glorot(embedding_dim, n_node_features)
    for i in range(n_convolutions - 1):
        sd[f"conv_layers.{i}.bias"] = torch.zeros(embedding_dim, dtype=dtype)
        sd[f"conv_layers.{i}.lin.weight"] = glorot(embedding_dim, embedding_dim)
    d = 2 * embedding_dim
    for k in range(readout_layers - 1):
        w, b = lin_default(d // 2, d)
        sd[f"readout.{k}.0.weight"], sd[f"readout.{k}.0.bias"] = w, b
        d //= 2
    w, b = lin_default(n_classes, d)
    sd[f"readout.{readout_layers - 1}.weight"], sd[f"readout.{readout_layers - 1}.bias"] = w, b
    return sd


def collate(graphs):
    """PyG ``Batch.from_data_list`` restated (call sites: reference call_methods.py:41,44,46):
    cat x, offset + cat edge_index, ``batch`` = node→graph id, ``ptr``, cat y, tensorise idx."""
    xs, eis, bs, ys, idxs, ptr = [], [], [], [], [], [0]
    off = 0
    for g, d in enumerate(graphs):
        n = d["x"].shape[0]
        xs.append(d["x"])
        eis.append(d["edge_index"] + off)
        bs.append(torch.full((n,), g, dtype=torch.int64))
        if "y" in d:
            ys.append(torch.as_tensor(d["y"], dtype=torch.float32).reshape(-1))
        if "idx" in d:
            idxs.append(int(d["idx"]))
        off += n
        ptr.append(off)
    out = {
        "x": torch.cat(xs, 0),
        "edge_index": torch.cat(eis, 1),
        "batch": torch.cat(bs, 0),
        "ptr": torch.tensor(ptr, dtype=torch.int64),
        "num_graphs": len(graphs),
    }
    if ys:
        out["y"] = torch.cat(ys, 0)
    if idxs:
        out["idx"] = torch.tensor(idxs, dtype=torch.int64)
    return out
