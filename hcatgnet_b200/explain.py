"""GNNExplainer on the CUDA kernels — what reference scripts_experiments/explain_gnn.py:39-57 and
utils/other_utils.py:23-88 obtain from ``torch_geometric.explain.Explainer(model, GNNExplainer(),
explanation_type='model', node_mask_type='attributes', edge_mask_type='object', mode='regression',
task_level='graph', return_type='raw')`` (only ``explanation.node_mask`` is consumed there).

The algorithm is the published one (Ying et al., NeurIPS 2019) with PyG's defaults: a soft mask on the node
attributes (``x * sigmoid(node_mask)``) and a soft mask on the edges, applied to the messages of every
conv layer, are trained with Adam so that the masked prediction reproduces the model's own prediction,
under size and entropy regularisers.  PyG injects the edge mask through hooks in ``MessagePassing``; here it
is the ``edge_mask`` input of ``GCN_explain.forward`` (``HcgGraph.edge_mask`` in include/hcg.h), whose
gradient — and the one w.r.t. ``x`` — come out of ``hcg_gcn_backward``.
"""
from __future__ import annotations

import math

import torch

COEFFS = {"edge_size": 0.005, "node_feat_size": 1.0, "edge_ent": 1.0, "node_feat_ent": 0.1, "EPS": 1e-15}


def init_masks(num_nodes: int, num_features: int, num_edges: int, device, generator: torch.Generator | None = None):
    """PyG's initialisation: node mask ~ N(0, 0.1²) per attribute, edge mask ~ N(0, std²) with
    std = gain('relu') * sqrt(2 / (2 N))."""
    std = torch.nn.init.calculate_gain("relu") * math.sqrt(2.0 / (2 * max(num_nodes, 1)))
    node = (torch.randn(num_nodes, num_features, generator=generator) * 0.1).to(device)
    edge = (torch.randn(num_edges, generator=generator) * std).to(device)
    return node.requires_grad_(True), edge.requires_grad_(True)


def _regularised_loss(y_hat, y, node_mask, edge_mask, hard_node, hard_edge):
    loss = torch.nn.functional.mse_loss(y_hat, y)
    eps = COEFFS["EPS"]
    m = edge_mask[hard_edge].sigmoid()
    loss = loss + COEFFS["edge_size"] * m.sum()
    ent = -m * torch.log(m + eps) - (1 - m) * torch.log(1 - m + eps)
    loss = loss + COEFFS["edge_ent"] * ent.mean()
    m = node_mask[hard_node].sigmoid()
    loss = loss + COEFFS["node_feat_size"] * m.mean()
    ent = -m * torch.log(m + eps) - (1 - m) * torch.log(1 - m + eps)
    loss = loss + COEFFS["node_feat_ent"] * ent.mean()
    return loss


def run_gnnexplainer(forward, x, num_edges: int, epochs: int = 100, lr: float = 0.01, masks=None,
                     generator: torch.Generator | None = None):
    """The optimisation loop.  ``forward(h, edge_mask_or_None)`` -> prediction ``[B, C]``; it is called once
    without masks for the target and once per epoch with ``h = x * sigmoid(node_mask)`` and
    ``sigmoid(edge_mask)``.  Returns ``(node_mask, edge_mask)`` post-processed as PyG does (sigmoid, entries
    that never received gradient zeroed)."""
    with torch.no_grad():
        target = forward(x, None).detach()
    node_mask, edge_mask = masks if masks is not None else init_masks(x.shape[0], x.shape[1], num_edges, x.device, generator)
    opt = torch.optim.Adam([node_mask, edge_mask], lr=lr)
    hard_node = hard_edge = None
    for it in range(epochs):
        opt.zero_grad()
        y_hat = forward(x * node_mask.sigmoid(), edge_mask.sigmoid())
        if hard_node is None:
            loss = torch.nn.functional.mse_loss(y_hat, target)   # first iteration: no regularisers (PyG: hard masks unset)
        else:
            loss = _regularised_loss(y_hat, target, node_mask, edge_mask, hard_node, hard_edge)
        loss.backward()
        opt.step()
        if it == 0:
            # the entries whose gradient is non-zero are the ones that take part in the prediction at all
            hard_node, hard_edge = node_mask.grad != 0.0, edge_mask.grad != 0.0
    with torch.no_grad():
        nm, em = node_mask.sigmoid(), edge_mask.sigmoid()
        if hard_node is not None:
            nm, em = nm * hard_node, em * hard_edge
    return nm, em


def gnn_explainer(model, x, edge_index, batch_index=None, epochs: int = 100, lr: float = 0.01, seed: int | None = None,
                  masks=None):
    """Explains the prediction of a ``GCN_explain`` module for the graph(s) ``(x, edge_index, batch_index)``.
    Returns ``{"node_mask": [N, F], "edge_mask": [E]}`` on the device of ``x``'s compute GPU."""
    from . import ops
    dev = ops.compute_device(x)
    x = x.detach().to(dev, torch.float32)
    edge_index = edge_index.to(dev)
    batch_index = None if batch_index is None else batch_index.to(dev)
    gen = None
    if seed is not None:
        gen = torch.Generator().manual_seed(seed)
    was_training = model.training
    model.eval()
    frozen = [(p, p.requires_grad) for p in model.parameters()]
    for p, _ in frozen:
        p.requires_grad_(False)   # only the masks are trained: no parameter-gradient work in the backward
    try:
        def fwd(h, m):
            return model(x=h, edge_index=edge_index, batch_index=batch_index, edge_mask=m)
        nm, em = run_gnnexplainer(fwd, x, int(edge_index.shape[1]), epochs=epochs, lr=lr, masks=masks, generator=gen)
    finally:
        for p, rg in frozen:
            p.requires_grad_(rg)
        model.train(was_training)
    return {"node_mask": nm, "edge_mask": em}
