"""GCN / GCN_explain — drop-in mirror of reference model/gcn.py:9-139.

Same constructor options, same ``forward`` signatures, same ``state_dict`` keys, shapes and
order (``conv1.bias``, ``conv1.lin.weight``, ``conv_layers.{i}.bias``,
``conv_layers.{i}.lin.weight``, ``readout.{k}.0.weight|bias``, ``readout.{R-1}.weight|bias``),
so the shipped ``model_params.pth`` load with ``strict=True``.  The sub-modules below only
HOLD parameters; the arithmetic of the whole network (GCNConv × L, LeakyReLU, global max‖mean
pooling, readout MLP; forward and backward) runs in libhcg.so (include/hcg.h) on the GPU.
There is no CPU implementation: tensors that arrive in host memory are staged to the current
CUDA device, computed there, and the results returned on the device the inputs came from
(that keeps reference utils/utils_model.py:82-111 ``predict_network``, which forces
``.to('cpu')``, working unchanged).
"""
from __future__ import annotations

import argparse
import math

import torch
import torch.nn as nn

from .. import _lib, ops
from .networks import BaseNetwork


class _DenseLin(nn.Module):
    """Parameter holder with the layout of PyG's bias-free ``Linear`` inside GCNConv
    (key ``weight`` [out, in], glorot-uniform init)."""

    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels))
        self.reset_parameters()

    def reset_parameters(self):
        a = math.sqrt(6.0 / (self.in_channels + self.out_channels))
        with torch.no_grad():
            self.weight.uniform_(-a, a)


class GCNConv(nn.Module):
    """Parameter holder for one graph convolution (replaces ``torch_geometric.nn.GCNConv`` as
    used at reference model/gcn.py:18-20,27-29).  Registration order gives the state_dict order
    ``bias`` then ``lin.weight``.  Like PyG, the weight is drawn twice at construction (once by
    the Linear's own init, once by GCNConv.reset_parameters), which keeps the global RNG stream
    in step with the reference for seeded from-scratch training."""

    def __init__(self, in_channels: int, out_channels: int, improved: bool = False):
        super().__init__()
        self.in_channels, self.out_channels, self.improved = in_channels, out_channels, improved
        self.bias = nn.Parameter(torch.empty(out_channels))
        self.lin = _DenseLin(in_channels, out_channels)
        self.reset_parameters()

    def reset_parameters(self):
        self.lin.reset_parameters()
        with torch.no_grad():
            self.bias.zero_()

    def forward(self, *args, **kwargs):
        raise RuntimeError("hcatgnet_b200.GCNConv only holds parameters; the fused network kernel is "
                           "invoked by GCN.forward / GCN_explain.forward")

    def extra_repr(self):
        return f"{self.in_channels}, {self.out_channels}, improved={self.improved}"


def _num_graphs_of(graph, batch, n_nodes: int) -> int:
    ng = getattr(graph, "num_graphs", None) if graph is not None else None
    if ng is not None:
        return int(ng)
    if batch is None:
        return 1 if n_nodes > 0 else 0
    return int(batch.max().item()) + 1 if batch.numel() else 0  # one sync, as PyG's pooling does


def _max_nodes_hint(graph) -> int:
    """Largest graph of the batch when that is known without touching the device (a host-side
    ``ptr``); 0 = unknown.  Only sizes a kernel's shared-memory capacity."""
    if graph is None:
        return 0
    known = int(getattr(graph, "max_graph_nodes", 0) or 0)
    if known:
        return known
    ptr = getattr(graph, "ptr", None)
    if ptr is None or ptr.is_cuda or ptr.numel() < 2:
        return 0
    return int((ptr[1:] - ptr[:-1]).max())


class _PlanHolder:
    """Carries a cached topology plan for tensor-level calls; pickles empty (``torch.save(model)``)."""

    def __getstate__(self):
        return {}


class _GCNBase(BaseNetwork):
    """Everything reference GCN and GCN_explain share (their constructors are identical,
    reference model/gcn.py:11-50 and :81-120)."""

    def __init__(self, opt: argparse.Namespace, n_node_features: int):
        super().__init__(opt=opt, n_node_features=n_node_features)
        self._name = "GCN"
        self.improved = opt.improved
        # read back hcg_collate's status word for every new batch (one stream sync): bad node ids or
        # an unsorted batch vector raise instead of being skipped
        self.validate_inputs = True

        self.conv1 = GCNConv(self.n_node_features, self.embedding_dim, improved=self.improved)
        self.relu1 = nn.LeakyReLU()

        self.conv_layers = nn.ModuleList([])
        for _ in range(self.n_convolutions - 1):
            self.conv_layers.append(GCNConv(self.embedding_dim, self.embedding_dim, self.improved))

        graph_embedding = self.embedding_dim * 2
        self.readout = nn.ModuleList([])
        for _ in range(self.readout_layers - 1):
            reduced_dim = int(graph_embedding / 2)
            self.readout.append(nn.Sequential(nn.Linear(graph_embedding, reduced_dim), nn.LeakyReLU()))
            graph_embedding = reduced_dim
        self.readout.append(nn.Linear(graph_embedding, self._n_classes))

        self._make_loss(opt.problem_type)
        self._make_optimizer(opt.optimizer, opt.lr)
        self._make_scheduler(scheduler=opt.scheduler, step_size=opt.step_size, gamma=opt.gamma, min_lr=opt.min_lr)

    # -- plumbing --------------------------------------------------------------------------
    def _dims(self) -> ops.NetDims:
        return ops.NetDims(F=self.n_node_features, H=self.embedding_dim, L=self.n_convolutions,
                           R=self.readout_layers, C=self._n_classes, slope=self.relu1.negative_slope)

    def _flat_params(self):
        ps = [self.conv1.lin.weight, self.conv1.bias]
        for conv in self.conv_layers:
            ps += [conv.lin.weight, conv.bias]
        for k in range(self.readout_layers - 1):
            ps += [self.readout[k][0].weight, self.readout[k][0].bias]
        ps += [self.readout[-1].weight, self.readout[-1].bias]
        return ps

    def _run(self, x, edge_index, batch, num_graphs, edge_weight=None, cache_on=None, edge_mask=None):
        """Stage to the GPU if needed, collate (cached on ``cache_on``), run the fused network."""
        home = x.device
        device = ops.compute_device(x)
        params = self._flat_params()
        need_grad = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params) or
                                                 (edge_mask is not None and edge_mask.requires_grad))
        if edge_weight is not None and edge_weight.requires_grad and torch.is_grad_enabled():
            return self._run_edge_weight_grad(x, edge_index, batch, num_graphs, edge_weight, home, device)
        if not need_grad and edge_weight is None and edge_mask is None and x.is_cuda and edge_index.is_cuda and \
                getattr(cache_on, "_hcg_plan", None) is None and ops.edge_list_hints_ok(cache_on) and \
                cache_on.ptr.is_cuda and cache_on.edge_ptr.is_cuda and x.dtype == torch.float32 and \
                edge_index.dtype == torch.int64 and cache_on.ptr.dtype == torch.int64 and \
                _lib.load().hcg_config_get(_lib.HCG_CFG_FORWARD_PATH) == _lib.PATH_AUTO:
            # inference on a batch that carries its offsets and size facts: no collate, the conv kernel reads the
            # edge list itself (refused -> None -> the general path below)
            pd = [ops.to_device(p, device, torch.float32) for p in params]
            res = ops.forward_edge_list(x.contiguous(), edge_index.contiguous(), cache_on.ptr, cache_on.edge_ptr,
                                        self._dims(), pd, validate=self.validate_inputs)
            if res is not None:
                return res
        plan = ops.cached_collate(cache_on, edge_index, batch, x.shape[0], num_graphs, device,
                                  edge_weight=edge_weight, improved=bool(self.improved), transpose=need_grad,
                                  validate=self.validate_inputs, max_graph_nodes=_max_nodes_hint(cache_on),
                                  grouped=getattr(cache_on, "edges_grouped", None))
        xd = ops.to_device(x, device, torch.float32)
        pd = [ops.to_device(p, device, torch.float32) for p in params]
        md = ops.to_device(edge_mask, device, torch.float32)
        out, emb = ops.gcn_apply(xd, plan, self._dims(), pd, edge_mask=md)
        if home != device:
            out, emb = out.to(home), emb.to(home)
        return out, emb


    def _run_edge_weight_grad(self, x, edge_index, batch, num_graphs, edge_weight, home, device):
        """Gradient w.r.t. ``edge_weight`` (reference model/gcn.py:124-127 hands it to conv1 only).
        The weights enter conv1's NORMALISATION (weighted degrees), so their derivative is not a per-edge dot product
        the aggregation kernels could emit: conv1 is written here with differentiable torch operations on the GPU
        (gcn_norm with weights: existing self-loops keep their weight, added ones get 2.0 when ``improved``; transform,
        weighted sum to dst, bias) and everything after it - the remaining conv layers, pooling, readout - runs in
        libhcg on conv1's output, whose backward returns the gradient w.r.t. that input.  A path for explainers; it
        is never taken when ``edge_weight`` does not require grad."""
        xd = ops.to_device(x, device, torch.float32)
        ei = ops.to_device(edge_index, device, torch.int64)
        bt = ops.to_device(batch, device, torch.int64)
        ew = edge_weight if edge_weight.device == device else edge_weight.to(device)
        n = xd.shape[0]
        row, col = ei[0], ei[1]
        loop = row == col
        keep = ~loop
        w_loop = torch.full((n,), 2.0 if self.improved else 1.0, dtype=torch.float32, device=device)
        if bool(loop.any()):
            w_loop = w_loop.index_put((row[loop],), ew[loop].to(torch.float32))     # an existing self-loop keeps its weight
        ar = torch.arange(n, dtype=torch.int64, device=device)
        w = torch.cat([ew[keep].to(torch.float32), w_loop])
        row2, col2 = torch.cat([row[keep], ar]), torch.cat([col[keep], ar])
        deg = torch.zeros(n, dtype=torch.float32, device=device).index_add(0, col2, w)
        dis = deg.pow(-0.5)
        dis = dis.masked_fill(dis == float("inf"), 0.0)
        norm = dis[row2] * w * dis[col2]
        params = [ops.to_device(p, device, torch.float32) for p in self._flat_params()]
        t = xd @ params[0].t()
        z1 = torch.zeros(n, self.embedding_dim, dtype=torch.float32, device=device).index_add(0, col2, t[row2] * norm[:, None]) + params[1]
        h1 = torch.nn.functional.leaky_relu(z1, self.relu1.negative_slope)
        dims = self._dims()
        if dims.L > 1:
            plan = ops.collate(ei, bt, n, num_graphs, transpose=True, validate=self.validate_inputs)
            sub = ops.NetDims(F=dims.H, H=dims.H, L=dims.L - 1, R=dims.R, C=dims.C, slope=dims.slope)
            out, emb = ops.gcn_apply(h1, plan, sub, params[2:])
        else:
            b = bt if bt is not None else torch.zeros(n, dtype=torch.int64, device=device)
            idx = b[:, None].expand(-1, h1.shape[1])
            mx = torch.full((num_graphs, h1.shape[1]), float("-inf"), device=device).scatter_reduce(0, idx, h1, reduce="amax", include_self=False)
            cnt = torch.zeros(num_graphs, device=device).index_add(0, b, torch.ones(n, device=device)).clamp(min=1)
            mean = torch.zeros(num_graphs, h1.shape[1], device=device).index_add(0, b, h1) / cnt[:, None]
            emb = torch.cat([mx, mean], dim=1)
            z = emb
            for k in range(dims.R):
                wk, bk = params[2 * dims.L + 2 * k], params[2 * dims.L + 2 * k + 1]
                z = z @ wk.t() + bk
                if k < dims.R - 1:
                    z = torch.nn.functional.leaky_relu(z, dims.slope)
            out = z
        if home != device:
            out, emb = out.to(home), emb.to(home)
        return out, emb


class GCN(_GCNBase):
    """reference model/gcn.py:9-76."""

    def forward(self, reaction_graph, return_graph_embedding=False):
        store = getattr(reaction_graph, "_packed_store", None)
        if store is not None and not torch.is_grad_enabled() and getattr(reaction_graph, "_real", None) is None and \
                _lib.load().hcg_config_get(_lib.HCG_CFG_FORWARD_PATH) == _lib.PATH_AUTO and \
                (self.validate_inputs or ops.edge_list_hints_ok(reaction_graph)):
            # a device-resident packed image that nobody has unpacked: inference straight from it
            dev = reaction_graph.ptr.device
            pd = [ops.to_device(p, dev, torch.float32) for p in self._flat_params()]
            res = ops.forward_packed(store.c_store, 0, reaction_graph.num_graphs, self._dims(), pd, dev,
                                     validate=self.validate_inputs)
            if res is not None:
                return res if return_graph_embedding == True else res[0]  # noqa: E712
        x, edge_index = reaction_graph.x, reaction_graph.edge_index
        # (our Batch builds its batch vector lazily from ptr: the kernels work from ptr, so leave it alone if it is not there yet)
        batch = ops.batch_vector_if_built(reaction_graph)
        # edge_weight is None on this path, so `improved` does not change the normalisation
        # (self-loop weight 1.0) — the behaviour the shipped golden outputs pin
        out, graph_emb = self._run(x, edge_index, batch, _num_graphs_of(reaction_graph, batch, x.shape[0]),
                                   cache_on=reaction_graph)
        if return_graph_embedding == True:  # noqa: E712  (same test as the reference, model/gcn.py:73)
            return out, graph_emb
        return out


class GCN_explain(_GCNBase):
    """reference model/gcn.py:79-139: tensor-level entry; ``edge_weight`` reaches conv1 only."""

    def forward(self, x=None, edge_index=None, batch_index=None, edge_weight=None, edge_mask=None):
        """``edge_mask`` ([E], optional, differentiable) is this package's stand-in for the hooks PyG's
        Explainer installs on every MessagePassing layer: it multiplies the messages of ALL conv layers,
        leaves added self-loops and the normalisation alone (see hcatgnet_b200/explain.py)."""
        out, _ = self._run(x, edge_index, batch_index, _num_graphs_of(None, batch_index, x.shape[0]),
                           edge_weight=edge_weight, cache_on=self._plan_cache(edge_index, edge_weight), edge_mask=edge_mask)
        return out

    def _plan_cache(self, edge_index, edge_weight):
        """An explainer calls forward hundreds of times on one graph: keep the topology plan on a small holder
        keyed by the edge_index tensor (cached_collate re-checks pointer, version and shape)."""
        if edge_weight is not None:
            return None
        holder = getattr(self, "_explain_holder", None)
        if holder is None:
            holder = _PlanHolder()
            object.__setattr__(self, "_explain_holder", holder)
        return holder
