"""Minimal batch container + loader with the surface the reference's loops touch.

The reference gets these from torch_geometric (``Batch.from_data_list`` via ``DataLoader``;
call sites reference call_methods.py:3,41,44,46 and scripts_experiments/predict_test.py:27).
Callers use ``batch.x / .edge_index / .batch / .y / .idx / .num_graphs / .to(device)``
(reference utils/utils_model.py:61-68,76-78,89-95) and ``len(loader.dataset)`` (:70,79).
This is host bookkeeping only (concatenate + offset); the CSR / normalisation work is done on
the GPU by ``hcg_collate``.
"""
from __future__ import annotations

import random
from typing import Iterable, Mapping, Sequence

import torch


# per-graph capacity of hcg_collate's graph-local path (csrc/hcg_collate.cu: kGrpNodes / kGrpEdges)
GROUPED_MAX_NODES, GROUPED_MAX_EDGES = 1024, 2048


def _field(g, name, default=None):
    if isinstance(g, Mapping):
        return g.get(name, default)
    return getattr(g, name, default)


class Batch:
    """A disjoint union of graphs: nodes concatenated, edge indices offset per graph."""

    def __init__(self, x, edge_index, batch, ptr=None, y=None, idx=None, num_graphs=None, edges_grouped=None):
        self.x, self.edge_index, self._batch, self.ptr, self.y, self.idx = x, edge_index, batch, ptr, y, idx
        # True when edge_index is known to list the edges graph by graph (this module's own collate
        # does); lets hcg_collate take its graph-local path without a status read-back
        self.edges_grouped = edges_grouped
        if num_graphs is None:
            num_graphs = (int(ptr.numel()) - 1) if ptr is not None else (int(batch.max()) + 1 if batch.numel() else 0)
        self.num_graphs = int(num_graphs)
        # size of the largest graph, kept as a host integer (it survives .to(device)); 0 = unknown
        self.max_graph_nodes = 0
        if ptr is not None and not ptr.is_cuda and ptr.numel() >= 2:
            self.max_graph_nodes = int((ptr[1:] - ptr[:-1]).max())
        # optional: first edge of each graph (PyG keeps the same numbers in Batch._slice_dict['edge_index']) and
        # the edge count of the largest graph, as a host integer; None / 0 = unknown
        self.edge_ptr = None
        self.max_graph_edges = 0
        # largest number of incoming edges of any node (host integer), None = unknown: with the sizes above it lets
        # inference run straight from the edge list (hcg_gcn_forward_coo: no collate)
        self.max_in_degree = None

    @classmethod
    def from_data_list(cls, graphs: Sequence) -> "Batch":
        sizes = [int(_field(g, "x").shape[0]) for g in graphs]
        ptr = torch.zeros(len(graphs) + 1, dtype=torch.int64)
        if sizes:
            ptr[1:] = torch.cumsum(torch.tensor(sizes, dtype=torch.int64), 0)
        x = torch.cat([_field(g, "x") for g in graphs], 0)
        ecount = torch.tensor([int(_field(g, "edge_index").shape[1]) for g in graphs], dtype=torch.int64)
        edge_index = torch.cat([_field(g, "edge_index") for g in graphs], 1).to(torch.int64)
        edge_index = edge_index + torch.repeat_interleave(ptr[:-1], ecount)[None, :]
        batch = torch.repeat_interleave(torch.arange(len(graphs), dtype=torch.int64),
                                        torch.tensor(sizes, dtype=torch.int64))
        ys = [_field(g, "y") for g in graphs]
        y = None if any(v is None for v in ys) else torch.cat(
            [torch.as_tensor(v, dtype=torch.float32).reshape(-1) for v in ys], 0)
        ids = [_field(g, "idx") for g in graphs]
        idx = None if any(v is None for v in ids) else torch.tensor([int(v) for v in ids], dtype=torch.int64)
        # The graph-local collate path (HCG_COLLATE_GROUPED) is only vouched for when the host has checked its
        # preconditions: the edges are listed graph by graph by construction here; self-loop edges and graphs
        # above the path's per-graph capacity are looked for explicitly (otherwise hcg_collate is left to
        # try the path under `validate` and fall back by itself).
        grouped = bool(edge_index.numel() == 0 or not bool((edge_index[0] == edge_index[1]).any()))
        grouped = grouped and (not sizes or max(sizes) <= GROUPED_MAX_NODES) and \
            (ecount.numel() == 0 or int(ecount.max()) <= GROUPED_MAX_EDGES)
        out = cls(x, edge_index, batch, ptr, y, idx, len(graphs), edges_grouped=True if grouped else None)
        out.edge_ptr = torch.zeros(len(graphs) + 1, dtype=torch.int64)
        if ecount.numel():
            out.edge_ptr[1:] = torch.cumsum(ecount, 0)
        out.max_graph_edges = int(ecount.max()) if ecount.numel() else 0
        out.max_in_degree = int(torch.bincount(edge_index[1]).max()) if edge_index.numel() and int(edge_index.min()) >= 0 else \
            (0 if not edge_index.numel() else None)
        return out

    @property
    def batch(self):
        """The N-entry graph id of every node.  It says what ``ptr`` says; loaders that ship ``ptr`` leave it out and it
        is rebuilt here on first use (on the device: ``hcg_batch_from_ptr``) - the kernels themselves work from ``ptr``."""
        if self._batch is None and self.ptr is not None and self.x is not None:
            if self.ptr.is_cuda:
                from . import ops
                self._batch = ops.batch_from_ptr(self.ptr, int(self.x.shape[0]))
            else:
                self._batch = torch.repeat_interleave(torch.arange(self.ptr.numel() - 1, dtype=torch.int64),
                                                      self.ptr[1:] - self.ptr[:-1])
        return self._batch

    @batch.setter
    def batch(self, value):
        self._batch = value

    @property
    def num_nodes(self) -> int:
        return int(self.x.shape[0])

    def to(self, device, non_blocking: bool = False) -> "Batch":
        """In place, returning self (like PyG's ``Data.to``), so a collate plan cached on the
        object by the model survives ``batch = batch.to(device)`` when nothing moves."""
        moved = False
        for name in ("x", "edge_index", "_batch", "ptr", "y", "idx", "edge_ptr"):
            t = getattr(self, name)
            if t is not None:
                t2 = t.to(device, non_blocking=non_blocking)
                moved |= t2 is not t
                setattr(self, name, t2)
        if moved and hasattr(self, "_hcg_plan"):
            del self._hcg_plan
        return self

    def pin_memory(self) -> "Batch":
        for name in ("x", "edge_index", "_batch", "ptr", "y", "idx", "edge_ptr"):
            t = getattr(self, name)
            if t is not None and not t.is_cuda:
                setattr(self, name, t.pin_memory())
        return self

    def __repr__(self):
        return (f"Batch(num_graphs={self.num_graphs}, x={tuple(self.x.shape)}, "
                f"edge_index={tuple(self.edge_index.shape)})")


class DataLoader:
    """List-of-graphs loader: ``batch_size`` graphs per step, optional reshuffle every epoch
    (python ``random``, so that ``seed_everything`` governs it)."""

    def __init__(self, dataset: Iterable, batch_size: int = 1, shuffle: bool = False):
        self.dataset = list(dataset)
        self.batch_size = int(batch_size)
        self.shuffle = bool(shuffle)

    def __len__(self):
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        order = list(range(len(self.dataset)))
        if self.shuffle:
            random.shuffle(order)
        for s in range(0, len(order), self.batch_size):
            yield Batch.from_data_list([self.dataset[i] for i in order[s:s + self.batch_size]])
