"""Synthetic reaction-graph batches of the shape BASELINE.json's configs name.

Per graph: ``n ~ U{n_min..n_max}`` atoms (default 40..120, mean 80); bonds = a random spanning
tree (each atom bonds to one of its two predecessors, giving chain-with-branches molecules)
plus ring-closure bonds up to ``round(1.5 n)`` bonds in total, no self-loops, no duplicates
=> ``e = 3 n`` directed entries stored as adjacent ``(i,j),(j,i)`` pairs in random bond order
(unsorted by src or dst, like reference data/rhcaa.py:157-164).  Node features ``x[n,F]`` are
U[0,1) fp32, targets ``y ~ N(0, 8^2)``.  Everything is produced the way PyG's collate would
deliver it: ``x`` fp32 [N,F], ``edge_index`` int64 [2,E] (batch-global node ids), ``batch``
int64 [N].  Fully vectorised numpy; deterministic in ``seed``.
"""
from __future__ import annotations

import numpy as np
import torch

from .data import Batch


def synth_batch_numpy(num_graphs: int, n_features: int = 25, seed: int = 20232023,
                      n_min: int = 40, n_max: int = 120, bonds_per_atom: float = 1.5, features: str = "uniform"):
    rng = np.random.default_rng(seed)
    B = int(num_graphs)
    n = rng.integers(n_min, n_max + 1, size=B).astype(np.int64)
    ptr = np.zeros(B + 1, dtype=np.int64)
    np.cumsum(n, out=ptr[1:])
    N = int(ptr[-1])
    gid = np.repeat(np.arange(B, dtype=np.int64), n)
    k = np.arange(N, dtype=np.int64) - ptr[gid]            # local atom index

    # spanning tree: atom k>=1 bonds to k-1 or k-2
    step = 1 + rng.integers(0, 2, size=N)
    step = np.minimum(step, k)
    tree_hi = np.nonzero(k >= 1)[0]
    tree_lo = tree_hi - step[tree_hi]

    # ring closures: `extra_g` distinct atoms k>=7 per graph, each bonded to k-d, d in 3..7
    target = np.rint(bonds_per_atom * n).astype(np.int64)
    extra = np.clip(target - (n - 1), 0, np.maximum(n - 7, 0))
    key = rng.random(N)
    key[k < 7] = 2.0                                       # ineligible atoms sort last
    order = np.lexsort((key, gid))                         # by graph, then random key
    rank = np.empty(N, dtype=np.int64)
    rank[order] = np.arange(N, dtype=np.int64) - ptr[gid[order]]
    ring_hi = np.nonzero(rank < extra[gid])[0]
    ring_lo = ring_hi - rng.integers(3, 8, size=ring_hi.size)

    hi = np.concatenate([tree_hi, ring_hi])
    lo = np.concatenate([tree_lo, ring_lo])
    # random bond order inside each graph, random orientation of each bond
    bg = gid[hi]
    perm = np.lexsort((rng.random(hi.size), bg))
    hi, lo = hi[perm], lo[perm]
    flip = rng.random(hi.size) < 0.5
    a = np.where(flip, lo, hi)
    b = np.where(flip, hi, lo)
    edge_index = np.empty((2, 2 * hi.size), dtype=np.int64)
    edge_index[0, 0::2], edge_index[1, 0::2] = a, b
    edge_index[0, 1::2], edge_index[1, 1::2] = b, a

    x = rng.random((N, n_features), dtype=np.float32)
    if features == "featuriser":
        # the block structure of the reference's node features (reference data/rhcaa.py:112-150): one-hot
        # element / degree / hybridisation blocks, binary flags, and the reaction temperature (one value per
        # graph, normalised) in the last column.  Drawn AFTER the uniform matrix so that the graphs, targets and
        # the generator state of the default variant do not change.
        widths = [10, 4, 4] + [1] * max(n_features - 19, 0)
        r2 = np.random.default_rng(seed + 1)
        x = np.zeros((N, n_features), dtype=np.float32)
        c = 0
        for w in widths:
            if c + w > n_features - 1:
                break
            if w > 1:
                x[np.arange(N), c + r2.integers(0, w, size=N)] = 1.0
            else:
                x[:, c] = (r2.random(N) < 0.3).astype(np.float32)
            c += w
        x[:, -1] = (r2.integers(1, 6, size=B).astype(np.float32) / 5.0)[gid]
    elif features != "uniform":
        raise ValueError(f"unknown feature style {features!r}")
    y = (rng.standard_normal(B) * 8.0).astype(np.float32)
    edge_ptr = np.zeros(B + 1, dtype=np.int64)          # bonds are listed graph by graph: 2 directed edges per bond
    np.cumsum(2 * np.bincount(bg, minlength=B), out=edge_ptr[1:])
    return {"x": x, "edge_index": edge_index, "batch": gid, "ptr": ptr, "edge_ptr": edge_ptr, "y": y, "num_graphs": B}


def synth_batch(num_graphs: int, n_features: int = 25, seed: int = 20232023, pin: bool = False, **kw) -> Batch:
    d = synth_batch_numpy(num_graphs, n_features, seed, **kw)
    b = Batch(torch.from_numpy(d["x"]), torch.from_numpy(d["edge_index"]), torch.from_numpy(d["batch"]),
              torch.from_numpy(d["ptr"]), torch.from_numpy(d["y"]),
              torch.arange(num_graphs, dtype=torch.int64), num_graphs, edges_grouped=True)
    b.edge_ptr = torch.from_numpy(d["edge_ptr"])
    b.max_graph_edges = int(np.diff(d["edge_ptr"]).max()) if num_graphs else 0
    b.max_in_degree = int(np.bincount(d["edge_index"][1]).max()) if d["edge_index"].size else 0
    return b.pin_memory() if pin else b
