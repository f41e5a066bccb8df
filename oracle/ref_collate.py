"""TEST INFRASTRUCTURE — integer reference for the batch-collate step (numpy).

PyG never materialises a CSR; the reference path works on the COO ``edge_index`` through
``gcn_norm`` / ``propagate`` (call sites: reference model/gcn.py:58,62).  The product builds
a dst-sorted CSR once per batch instead, so the integer contract is defined HERE and the
CUDA collate kernel must reproduce it bit-for-bit:

* self-loop edges (src == dst) are dropped from the list (``add_remaining_self_loops``
  removes them and re-adds exactly one loop per node; the loop is implicit in the CSR);
* kept edges are ordered by (dst, original edge id)  — a *stable* sort by dst;
* ``rowptr[i]`` = number of kept edges with dst < i; ``col[k]`` = src of the k-th kept edge;
  ``eid[k]`` = its position in the original ``edge_index``;
* ``deg[i]`` = in-degree over kept edges (the reference's ``deg`` is this + self-loop weight);
* ``graph_ptr[g]`` = first node of graph g (``batch`` is sorted; PyG ``Batch.ptr``).

PARITY STATUS: pinned by construction against numpy's stable argsort and, for ``deg``,
against ``ref_gcn.gcn_norm`` (tests/test_oracle_golden.py).
"""
from __future__ import annotations

import numpy as np


def csr_by_dst(edge_index: np.ndarray, num_nodes: int, by: str = "dst"):
    ei = np.asarray(edge_index, dtype=np.int64)
    src, dst = (ei[0], ei[1]) if by == "dst" else (ei[1], ei[0])
    keep = np.nonzero(src != dst)[0]
    order = keep[np.argsort(dst[keep], kind="stable")]
    col = src[order].astype(np.int32)
    eid = order.astype(np.int32)
    deg = np.bincount(dst[keep], minlength=num_nodes).astype(np.int32)
    rowptr = np.zeros(num_nodes + 1, dtype=np.int32)
    np.cumsum(deg, out=rowptr[1:])
    return rowptr, col, eid, deg


def graph_ptr(batch: np.ndarray, num_graphs: int):
    b = np.asarray(batch, dtype=np.int64)
    return np.searchsorted(b, np.arange(num_graphs + 1), side="left").astype(np.int32)


def dis_from_deg(deg: np.ndarray, fill: float = 1.0):
    d = deg.astype(np.float32) + np.float32(fill)
    with np.errstate(divide="ignore"):
        out = (np.float32(1.0) / np.sqrt(d, dtype=np.float32)).astype(np.float32)
    out[~np.isfinite(out)] = 0.0
    return out
