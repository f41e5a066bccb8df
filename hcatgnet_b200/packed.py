"""Packed on-disk / on-wire graph format and its device-side loader.

The reference keeps one pickled ``Data`` object per reaction (``reaction_<i>.pt``, reference
data/datasets.py:74-78, data/rhcaa.py:90-92) and collates with Python per batch
(``DataLoader``, reference call_methods.py:15-47): 668 file opens + unpickles per pass, and a host
batch of 100 B per node + 16 B per directed edge.  Here the whole set is ONE binary image:

    header (128 B) | col_slot i32[F] | node_ptr i64[G+1] | edge_ptr i64[G+1] | y f32[G] | idx i64[G] |
    fold i32[G] | xu8 u8[N, Fu] | xf32 f32[N, Ff] | edges uK[E, 2] | xbits u8[N, ceil(Fb / 8)]
                                                                          (every section 64-byte aligned)

* node features are split per COLUMN: a column whose every value is 0 or 1 (the one-hot and flag columns of the
  featuriser, reference data/rhcaa.py:112-150) takes ONE BIT per node, a column of integers in 0..255 (counts) one
  byte, the others (temperature) stay f32 — lossless by construction, checked at pack time;
* edges are (src, dst) pairs LOCAL to their graph in u8 (graphs <= 256 nodes), u16 or i32, graph by graph.

RhCAA-shaped graphs take ~1.0 KB instead of ~11.9 KB (2.7 KB before the bit columns).  The image is consumed on the GPU:
``DeviceGraphStore`` keeps it resident (the 10 M-graph screening set is ~27 GB of 180 GB HBM) and
``batch(ids)`` assembles any selection of graphs — a shuffled mini-batch, a cross-validation fold — with one
kernel (``hcg_unpack_batch``, csrc/hcg_packed.cu) into exactly the tensors PyG's collate would have produced
(x f32, edge_index i64 with node offsets, batch, ptr, y): no host collate, no per-step PCIe traffic beyond
the id list.  For sets larger than HBM the same image format is the wire format: ship a chunk, unpack it.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
import re
import sys
import types
from typing import Sequence

import numpy as np
import torch

from . import _lib, ops
from .data import GROUPED_MAX_EDGES, GROUPED_MAX_NODES, Batch, _field

FLAG_GROUPABLE = 1      # header flag: the packer checked the preconditions of hcg_collate's graph-local path
MAGIC = b"HCGPACK1"
HEADER_BYTES = 128
_ALIGN = 64
_EDGE_DTYPES = {1: np.uint8, 2: np.uint16, 4: np.int32}


def _aligned(n: int) -> int:
    return (n + _ALIGN - 1) // _ALIGN * _ALIGN


BIT_SLOT0 = 65536       # col_slot values from here on: bit (slot - BIT_SLOT0) of the node's xbits row


def _sections(G: int, N: int, E: int, F: int, Fu: int, Ff: int, ek: int, Fb: int = 0):
    """(name, dtype, shape) in file order.  ``xbits`` (columns whose every value is 0 or 1, one bit each) comes last:
    images of writers that did not know it (Fb = 0 in the header) keep their layout."""
    return [("col_slot", np.int32, (F,)), ("node_ptr", np.int64, (G + 1,)), ("edge_ptr", np.int64, (G + 1,)),
            ("y", np.float32, (G,)), ("idx", np.int64, (G,)), ("fold", np.int32, (G,)),
            ("xu8", np.uint8, (N, Fu)), ("xf32", np.float32, (N, Ff)), ("edges", _EDGE_DTYPES[ek], (E, 2)),
            ("xbits", np.uint8, (N, (Fb + 7) // 8))]


def _layout(G, N, E, F, Fu, Ff, ek, Fb=0):
    off, out = HEADER_BYTES, {}
    for name, dt, shape in _sections(G, N, E, F, Fu, Ff, ek, Fb):
        nbytes = int(np.prod(shape, dtype=np.int64)) * np.dtype(dt).itemsize
        out[name] = (off, dt, shape, nbytes)
        off = _aligned(off + nbytes)
    return out, off


class PackedDataset:
    """The packed image as typed numpy views over one byte buffer (in memory or memory-mapped)."""

    def __init__(self, blob: np.ndarray):
        assert blob.dtype == np.uint8 and blob.ndim == 1
        if bytes(blob[:8]) != MAGIC:
            raise ValueError("not a packed graph image (bad magic)")
        h = blob[8:8 + 8 * 8].view(np.int64)
        self.G, self.N, self.E, self.F, self.Fu, self.Ff, self.edge_kind, self.max_graph_nodes = (int(v) for v in h)
        h2 = blob[72:72 + 2 * 8].view(np.int64)
        # written by from_arrays (zero in images of older writers): size of the largest edge list, and whether the
        # packer verified what hcg_collate's graph-local path takes for granted (no self-loop edges; every graph
        # within its per-graph capacity).  Unverified images leave the choice to hcg_collate's own device check.
        self.max_graph_edges, self.flags = int(h2[0]), int(h2[1])
        # 1 + the largest in-degree of any node (0 in images of older writers: unknown)
        self.max_in_degree = int(blob[88:96].view(np.int64)[0]) - 1
        if self.max_in_degree < 0:
            self.max_in_degree = None
        self.Fb = int(blob[96:104].view(np.int64)[0])       # bit columns (0 in images of older writers)
        self.layout, total = _layout(self.G, self.N, self.E, self.F, self.Fu, self.Ff, self.edge_kind, self.Fb)
        if blob.size < total:
            raise ValueError(f"packed graph image truncated: {blob.size} of {total} bytes")
        self.blob = blob[:total]
        for name, (off, dt, shape, nbytes) in self.layout.items():
            setattr(self, name, self.blob[off:off + nbytes].view(dt).reshape(shape))

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_arrays(cls, x, edge_index, ptr, y=None, idx=None, fold=None, pinned: bool = False,
                    bits: bool = True) -> "PackedDataset":
        """From a collated set: ``x[N,F]`` f32, ``edge_index[2,E]`` with GLOBAL node numbers listed graph by
        graph (what ``Batch.from_data_list`` / PyG's collate produce), ``ptr[G+1]``."""
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
        ei = np.asarray(edge_index, dtype=np.int64)
        ptr = np.asarray(ptr, dtype=np.int64)
        G, N, E, F = ptr.size - 1, x.shape[0], ei.shape[1], x.shape[1]
        assert ptr[0] == 0 and ptr[-1] == N, "ptr does not cover x"
        gsrc = np.searchsorted(ptr, ei[0], side="right") - 1 if E else np.zeros(0, np.int64)
        gdst = np.searchsorted(ptr, ei[1], side="right") - 1 if E else np.zeros(0, np.int64)
        if E and (np.any(gsrc != gdst) or np.any(np.diff(gsrc) < 0) or ei.min() < 0 or ei.max() >= N):
            raise ValueError("edges must stay inside their graph and be listed graph by graph")
        edge_ptr = np.zeros(G + 1, np.int64)
        np.cumsum(np.bincount(gsrc, minlength=G), out=edge_ptr[1:])
        sizes = np.diff(ptr)
        nmax = int(sizes.max()) if G else 0
        ek = 1 if nmax <= 256 else 2 if nmax <= 65536 else 4
        # lossless column split: one bit (every value 0 or 1: one-hot / flag columns), one byte (integers 0..255), f32
        # (-0.0 compares equal to 0 but would come back as +0.0: such a column stays f32)
        as_bit = np.array([bool(np.all(((x[:, c] == 0) & ~np.signbit(x[:, c])) | (x[:, c] == 1))) for c in range(F)],
                          dtype=bool) if N else np.ones(F, bool)
        if not bits:
            as_bit[:] = False
        as_u8 = np.array([bool(np.all((x[:, c] >= 0) & (x[:, c] <= 255) & (x[:, c] == np.floor(x[:, c])) & ~np.signbit(x[:, c])))
                          for c in range(F)], dtype=bool) if N else np.zeros(F, bool)
        as_u8 &= ~as_bit
        as_f32 = ~(as_bit | as_u8)
        Fb, Fu, Ff = int(as_bit.sum()), int(as_u8.sum()), int(as_f32.sum())
        col_slot = np.empty(F, np.int32)
        col_slot[as_bit] = BIT_SLOT0 + np.arange(Fb, dtype=np.int32)
        col_slot[as_u8] = np.arange(Fu, dtype=np.int32)
        col_slot[as_f32] = ~np.arange(Ff, dtype=np.int32)      # negative: ~slot in the f32 block
        layout, total = _layout(G, N, E, F, Fu, Ff, ek, Fb)
        blob_t = torch.zeros(total, dtype=torch.uint8)
        if pinned:
            blob_t = blob_t.pin_memory()
        blob = blob_t.numpy()
        blob[:8] = np.frombuffer(MAGIC, np.uint8)
        blob[8:72].view(np.int64)[:] = [G, N, E, F, Fu, Ff, ek, nmax]
        emax = int(np.diff(edge_ptr).max()) if G else 0
        groupable = (not E or not bool(np.any(ei[0] == ei[1]))) and nmax <= GROUPED_MAX_NODES and emax <= GROUPED_MAX_EDGES
        blob[72:88].view(np.int64)[:] = [emax, FLAG_GROUPABLE if groupable else 0]
        blob[88:96].view(np.int64)[:] = [1 + (int(np.bincount(ei[1]).max()) if E else 0)]
        blob[96:104].view(np.int64)[:] = [Fb]
        self = cls(blob)
        self._owner = blob_t
        self.col_slot[:] = col_slot
        self.node_ptr[:] = ptr
        self.edge_ptr[:] = edge_ptr
        self.y[:] = 0 if y is None else np.asarray(y, np.float32).reshape(-1)
        self.idx[:] = np.arange(G) if idx is None else np.asarray(idx, np.int64).reshape(-1)
        self.fold[:] = -1 if fold is None else np.asarray(fold, np.int32).reshape(-1)
        self.xu8[:] = x[:, as_u8].astype(np.uint8)
        self.xf32[:] = x[:, as_f32]
        if Fb:
            self.xbits[:] = np.packbits(x[:, as_bit].astype(np.uint8), axis=1, bitorder="little")
        if E:
            self.edges[:, 0] = (ei[0] - ptr[gsrc]).astype(self.edges.dtype)
            self.edges[:, 1] = (ei[1] - ptr[gsrc]).astype(self.edges.dtype)
        return self

    @classmethod
    def from_graphs(cls, graphs: Sequence, pinned: bool = False) -> "PackedDataset":
        """From a list of per-graph records (``x``, ``edge_index`` local, optional ``y`` / ``idx`` / ``fold``)."""
        b = Batch.from_data_list(graphs)
        folds = [_field(g, "fold") for g in graphs]
        fold = None if any(f is None for f in folds) else [int(f) for f in folds]
        return cls.from_arrays(b.x.numpy(), b.edge_index.numpy(), b.ptr.numpy(), None if b.y is None else b.y.numpy(),
                               None if b.idx is None else b.idx.numpy(), fold, pinned=pinned)

    # ------------------------------------------------------------------ file I/O
    def save(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(self.blob.tobytes())

    @classmethod
    def load(cls, path: str, mmap: bool = True) -> "PackedDataset":
        return cls(np.memmap(path, dtype=np.uint8, mode="r") if mmap else np.fromfile(path, dtype=np.uint8))

    @property
    def nbytes(self) -> int:
        return int(self.blob.size)

    # ------------------------------------------------------------------ host decoder (format reference)
    def unpack_host(self, ids=None) -> Batch:
        """numpy decoder of the format: the batch ``Batch.from_data_list`` would build from graphs ``ids``."""
        ids = np.arange(self.G, dtype=np.int64) if ids is None else np.asarray(ids, dtype=np.int64).reshape(-1)
        n = self.node_ptr[ids + 1] - self.node_ptr[ids]
        e = self.edge_ptr[ids + 1] - self.edge_ptr[ids]
        ptr = np.zeros(ids.size + 1, np.int64)
        np.cumsum(n, out=ptr[1:])
        rows = np.concatenate([np.arange(self.node_ptr[g], self.node_ptr[g + 1]) for g in ids]) if ids.size else np.zeros(0, np.int64)
        slots = np.concatenate([np.arange(self.edge_ptr[g], self.edge_ptr[g + 1]) for g in ids]) if ids.size else np.zeros(0, np.int64)
        x = np.empty((rows.size, self.F), np.float32)
        bitc = self.col_slot >= BIT_SLOT0
        u8c = (self.col_slot >= 0) & ~bitc
        f32c = self.col_slot < 0
        x[:, u8c] = self.xu8[rows][:, self.col_slot[u8c]]
        x[:, f32c] = self.xf32[rows][:, ~self.col_slot[f32c]]
        if self.Fb:
            x[:, bitc] = np.unpackbits(self.xbits[rows], axis=1, bitorder="little")[:, self.col_slot[bitc] - BIT_SLOT0]
        ei = self.edges[slots].astype(np.int64).T + np.repeat(ptr[:-1], e)[None, :]
        batch = np.repeat(np.arange(ids.size, dtype=np.int64), n)
        return Batch(torch.from_numpy(x), torch.from_numpy(np.ascontiguousarray(ei)), torch.from_numpy(batch),
                     torch.from_numpy(ptr), torch.from_numpy(self.y[ids].copy()), torch.from_numpy(self.idx[ids].copy()),
                     ids.size, edges_grouped=True if (self.flags & FLAG_GROUPABLE) else None)


# ---------------------------------------------------------------------------------------------------
# converter from the reference's processed directory (one torch zip-pickle per reaction)
# ---------------------------------------------------------------------------------------------------
class _Stub:
    """Stand-in for the torch_geometric classes inside the pickles: keeps whatever state arrives."""

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        self.__dict__["_state"] = state
        if isinstance(state, dict):
            self.__dict__.update(state)


_PYG_CLASSES = {"torch_geometric.data.data": ("Data", "DataEdgeAttr", "DataTensorAttr"),
                "torch_geometric.data.storage": ("GlobalStorage", "BaseStorage", "NodeStorage", "EdgeStorage")}


def load_reaction_file(path: str) -> dict:
    """One ``reaction_<i>.pt`` -> its field dict (x, edge_index, y, idx, fold, ...), without torch_geometric:
    the pickles' only foreign globals are PyG's Data / storage classes, replaced by stubs while loading."""
    added = []
    try:
        for mod, names in _PYG_CLASSES.items():
            parts = mod.split(".")
            for k in range(1, len(parts) + 1):
                name = ".".join(parts[:k])
                if name not in sys.modules:
                    sys.modules[name] = types.ModuleType(name)
                    added.append(name)
            for n in names:
                if not hasattr(sys.modules[mod], n):
                    setattr(sys.modules[mod], n, type(n, (_Stub,), {"__module__": mod}))
        obj = torch.load(path, map_location="cpu", weights_only=False)
    finally:
        for name in added:
            sys.modules.pop(name, None)
    store = getattr(obj, "_store", obj)
    mapping = getattr(store, "_mapping", None)
    if mapping is None and isinstance(getattr(store, "_state", None), dict):
        mapping = store._state.get("_mapping")
    if not isinstance(mapping, dict):
        raise ValueError(f"{path}: no field mapping found in the pickled Data object")
    return mapping


def convert_processed_dir(processed_dir: str, out_path: str | None = None) -> PackedDataset:
    """All ``reaction_<i>.pt`` of a reference ``processed/`` directory, in index order (stray copies such as
    ``reaction_12 2.pt`` are skipped), into one packed image (optionally written to ``out_path``)."""
    files = []
    for f in glob.glob(os.path.join(processed_dir, "reaction_*.pt")):
        m = re.fullmatch(r"reaction_(\d+)\.pt", os.path.basename(f))
        if m:
            files.append((int(m.group(1)), f))
    if not files:
        raise FileNotFoundError(f"no reaction_<i>.pt under {processed_dir}")
    graphs = []
    for i, f in sorted(files):
        d = load_reaction_file(f)
        graphs.append({"x": d["x"], "edge_index": d["edge_index"], "y": d.get("y"), "idx": d.get("idx", i),
                       "fold": d.get("fold", -1)})
    ds = PackedDataset.from_graphs(graphs)
    if out_path:
        ds.save(out_path)
    return ds


# ---------------------------------------------------------------------------------------------------
# device side
# ---------------------------------------------------------------------------------------------------
class DeviceGraphStore:
    """A packed image resident in device memory + ``batch(ids)``: any selection of its graphs as a collated
    device ``Batch`` (one ``hcg_unpack_batch`` launch; the only host->device traffic is the id / offset lists).

    ``from_blob`` wraps an image that is ALREADY on the device (a chunk shipped with one ``copy_``), with the
    host copy supplying the sizes."""

    def __init__(self, ds: PackedDataset, device=None, blob_dev: torch.Tensor | None = None):
        self.ds = ds
        self.device = torch.device(device if device is not None else "cuda")
        _lib.load()     # raises if libhcg.so is missing: there is no host fallback for the loader
        if blob_dev is None:
            host = torch.from_numpy(np.ascontiguousarray(ds.blob))
            blob_dev = host.to(self.device)
        assert blob_dev.dtype == torch.uint8 and blob_dev.numel() >= ds.nbytes and blob_dev.data_ptr() % _ALIGN == 0
        self.blob_dev = blob_dev
        self._bind()

    @classmethod
    def from_blob(cls, ds: PackedDataset, blob_dev: torch.Tensor) -> "DeviceGraphStore":
        return cls(ds, blob_dev.device, blob_dev)

    def _bind(self):
        base = self.blob_dev.data_ptr()
        s = _lib.HcgPackedStore()
        s.G, s.N, s.E = self.ds.G, self.ds.N, self.ds.E
        s.F, s.Fu, s.Ff, s.edge_kind = self.ds.F, self.ds.Fu, self.ds.Ff, self.ds.edge_kind
        s.Fb, s.bit_stride = self.ds.Fb, (self.ds.Fb + 7) // 8
        for name in ("col_slot", "node_ptr", "edge_ptr", "y", "xu8", "xf32", "edges", "xbits"):
            setattr(s, name, base + self.ds.layout[name][0])
        self.c_store = s

    def rebind(self, ds: PackedDataset, blob_dev: torch.Tensor | None = None) -> "DeviceGraphStore":
        """Point the store at another image of the stream (same device buffer refilled, or a new one)."""
        self.ds = ds
        if blob_dev is not None:
            self.blob_dev = blob_dev
        self._bind()
        return self

    def batch(self, ids=None, with_batch_vector: bool = True) -> Batch:
        """``ids``: sequence / numpy / CPU tensor of graph numbers (repeats allowed); None = all, in order."""
        ds, dev = self.ds, self.device
        if ids is None:
            return self._batch_all(with_batch_vector)
        else:
            ids_np = np.ascontiguousarray(np.asarray(ids, dtype=np.int64).reshape(-1))
            if ids_np.size and (ids_np.min() < 0 or ids_np.max() >= ds.G):
                raise IndexError(f"graph id out of range [0, {ds.G})")
            n = ds.node_ptr[ids_np + 1] - ds.node_ptr[ids_np]
            e = ds.edge_ptr[ids_np + 1] - ds.edge_ptr[ids_np]
        B = int(n.size)
        meta = torch.empty(3, B + 1, dtype=torch.int64, pin_memory=dev.type == "cuda")
        m = meta.numpy()
        m[0, 0] = m[1, 0] = 0
        np.cumsum(n, out=m[0, 1:])
        np.cumsum(e, out=m[1, 1:])
        m[2, :B] = ids_np if ids_np is not None else 0
        N_out, E_out = int(m[0, B]), int(m[1, B])
        meta_d = meta.to(dev, non_blocking=True)
        return self._assemble(meta_d, ids_np, B, N_out, E_out, int(n.max()) if B else 0, int(e.max()) if B else 0,
                              with_batch_vector)

    def _assemble(self, meta_d, ids_np, B, N_out, E_out, nmax, emax, with_batch_vector=True) -> Batch:
        """The device half of ``batch(ids)``: ``meta_d`` int64 [3, B+1] on the device = (node offsets, edge offsets,
        graph ids) of the selection.  One hcg_unpack_batch launch."""
        ds, dev = self.ds, self.device
        x = ops.bucketed_empty((N_out, ds.F), torch.float32, dev)
        ei = ops.bucketed_empty((2, E_out), torch.int64, dev)
        bvec = ops.bucketed_empty((N_out,), torch.int64, dev) if with_batch_vector else None
        y = torch.empty(B, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().hcg_unpack_batch(
                C.byref(self.c_store), meta_d[2].data_ptr() if ids_np is not None else None, 0, B,
                meta_d[0].data_ptr(), meta_d[1].data_ptr(), x.data_ptr(), ei.data_ptr(), E_out,
                bvec.data_ptr() if bvec is not None else None, y.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        idx = torch.from_numpy(ds.idx[ids_np].copy() if ids_np is not None else np.array(ds.idx))
        out = Batch(x, ei, bvec, meta_d[0], y, idx, B, edges_grouped=True if (ds.flags & FLAG_GROUPABLE) else None)
        out.max_graph_nodes, out.max_graph_edges = nmax, emax
        out.max_in_degree = ds.max_in_degree
        out.edge_ptr = meta_d[1]
        out._keepalive = meta_d
        return out

    def epoch_batches(self, order, batch_size: int):
        """All mini-batches of one pass over ``order`` (graph ids, already shuffled): the offset / id tables of
        EVERY batch are built in one go and cross PCIe as one pinned copy; per batch only views are sliced and one
        kernel is launched.  (The reference's DataLoader does this work in Python per batch, reference
        call_methods.py:41-46.)"""
        ds, dev = self.ds, self.device
        order = np.ascontiguousarray(np.asarray(order, dtype=np.int64).reshape(-1))
        if order.size and (order.min() < 0 or order.max() >= ds.G):
            raise IndexError(f"graph id out of range [0, {ds.G})")
        n = ds.node_ptr[order + 1] - ds.node_ptr[order]
        e = ds.edge_ptr[order + 1] - ds.edge_ptr[order]
        starts = list(range(0, order.size, batch_size))
        width = batch_size + 1
        meta = torch.zeros(len(starts), 3, width, dtype=torch.int64, pin_memory=dev.type == "cuda")
        m = meta.numpy()
        sizes = []
        for k, s0 in enumerate(starts):
            B = min(batch_size, order.size - s0)
            np.cumsum(n[s0:s0 + B], out=m[k, 0, 1:B + 1])
            np.cumsum(e[s0:s0 + B], out=m[k, 1, 1:B + 1])
            m[k, 2, :B] = order[s0:s0 + B]
            sizes.append((B, int(m[k, 0, B]), int(m[k, 1, B]), int(n[s0:s0 + B].max()), int(e[s0:s0 + B].max())))
        meta_d = meta.to(dev, non_blocking=True)
        for k, s0 in enumerate(starts):
            B, N_out, E_out, nmax, emax = sizes[k]
            yield self._assemble(meta_d[k, :, :B + 1], order[s0:s0 + B], B, N_out, E_out, nmax, emax)   # (rows stay contiguous)


    def _ptr_views(self):
        """The image's own node_ptr / edge_ptr sections as device int64 tensors (no copy: views of the blob)."""
        def view(name):
            off, _dt, shape, nbytes = self.ds.layout[name]
            return self.blob_dev[off:off + nbytes].view(torch.int64)
        return view("node_ptr"), view("edge_ptr")

    def _idx_tensor(self) -> torch.Tensor:
        ds = self.ds
        idx = getattr(ds, "_idx_t", None)
        if idx is None:
            idx = torch.from_numpy(np.array(ds.idx))
            try:
                ds._idx_t = idx
            except AttributeError:
                pass
        return idx

    def lazy_batch(self) -> "LazyPackedBatch":
        """Every graph of the image, in order, unpacked only on demand (see ``LazyPackedBatch``)."""
        return LazyPackedBatch(self)

    def _batch_all(self, with_batch_vector: bool) -> Batch:
        """``batch(None)``: every graph of the image, in order (a wire-format chunk).  The exclusive prefix sums
        hcg_unpack_batch needs ARE the image's node_ptr / edge_ptr sections, already on the device; sizes come
        from the header.  Nothing is computed or allocated on the host per chunk (no pinned staging buffer, no
        cumulative sums): the chunk costs its one PCIe copy and one kernel."""
        ds, dev = self.ds, self.device
        B, N_out, E_out = ds.G, ds.N, ds.E
        node_ptr_d, edge_ptr_d = self._ptr_views()
        x = ops.bucketed_empty((N_out, ds.F), torch.float32, dev)
        ei = ops.bucketed_empty((2, E_out), torch.int64, dev)
        bvec = ops.bucketed_empty((N_out,), torch.int64, dev) if with_batch_vector else None
        y = torch.empty(B, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().hcg_unpack_batch(
                C.byref(self.c_store), None, 0, B, node_ptr_d.data_ptr(), edge_ptr_d.data_ptr(), x.data_ptr(), ei.data_ptr(),
                E_out, bvec.data_ptr() if bvec is not None else None, y.data_ptr(), torch.cuda.current_stream(dev).cuda_stream))
        idx = self._idx_tensor()
        out = Batch(x, ei, bvec, node_ptr_d, y, idx, B, edges_grouped=True if (ds.flags & FLAG_GROUPABLE) else None)
        out.max_graph_nodes = int(ds.max_graph_nodes)
        out.max_graph_edges = int(ds.max_graph_edges)
        out.max_in_degree = ds.max_in_degree
        out.edge_ptr = edge_ptr_d
        out._keepalive = self.blob_dev
        return out


class LazyPackedBatch(Batch):
    """Every graph of a device-resident packed image as a batch whose ``x`` / ``edge_index`` / ``y`` are only unpacked
    (``hcg_unpack_batch``) if somebody reads them: inference through ``GCN.forward`` runs straight from the image
    (``hcg_gcn_forward_packed``) and never does."""

    def __init__(self, store: "DeviceGraphStore"):
        ds = store.ds
        self._packed_store, self._real = store, None
        self.ptr, self.edge_ptr = store._ptr_views()
        self._batch = None
        self.num_graphs = int(ds.G)
        self.idx = store._idx_tensor()
        self.edges_grouped = True if (ds.flags & FLAG_GROUPABLE) else None
        self.max_graph_nodes, self.max_graph_edges = int(ds.max_graph_nodes), int(ds.max_graph_edges)
        self.max_in_degree = ds.max_in_degree
        self._keepalive = store.blob_dev

    def _materialise(self) -> Batch:
        if self._real is None:
            self._real = self._packed_store._batch_all(False)
        return self._real

    x = property(lambda self: self._materialise().x)
    edge_index = property(lambda self: self._materialise().edge_index)
    y = property(lambda self: self._materialise().y)

    @property
    def num_nodes(self) -> int:
        return int(self._packed_store.ds.N)

    def to(self, device, non_blocking: bool = False):
        dev = torch.device(device)
        if dev.type == "cuda" and (dev.index is None or dev.index == self.ptr.device.index):
            return self
        return self._materialise().to(device, non_blocking=non_blocking)

    def pin_memory(self):
        return self

    def __repr__(self):
        ds = self._packed_store.ds
        return f"LazyPackedBatch(num_graphs={ds.G}, nodes={ds.N}, edges={ds.E}, unpacked={self._real is not None})"


class StoreLoader:
    """``DataLoader`` over graphs of a ``DeviceGraphStore`` (reference call_methods.py:41-46 builds these with
    PyG's DataLoader over lists of Data): yields collated DEVICE batches assembled by ``store.batch``.
    ``len(loader.dataset)`` is the number of graphs, as the reference's loops expect
    (reference utils/utils_model.py:70,79).  Shuffling uses python ``random`` like ``data.DataLoader``.
    A loader that does not shuffle yields the same batches every epoch: they (and the collate plans the model
    caches on them) are assembled once and kept."""

    def __init__(self, store: DeviceGraphStore, ids, batch_size: int = 1, shuffle: bool = False, cache: bool | None = None):
        self.store = store
        self.dataset = [int(i) for i in ids]
        self.batch_size = int(batch_size)
        self.shuffle = bool(shuffle)
        self.cache = (not self.shuffle) if cache is None else bool(cache) and not self.shuffle
        self._kept = None

    def __len__(self):
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        if self._kept is not None:
            yield from self._kept
            return
        import random
        order = list(self.dataset)
        if self.shuffle:
            random.shuffle(order)
        kept = []
        for b in self.store.epoch_batches(order, self.batch_size):
            if self.cache:
                kept.append(b)
            yield b
        if self.cache:
            self._kept = kept


def create_store_loaders(store: DeviceGraphStore, opt):
    """The nested cross-validation generator of reference call_methods.py:15-47 over a device-resident store:
    for every outer fold (test set) and every remaining inner fold (validation set) yields
    ``(train_loader, val_loader, test_loader)``; graphs keep the reference's order inside each fold and the
    training set is the remaining folds concatenated in fold order, reshuffled every epoch."""
    n_folds = int(opt.folds)
    fold = np.asarray(store.ds.fold)
    if fold.size and (fold.min() < 0 or fold.max() >= n_folds):
        raise ValueError(f"fold labels outside [0, {n_folds}) in the packed image")
    folds = [np.nonzero(fold == f)[0].tolist() for f in range(n_folds)]
    for outer in range(n_folds):
        proxy = folds[:outer] + folds[outer + 1:]
        for inner in range(len(proxy)):
            rest = proxy[:inner] + proxy[inner + 1:]
            train = [i for sub in rest for i in sub]
            yield (StoreLoader(store, train, opt.batch_size, shuffle=True),
                   StoreLoader(store, proxy[inner], opt.batch_size, shuffle=False),
                   StoreLoader(store, folds[outer], opt.batch_size, shuffle=False))


if __name__ == "__main__":  # python -m hcatgnet_b200.packed <processed_dir> <out.hcgp>
    if len(sys.argv) != 3:
        raise SystemExit("usage: python -m hcatgnet_b200.packed <reference processed/ directory> <out.hcgp>")
    _ds = convert_processed_dir(sys.argv[1], sys.argv[2])
    print(f"{_ds.G} graphs, {_ds.N} nodes, {_ds.E} directed edges, F={_ds.F} ({_ds.Fu} u8 + {_ds.Ff} f32 columns), "
          f"{_ds.edge_kind}-byte endpoints -> {_ds.nbytes} bytes")
