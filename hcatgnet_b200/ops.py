"""Host-side glue between torch tensors and the C-ABI (include/hcg.h).

Nothing here computes: tensors are only allocated, their device pointers handed to libhcg.so
on torch's current CUDA stream.  Inputs living in host memory are staged to the GPU first
(that is the end-to-end path bench.py times); without a CUDA device every call raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch

from . import _lib

LEAKY_SLOPE = 0.01  # nn.LeakyReLU() default used throughout reference model/gcn.py


def _require_cuda() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("hcatgnet_b200 needs a CUDA device (B200): there is no CPU implementation of the GCN path")
    return torch.device("cuda", torch.cuda.current_device())


def compute_device(t: torch.Tensor) -> torch.device:
    """The GPU the work runs on: the tensor's own device if it is a CUDA tensor, else the
    current CUDA device (host tensors get staged there)."""
    return t.device if t.is_cuda else _require_cuda()


def to_device(t: torch.Tensor | None, device: torch.device, dtype=None) -> torch.Tensor | None:
    if t is None:
        return None
    if dtype is not None and t.dtype != dtype:
        t = t.to(dtype)
    if t.device != device:
        t = t.to(device, non_blocking=True)
    return t.contiguous()


def _ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def _stream(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def bucketed_empty(shape, dtype, device) -> torch.Tensor:
    """``torch.empty(shape)`` whose allocation is rounded up to one of eight sizes per power of two.  Batches of a
    stream differ in size by a few per cent; with exact sizes torch's caching allocator misses on most of them (a
    cudaMalloc of 100+ MB per tensor and step until every size has been seen), with buckets the second batch already
    reuses the first one's blocks.  At most 12.5 % of slack, never touched."""
    shape = tuple(int(v) for v in shape)
    numel = 1
    for v in shape:
        numel *= v
    if numel < (1 << 16):
        return torch.empty(shape, dtype=dtype, device=device)
    step = 1 << (numel.bit_length() - 4)
    cap = (numel + step - 1) // step * step
    return torch.empty(cap, dtype=dtype, device=device)[:numel].view(shape)


def _workspace(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


@dataclass
class GraphPlan:
    """Device-resident batch topology (the HcgGraph of include/hcg.h) built once per batch."""
    N: int
    E: int
    B: int
    rowptr: torch.Tensor
    col: torch.Tensor
    eid: torch.Tensor
    dis: torch.Tensor
    graph_ptr: torch.Tensor
    fill: float = 1.0
    t_rowptr: torch.Tensor | None = None
    t_col: torch.Tensor | None = None
    t_eid: torch.Tensor | None = None
    ew0: torch.Tensor | None = None
    t_ew0: torch.Tensor | None = None
    selfw0: torch.Tensor | None = None
    dis0: torch.Tensor | None = None
    flags: int = 0             # HCG_GRAPH_* bits reported by hcg_collate_status
    max_graph_nodes: int = 0   # hint (0 = unknown): sizes the smem capacity of the graph-resident kernel
    max_graph_edges: int = 0   # hint (0 = unknown): edge count of the largest graph
    _keep: list = field(default_factory=list, repr=False)

    @property
    def has_transpose(self) -> bool:
        return self.t_rowptr is not None

    @property
    def device(self):
        return self.rowptr.device

    def c_struct(self, edge_mask: torch.Tensor | None = None) -> _lib.HcgGraph:
        return _lib.HcgGraph(
            self.N, self.E, self.B, _ptr(self.rowptr), _ptr(self.col), _ptr(self.eid), _ptr(self.dis),
            _ptr(self.graph_ptr), _ptr(self.t_rowptr), _ptr(self.t_col), _ptr(self.t_eid), self.fill,
            _ptr(self.ew0), _ptr(self.t_ew0), _ptr(self.selfw0), _ptr(self.dis0), int(self.max_graph_nodes), int(self.flags),
            _ptr(edge_mask), int(self.max_graph_edges), 0)


def collate(edge_index: torch.Tensor, batch: torch.Tensor | None, num_nodes: int, num_graphs: int,
            edge_weight: torch.Tensor | None = None, improved: bool = False, transpose: bool = False,
            fill: float = 1.0, validate: bool = True, max_graph_nodes: int = 0,
            grouped: bool | None = None, ptr: torch.Tensor | None = None, edge_ptr: torch.Tensor | None = None,
            max_graph_edges: int = 0) -> GraphPlan:
    """COO ``edge_index`` [2,E] int64 + sorted ``batch`` [N] int64 -> :class:`GraphPlan`.

    Runs ``hcg_collate``.  ``improved`` only matters together with ``edge_weight`` (self-loop
    fill 2.0), mirroring the behaviour of the reference's PyG path.
    ``validate`` reads back one status word (a stream sync) and raises on bad indices / unsorted
    batch; the kernels are memory-safe either way.
    ``grouped`` selects the graph-local fast path of ``hcg_collate`` (HCG_COLLATE_GROUPED: edges listed
    graph by graph, as ``Batch.from_data_list`` / PyG's collate produce them; same outputs bit for bit):
    ``True`` = the caller vouches for it, ``None`` = try it when the status word is read back anyway
    (``validate``) and fall back to the general kernels if the device check refuses, ``False`` = never.
    ``ptr`` / ``edge_ptr`` (int64 [B+1]: first node / first edge of every graph, as PyG's ``Batch.ptr`` and
    ``Batch._slice_dict['edge_index']`` hold them): with both, the graph-local path takes its offsets from
    them (``hcg_collate_offsets``) instead of deriving them from ``batch`` and searching the edge list.
    """
    device = _require_cuda() if not edge_index.is_cuda else edge_index.device
    lib = _lib.load()
    if edge_index.dim() != 2 or edge_index.shape[0] != 2:
        raise ValueError(f"edge_index must have shape [2, E], got {tuple(edge_index.shape)}")
    ei = to_device(edge_index, device, torch.int64)
    bt = to_device(batch, device, torch.int64)
    ew = to_device(edge_weight, device, torch.float32)
    N, E, B = int(num_nodes), int(ei.shape[1]), int(num_graphs)
    if bt is not None and bt.numel() != N:
        raise ValueError(f"batch has {bt.numel()} entries for {N} nodes")
    if ew is not None and ew.numel() != E:
        raise ValueError(f"edge_weight has {ew.numel()} entries for {E} edges")
    if bt is None and B > 1 and ptr is None:
        raise ValueError("batch vector (or ptr) required when num_graphs > 1")
    i32 = dict(dtype=torch.int32, device=device)
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        plan = GraphPlan(N, E, B, torch.empty(N + 1, **i32), torch.empty(max(E, 1), **i32),
                         torch.empty(max(E, 1), **i32), torch.empty(max(N, 1), **f32), torch.empty(B + 1, **i32),
                         fill=float(fill), max_graph_nodes=int(max_graph_nodes), max_graph_edges=int(max_graph_edges))
        flags = 0
        have_offsets = ptr is not None and edge_ptr is not None and ptr.numel() == B + 1 and edge_ptr.numel() == B + 1
        use_grouped = (bt is not None or have_offsets) and ew is None and B > 0 and N > 0 and E > 0 and \
            (grouped is True or (grouped is None and validate))
        if use_grouped:
            flags |= _lib.HCG_COLLATE_GROUPED
        if transpose:
            flags |= _lib.HCG_COLLATE_TRANSPOSE
            plan.t_rowptr = torch.empty(N + 1, **i32)
            plan.t_col = torch.empty(max(E, 1), **i32)
            plan.t_eid = torch.empty(max(E, 1), **i32)
        if ew is not None:
            plan.ew0 = torch.empty(max(E, 1), **f32)
            plan.selfw0 = torch.empty(max(N, 1), **f32)
            plan.dis0 = torch.empty(max(N, 1), **f32)
            if transpose:
                plan.t_ew0 = torch.empty(max(E, 1), **f32)
        ws = _workspace(lib.hcg_collate_workspace_bytes(N, E, B, flags), device)
        st = _stream(device)
        if use_grouped and have_offsets:
            ptr_d, eptr_d = to_device(ptr, device, torch.int64), to_device(edge_ptr, device, torch.int64)
            _lib.check(lib.hcg_collate_offsets(
                _ptr(ei), E, _ptr(ptr_d), _ptr(eptr_d), N, B, float(fill), flags & _lib.HCG_COLLATE_TRANSPOSE,
                int(max_graph_nodes), int(max_graph_edges),
                _ptr(plan.rowptr), _ptr(plan.col), _ptr(plan.eid), _ptr(plan.dis), _ptr(plan.graph_ptr),
                _ptr(plan.t_rowptr), _ptr(plan.t_col), _ptr(plan.t_eid), _ptr(ws), ws.numel(), st))
            plan._keep += [ptr_d, eptr_d]
        else:
            if bt is None and B > 1:       # the general kernels work from the N-entry batch vector
                bt = batch_from_ptr(to_device(ptr, device, torch.int64), N)
            _lib.check(lib.hcg_collate(
                _ptr(ei), E, _ptr(bt), N, B, float(fill), _ptr(ew), 2.0 if improved else 1.0, flags,
                _ptr(plan.rowptr), _ptr(plan.col), _ptr(plan.eid), _ptr(plan.dis), _ptr(plan.graph_ptr),
                _ptr(plan.t_rowptr), _ptr(plan.t_col), _ptr(plan.t_eid),
                _ptr(plan.ew0), _ptr(plan.t_ew0), _ptr(plan.selfw0), _ptr(plan.dis0),
                _ptr(ws), ws.numel(), st))
        if validate:
            gflags = C.c_int32(0)
            rc = lib.hcg_collate_status(_ptr(ws), st, C.byref(gflags))
            if rc == _lib.HCG_ERR_NOT_GROUPED and use_grouped:  # the device check refused the fast path
                if bt is None and B > 1:
                    bt = batch_from_ptr(to_device(ptr, device, torch.int64), N)
                return collate(ei, bt, N, B, edge_weight=ew, improved=improved, transpose=transpose, fill=fill,
                               validate=True, max_graph_nodes=max_graph_nodes, grouped=False)   # (general kernels: no offsets)
            _lib.check(rc)
            plan.flags = int(gflags.value)
        else:
            ws.record_stream(torch.cuda.current_stream(device))
    return plan


def batch_from_ptr(ptr: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """PyG's ``Batch.batch`` [N] rebuilt on the GPU from ``Batch.ptr`` [B+1] (``hcg_batch_from_ptr``)."""
    if not ptr.is_cuda:
        raise RuntimeError("batch_from_ptr expects the ptr tensor on a CUDA device")
    ptr = ptr.to(torch.int64).contiguous()
    out = torch.empty(int(num_nodes), dtype=torch.int64, device=ptr.device)
    with torch.cuda.device(ptr.device):
        _lib.check(_lib.load().hcg_batch_from_ptr(_ptr(ptr), int(ptr.numel()) - 1, int(num_nodes), _ptr(out), _stream(ptr.device)))
    return out


def batch_vector_if_built(holder):
    """``holder.batch`` unless the holder is one of our lazily-built batches that carries ``ptr`` / ``edge_ptr`` and has
    not built its N-entry batch vector yet (the collate kernels then work from the offsets: no need to build it)."""
    if hasattr(holder, "_batch") and getattr(holder, "ptr", None) is not None and getattr(holder, "edge_ptr", None) is not None:
        return holder._batch
    return holder.batch


def cached_collate(holder, edge_index, batch, num_nodes, num_graphs, device, edge_weight=None,
                   improved=False, transpose=False, validate=True, max_graph_nodes=0, grouped=None) -> GraphPlan:
    # graph offsets the batch object carries (ours do; so does PyG's Batch): spare hcg_collate the search for them
    ptr, edge_ptr = getattr(holder, "ptr", None), getattr(holder, "edge_ptr", None)
    max_graph_edges = int(getattr(holder, "max_graph_edges", 0) or 0)
    return _cached_collate(holder, edge_index, batch, num_nodes, num_graphs, device, edge_weight, improved, transpose,
                           validate, max_graph_nodes, grouped, ptr, edge_ptr, max_graph_edges)


def _cached_collate(holder, edge_index, batch, num_nodes, num_graphs, device, edge_weight, improved, transpose, validate,
                    max_graph_nodes, grouped, ptr, edge_ptr, max_graph_edges) -> GraphPlan:
    """``collate`` memoised on ``holder`` (the batch object the caller passed to ``forward``).

    The reference recomputes the normalisation inside every GCNConv call of every forward
    (``cached=False``); here the topology is built once per batch object and reused by all
    layers, by the backward pass and by later epochs that present the same batch again.
    Weighted calls (explainer masks change every step) are never cached.
    """
    key = None
    if holder is not None and edge_weight is None:
        # Identity of the caller's tensors, not their addresses: the cache entry keeps references to them, so
        # the allocator cannot hand the same storage to another graph's edge_index while the entry is alive
        # (two same-shape molecules explained back to back would otherwise share one plan).
        key = (edge_index._version, None if batch is None else batch._version, int(num_nodes), int(num_graphs),
               str(device))
        hit = getattr(holder, "_hcg_plan", None)
        if hit is not None and hit[0] == key and hit[2] is edge_index and hit[3] is batch and \
                (hit[1].has_transpose or not transpose):
            return hit[1]
    with torch.cuda.device(device):
        plan = collate(to_device(edge_index, device, torch.int64), to_device(batch, device, torch.int64),
                       num_nodes, num_graphs, edge_weight=to_device(edge_weight, device, torch.float32),
                       improved=improved, transpose=transpose, validate=validate,
                       max_graph_nodes=max_graph_nodes, grouped=grouped, ptr=ptr, edge_ptr=edge_ptr,
                       max_graph_edges=max_graph_edges)
    if key is not None:
        try:
            object.__setattr__(holder, "_hcg_plan", (key, plan, edge_index, batch))
        except (AttributeError, TypeError):
            pass
    return plan


@dataclass
class NetDims:
    F: int
    H: int
    L: int
    R: int
    C: int
    slope: float = LEAKY_SLOPE

    def ro_in(self, k: int) -> int:
        d = 2 * self.H
        for _ in range(k):
            d //= 2
        return d

    def ro_out(self, k: int) -> int:
        return self.C if k == self.R - 1 else self.ro_in(k) // 2


def _params_struct(dims: NetDims, tensors) -> _lib.HcgParams:
    """tensors: conv_w0, conv_b0, ..., ro_w0, ro_b0, ... (contiguous fp32 on the GPU)."""
    p = _lib.HcgParams()
    p.F, p.H, p.L, p.R, p.C, p.slope = dims.F, dims.H, dims.L, dims.R, dims.C, dims.slope
    for l in range(dims.L):
        p.conv_w[l] = tensors[2 * l].data_ptr()
        p.conv_b[l] = tensors[2 * l + 1].data_ptr()
    for k in range(dims.R):
        p.ro_w[k] = tensors[2 * dims.L + 2 * k].data_ptr()
        p.ro_b[k] = tensors[2 * dims.L + 2 * k + 1].data_ptr()
    return p


def _check_param_shapes(dims: NetDims, tensors) -> None:
    if len(tensors) != 2 * (dims.L + dims.R):
        raise ValueError(f"expected {2 * (dims.L + dims.R)} parameter tensors, got {len(tensors)}")
    for l in range(dims.L):
        want = (dims.H, dims.F if l == 0 else dims.H)
        if tuple(tensors[2 * l].shape) != want or tuple(tensors[2 * l + 1].shape) != (dims.H,):
            raise ValueError(f"conv layer {l}: weight {tuple(tensors[2 * l].shape)} / bias "
                             f"{tuple(tensors[2 * l + 1].shape)} do not match {want} / {(dims.H,)}")
    for k in range(dims.R):
        w, b = tensors[2 * dims.L + 2 * k], tensors[2 * dims.L + 2 * k + 1]
        want = (dims.ro_out(k), dims.ro_in(k))
        if tuple(w.shape) != want or tuple(b.shape) != (want[0],):
            raise ValueError(f"readout layer {k}: weight {tuple(w.shape)} / bias {tuple(b.shape)} do not match {want}")


def _check_mask(plan: GraphPlan, edge_mask):
    if edge_mask is not None and (edge_mask.numel() != plan.E or edge_mask.dtype != torch.float32 or not edge_mask.is_cuda):
        raise ValueError(f"edge_mask must be a CUDA float32 tensor with one entry per edge ({plan.E})")


def forward_raw(plan: GraphPlan, x: torch.Tensor, dims: NetDims, params, save: bool, edge_mask=None):
    """One ``hcg_gcn_forward`` call.  Returns (out[B,C], emb[B,2H], saved tensors or None).
    ``edge_mask``: optional [E] multiplier of every message of every conv layer (original edge order)."""
    _check_mask(plan, edge_mask)
    lib = _lib.load()
    device = plan.device
    if x.shape != (plan.N, dims.F):
        raise ValueError(f"x has shape {tuple(x.shape)}, expected {(plan.N, dims.F)}")
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        out = torch.empty(plan.B, dims.C, **f32)
        emb = torch.empty(plan.B, 2 * dims.H, **f32)
        pstruct = _params_struct(dims, params)
        saved_t = None
        saved_c = None
        if save:
            acts = [torch.empty(plan.N, dims.H, **f32) for _ in range(dims.L)]
            hid = [torch.empty(plan.B, dims.ro_out(k), **f32) for k in range(dims.R - 1)]
            saved_t = (acts, hid)
            saved_c = _lib.HcgSaved()
            for l, a in enumerate(acts):
                saved_c.acts[l] = a.data_ptr()
            for k, h in enumerate(hid):
                saved_c.hid[k] = h.data_ptr()
        ws = _workspace(lib.hcg_forward_workspace_bytes(plan.N, plan.E, plan.B, C.byref(pstruct)), device)
        g = plan.c_struct(edge_mask)
        _lib.check(lib.hcg_gcn_forward(C.byref(g), _ptr(x), C.byref(pstruct), _ptr(out), _ptr(emb),
                                       C.byref(saved_c) if saved_c is not None else None,
                                       _ptr(ws), ws.numel(), _stream(device)))
        ws.record_stream(torch.cuda.current_stream(device))
    return out, emb, saved_t


# capacity of hcg_gcn_forward_coo (include/hcg.h): nodes / edges per graph, incoming edges per node
COO_MAX_NODES, COO_MAX_EDGES, COO_MAX_IN_DEGREE = 128, 512, 16


def edge_list_hints_ok(holder) -> bool:
    """The batch object vouches (host-side integers, no device read) for what ``hcg_gcn_forward_coo`` requires."""
    if holder is None or getattr(holder, "edges_grouped", None) is not True:
        return False
    if getattr(holder, "ptr", None) is None or getattr(holder, "edge_ptr", None) is None:
        return False
    nmax, emax = int(getattr(holder, "max_graph_nodes", 0) or 0), int(getattr(holder, "max_graph_edges", 0) or 0)
    dmax = getattr(holder, "max_in_degree", None)
    return 0 < nmax <= COO_MAX_NODES and emax <= COO_MAX_EDGES and dmax is not None and 0 <= int(dmax) <= COO_MAX_IN_DEGREE


def forward_edge_list(x: torch.Tensor, edge_index: torch.Tensor, ptr: torch.Tensor, edge_ptr: torch.Tensor, dims: NetDims,
                      params, fill: float = 1.0, validate: bool = True):
    """Inference straight from the batch's edge list: one ``hcg_gcn_forward_coo`` call, no collate, no ``GraphPlan``
    (the conv kernel builds each graph's normalised neighbour lists in shared memory).  Same bits as
    ``collate`` + ``forward_raw``.  Returns ``(out, emb)``, or ``None`` when the network's shape cannot take the path
    or - with ``validate`` (one small read-back, a stream sync) - when the device refused the batch (oversize graph,
    in-degree above 16, self-loop, edge between graphs): the caller then goes through ``collate``.  Without
    ``validate`` the caller vouches for the batch (``edge_list_hints_ok``)."""
    lib = _lib.load()
    device = x.device
    _check_param_shapes(dims, params)
    pstruct = _params_struct(dims, params)
    if not lib.hcg_forward_coo_supported(C.byref(pstruct)):
        return None
    if edge_index.dim() != 2 or edge_index.shape[0] != 2 or edge_index.dtype != torch.int64 or not edge_index.is_contiguous():
        raise ValueError("edge_index must be a contiguous int64 [2, E] tensor")
    B, N, E = int(ptr.numel()) - 1, int(x.shape[0]), int(edge_index.shape[1])
    if x.shape != (N, dims.F) or x.dtype != torch.float32 or not x.is_contiguous():
        raise ValueError(f"x must be a contiguous float32 [N, {dims.F}] tensor")
    if edge_ptr.numel() != B + 1 or ptr.dtype != torch.int64 or edge_ptr.dtype != torch.int64:
        raise ValueError("ptr / edge_ptr must be int64 [B + 1]")
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        out = torch.empty(B, dims.C, **f32)
        emb = torch.empty(B, 2 * dims.H, **f32)
        ws = _workspace(lib.hcg_forward_coo_workspace_bytes(B, C.byref(pstruct)), device)
        st = _stream(device)
        _lib.check(lib.hcg_gcn_forward_coo(_ptr(x), _ptr(edge_index), _ptr(ptr), _ptr(edge_ptr), N, E, B, float(fill),
                                           C.byref(pstruct), _ptr(out), _ptr(emb), _ptr(ws), ws.numel(), st))
        ws.record_stream(torch.cuda.current_stream(device))
        if validate:
            flags = C.c_int32(0)
            _lib.check(lib.hcg_forward_coo_status(_ptr(ws), B, st, C.byref(flags)))
            if flags.value:
                return None
    return out, emb


def forward_packed(c_store, first: int, B: int, dims: NetDims, params, device, fill: float = 1.0, validate: bool = True):
    """Inference on graphs ``[first, first + B)`` of a packed image resident on ``device`` (``hcg_gcn_forward_packed``: no
    unpack, no collate; x and edge_index never exist in HBM).  Returns ``(out, emb)`` or ``None`` when the image / network
    cannot take the path or - with ``validate`` - the device refused the graphs: unpack and go through the other paths."""
    lib = _lib.load()
    _check_param_shapes(dims, params)
    pstruct = _params_struct(dims, params)
    if not lib.hcg_forward_packed_supported(C.byref(c_store), C.byref(pstruct)):
        return None
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        out = torch.empty(B, dims.C, **f32)
        emb = torch.empty(B, 2 * dims.H, **f32)
        ws = _workspace(lib.hcg_forward_coo_workspace_bytes(B, C.byref(pstruct)), device)
        st = _stream(device)
        _lib.check(lib.hcg_gcn_forward_packed(C.byref(c_store), int(first), int(B), float(fill), C.byref(pstruct), _ptr(out), _ptr(emb),
                                              _ptr(ws), ws.numel(), st))
        ws.record_stream(torch.cuda.current_stream(device))
        if validate:
            flags = C.c_int32(0)
            _lib.check(lib.hcg_forward_coo_status(_ptr(ws), B, st, C.byref(flags)))
            if flags.value:
                return None
    return out, emb


class EnsembleWeights:
    """The conv / readout weights of M same-architecture networks stacked along a new first axis (what
    ``hcg_gcn_forward_coo_ensemble`` takes).  Built once per ensemble from the models' flat parameter lists
    (``_GCNBase._flat_params`` order: conv1 w, b, conv2 w, b, readout0 w, b, readout1 w, b)."""

    def __init__(self, dims: NetDims, params_per_model, device):
        if dims.L != 2 or dims.R != 2 or dims.H != 64 or dims.F > 32:
            raise ValueError("one-pass ensemble needs the default architecture (2 conv layers, H = 64, readout 128-64-C, F <= 32)")
        for ps in params_per_model:
            _check_param_shapes(dims, ps)
        self.dims, self.M = dims, len(params_per_model)
        self.stacked = [torch.stack([to_device(ps[k].detach(), device, torch.float32) for ps in params_per_model]).contiguous()
                        for k in range(8)]

    @staticmethod
    def supported(dims: NetDims) -> bool:
        return dims.L == 2 and dims.R == 2 and dims.H == 64 and dims.F <= 32 and dims.C <= 64 and 0.0 <= dims.slope <= 1.0

    def c_struct(self) -> _lib.HcgEnsemble:
        e = _lib.HcgEnsemble()
        e.M, e.F, e.H, e.C, e.slope = self.M, self.dims.F, self.dims.H, self.dims.C, float(self.dims.slope)
        (e.conv_w0, e.conv_b0, e.conv_w1, e.conv_b1, e.ro_w0, e.ro_b0, e.ro_w1, e.ro_b1) = (t.data_ptr() for t in self.stacked)
        return e


def forward_edge_list_ensemble(x, edge_index, ptr, edge_ptr, weights: EnsembleWeights, fill: float = 1.0, validate: bool = True):
    """M networks over one batch in ONE pass (``hcg_gcn_forward_coo_ensemble``): returns ``(out[M,B,C], emb[M,B,2H])``
    or ``None`` when the device refused the batch (``validate``; see ``forward_edge_list``).  Per model the same bits as
    ``forward_edge_list``."""
    lib = _lib.load()
    device, dims = x.device, weights.dims
    if edge_index.dim() != 2 or edge_index.shape[0] != 2 or edge_index.dtype != torch.int64 or not edge_index.is_contiguous():
        raise ValueError("edge_index must be a contiguous int64 [2, E] tensor")
    B, N, E = int(ptr.numel()) - 1, int(x.shape[0]), int(edge_index.shape[1])
    if x.shape != (N, dims.F) or x.dtype != torch.float32 or not x.is_contiguous():
        raise ValueError(f"x must be a contiguous float32 [N, {dims.F}] tensor")
    if edge_ptr.numel() != B + 1 or ptr.dtype != torch.int64 or edge_ptr.dtype != torch.int64:
        raise ValueError("ptr / edge_ptr must be int64 [B + 1]")
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        out = torch.empty(weights.M, B, dims.C, **f32)
        emb = torch.empty(weights.M, B, 2 * dims.H, **f32)
        ws = _workspace(lib.hcg_ensemble_workspace_bytes(B, weights.M), device)
        st = _stream(device)
        e = weights.c_struct()
        _lib.check(lib.hcg_gcn_forward_coo_ensemble(_ptr(x), _ptr(edge_index), _ptr(ptr), _ptr(edge_ptr), N, E, B, float(fill),
                                                    C.byref(e), _ptr(out), _ptr(emb), _ptr(ws), ws.numel(), st))
        ws.record_stream(torch.cuda.current_stream(device))
        if validate:
            flags = C.c_int32(0)
            _lib.check(lib.hcg_forward_coo_status(_ptr(ws), B, st, C.byref(flags)))
            if flags.value:
                return None
    return out, emb


def backward_raw(plan: GraphPlan, x: torch.Tensor, dims: NetDims, params, saved_t, emb, g_out, g_emb, need_gx: bool,
                 flat_out: torch.Tensor | None = None, edge_mask=None, need_gmask: bool = False):
    """One ``hcg_gcn_backward`` call.  Returns (g_x or None, list of parameter grads) — and the gradient
    w.r.t. ``edge_mask`` as a third element when ``need_gmask``.
    ``flat_out``: an existing fp32 slab (>= sum of parameter sizes) to write the gradients into."""
    _check_mask(plan, edge_mask)
    lib = _lib.load()
    device = plan.device
    f32 = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        pstruct = _params_struct(dims, params)
        acts, hid = saved_t
        saved_c = _lib.HcgSaved()
        for l, a in enumerate(acts):
            saved_c.acts[l] = a.data_ptr()
        for k, h in enumerate(hid):
            saved_c.hid[k] = h.data_ptr()
        # one flat slab for all parameter gradients (a single all-reduce buffer for data parallelism)
        sizes = [p.numel() for p in params]
        flat = flat_out if flat_out is not None else torch.empty(sum(sizes), **f32)
        grads, off = [], 0
        for p_, n in zip(params, sizes):
            grads.append(flat[off:off + n].view_as(p_))
            off += n
        gc = _lib.HcgGrads()
        for l in range(dims.L):
            gc.conv_w[l] = grads[2 * l].data_ptr()
            gc.conv_b[l] = grads[2 * l + 1].data_ptr()
        for k in range(dims.R):
            gc.ro_w[k] = grads[2 * dims.L + 2 * k].data_ptr()
            gc.ro_b[k] = grads[2 * dims.L + 2 * k + 1].data_ptr()
        g_x = torch.empty(plan.N, dims.F, **f32) if need_gx else None
        g_mask = None
        if need_gmask and edge_mask is not None:
            g_mask = torch.empty(max(plan.E, 1), **f32)
            gc.edge_mask = g_mask.data_ptr()
        ws = _workspace(lib.hcg_backward_workspace_bytes(plan.N, plan.E, plan.B, C.byref(pstruct)), device)
        g = plan.c_struct(edge_mask)
        _lib.check(lib.hcg_gcn_backward(C.byref(g), _ptr(x), C.byref(pstruct), C.byref(saved_c), _ptr(emb),
                                        _ptr(g_out), _ptr(g_emb), C.byref(gc), _ptr(g_x),
                                        _ptr(ws), ws.numel(), _stream(device)))
        ws.record_stream(torch.cuda.current_stream(device))
    if need_gmask:
        return g_x, grads, (None if g_mask is None else g_mask[:plan.E])
    return g_x, grads


class _GCNFunction(torch.autograd.Function):
    """autograd node whose forward/backward are single C-ABI calls."""

    @staticmethod
    def forward(ctx, x, plan, dims, need_grad, edge_mask, *params):
        # `need_grad` is decided by the caller: grad mode is always off inside Function.forward and
        # ctx.needs_input_grad reports requires_grad flags even under torch.no_grad()
        params = tuple(p.contiguous() for p in params)
        if need_grad and not plan.has_transpose:
            raise RuntimeError("this GraphPlan was collated without the transposed CSR; backward needs transpose=True")
        if edge_mask is not None:
            edge_mask = edge_mask.contiguous()
        out, emb, saved_t = forward_raw(plan, x, dims, params, save=need_grad, edge_mask=edge_mask)
        if need_grad:
            ctx.plan, ctx.dims, ctx.saved_t = plan, dims, saved_t
            ctx.has_mask = edge_mask is not None
            if edge_mask is not None:
                ctx.save_for_backward(x, emb, edge_mask, *params)
            else:
                ctx.save_for_backward(x, emb, *params)
        return out, emb

    @staticmethod
    def backward(ctx, g_out, g_emb):
        if ctx.has_mask:
            x, emb, edge_mask, *params = ctx.saved_tensors
        else:
            x, emb, *params = ctx.saved_tensors
            edge_mask = None
        g_out = g_out.contiguous() if g_out is not None else None
        g_emb = g_emb.contiguous() if g_emb is not None else None
        want_gmask = bool(ctx.has_mask and ctx.needs_input_grad[4])
        res = backward_raw(ctx.plan, x, ctx.dims, params, ctx.saved_t, emb, g_out, g_emb,
                           need_gx=ctx.needs_input_grad[0], edge_mask=edge_mask, need_gmask=want_gmask)
        g_x, grads = res[0], res[1]
        g_mask = res[2] if want_gmask else None
        return (g_x, None, None, None, g_mask, *grads)


def gcn_apply(x: torch.Tensor, plan: GraphPlan, dims: NetDims, params, edge_mask: torch.Tensor | None = None):
    """Differentiable GCN hot path: returns (out[B,C], emb[B,2H]).  ``edge_mask`` ([E], original edge order)
    multiplies every message of every conv layer and is differentiable (explainers)."""
    _check_param_shapes(dims, params)
    need_grad = torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in params) or
                                             (edge_mask is not None and edge_mask.requires_grad))
    return _GCNFunction.apply(x, plan, dims, need_grad, edge_mask, *params)
