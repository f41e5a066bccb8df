"""Streaming screening / prediction driver: the working counterpart of the reference's
``scripts_experiments/predict.py:11-60`` (dead code there) and of ``predict_network``
(reference utils/utils_model.py:82-111), for batches that live in HOST memory.

``DevicePrefetcher`` stages batch i+1 to the GPU on a side stream while batch i computes, so the
PCIe copy (the bound of end-to-end screening: ~12.5 KB per 80-atom graph) overlaps the kernels;
``predict_stream`` runs a model over such a stream and hands back host results one batch late, so
the host never waits on the step it has just queued.
"""
from __future__ import annotations

from collections import deque
from typing import Iterable, Iterator

import torch

from . import ops
from .data import Batch
from .packed import DeviceGraphStore, PackedDataset


_COPY_STREAMS: dict = {}


class DevicePrefetcher:
    """Iterates an iterable of host ``Batch`` objects as device batches, ``depth`` copies ahead.
    An item may also be a ``PackedDataset`` chunk (packed wire format, ~11x fewer bytes per graph): its image
    crosses PCIe as ONE copy and is expanded on the device by ``hcg_unpack_batch`` - or, with ``lazy_packed``, consumed
    as it is by ``hcg_gcn_forward_packed``.

    Copies are issued on a private stream (pinned host memory makes them truly asynchronous);
    the consumer's current stream is made to wait on each batch's copy event when it is handed
    out, and the device tensors are marked as used by that stream for the caching allocator.
    """

    def __init__(self, batches: Iterable[Batch], device=None, depth: int = 2, lazy_packed: bool = False):
        self.device = torch.device(device) if device is not None else ops._require_cuda()
        # lazy_packed: a PackedDataset chunk is handed out as a LazyPackedBatch (unpacked only when x / edge_index are
        # read): what inference wants; training reads x and y anyway
        self.lazy_packed = bool(lazy_packed)
        self.batches = iter(batches)
        self.depth = max(1, int(depth))
        # ONE copy stream per device for the whole process: torch's caching allocator keeps a pool per stream, so a new
        # stream per prefetcher would strand the previous one's staging buffers and cudaMalloc its own again
        key = (self.device.type, self.device.index if self.device.index is not None else torch.cuda.current_device())
        if key not in _COPY_STREAMS:
            _COPY_STREAMS[key] = torch.cuda.Stream(self.device)
        self.copy_stream = _COPY_STREAMS[key]
        self.queue: deque = deque()

    def _stage_one(self) -> bool:
        try:
            h = next(self.batches)
        except StopIteration:
            return False
        if isinstance(h, PackedDataset):
            with torch.cuda.stream(self.copy_stream):
                owner = getattr(h, "_owner", None)     # the pinned torch buffer behind an image packed with pinned=True
                blob_h = owner[:h.nbytes] if owner is not None else torch.from_numpy(h.blob)
                blob_d = ops.bucketed_empty((blob_h.numel(),), torch.uint8, self.device)
                blob_d.copy_(blob_h, non_blocking=True)
                store = DeviceGraphStore.from_blob(h, blob_d)
                # unpacked only if somebody reads x / edge_index: inference runs straight from the image
                d = store.lazy_batch() if self.lazy_packed else store.batch(None, with_batch_vector=False)
                d._store = store
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
            self.queue.append((d, ev, h))
            return True
        with torch.cuda.stream(self.copy_stream):
            def mv(t):
                if t is None or t.is_cuda:
                    return t
                t_d = ops.bucketed_empty(t.shape, t.dtype, self.device)     # (size buckets: see ops.bucketed_empty)
                t_d.copy_(t, non_blocking=True)
                return t_d

            ptr = getattr(h, "ptr", None)
            if ptr is not None and not ptr.is_cuda and ptr.numel() == h.num_graphs + 1 and h.x is not None:
                # `batch` and `ptr` carry the same information: ship the (B+1)-entry ptr and rebuild the
                # N-entry batch vector on the device (5 % fewer bytes over PCIe for a reaction graph)
                ptr_d = mv(ptr)
                bt = None          # (Batch.batch rebuilds it on the device if anybody asks)
            else:
                ptr_d, bt = None, mv(h.batch)
            d = Batch(mv(h.x), mv(h.edge_index), bt, ptr_d, mv(h.y), h.idx, h.num_graphs,
                      edges_grouped=getattr(h, "edges_grouped", None))
            d.max_graph_nodes = getattr(h, "max_graph_nodes", 0)
            d.max_graph_edges = getattr(h, "max_graph_edges", 0)
            d.max_in_degree = getattr(h, "max_in_degree", None)
            d.edge_ptr = mv(getattr(h, "edge_ptr", None)) if ptr_d is not None else None
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        self.queue.append((d, ev, h))
        return True

    def __iter__(self) -> Iterator[Batch]:
        while len(self.queue) < self.depth and self._stage_one():
            pass
        while self.queue:
            d, ev, h = self.queue.popleft()
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            store = getattr(d, "_store", None)
            lazy = getattr(d, "_packed_store", None) is not None
            for t in ((None, None, None) if lazy else (d.x, d.edge_index, d.y)) + \
                    (d._batch, d.ptr, getattr(d, "edge_ptr", None), getattr(d, "_keepalive", None),
                     store.blob_dev if store is not None else None):
                if t is not None and t.is_cuda:
                    t.record_stream(cur)
            self._stage_one()  # refill behind the batch being handed out
            d._host = h
            yield d


def predict_stream(model, batches: Iterable[Batch], return_embedding: bool = False, device=None,
                   depth: int = 2, validate: bool = False):
    """Yields ``(out, emb_or_None, host_batch)`` per batch, ``out``/``emb`` as pinned HOST tensors.

    Inference only (``torch.no_grad``).  ``validate=True`` keeps the per-batch status read-back of
    ``hcg_collate`` (one stream sync per batch; bad indices raise); with ``False`` the kernels stay
    memory-safe but malformed input is not reported.
    """
    prev_validate = getattr(model, "validate_inputs", True)
    model.validate_inputs = bool(validate)
    was_training = model.training
    model.eval()
    pending: deque = deque()
    try:
        with torch.no_grad():
            for d in DevicePrefetcher(batches, device=device, depth=depth, lazy_packed=True):
                out, emb = model(d, True)
                o_h = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
                o_h.copy_(out, non_blocking=True)
                e_h = None
                if return_embedding:
                    e_h = torch.empty(emb.shape, dtype=emb.dtype, pin_memory=True)
                    e_h.copy_(emb, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(out.device))   # the stream the copies above were queued on
                pending.append((o_h, e_h, d._host, ev, (out, emb)))
                if len(pending) > 1:  # hand back the previous batch: the GPU queue stays one step ahead
                    o, e, h, pev, _keep = pending.popleft()
                    pev.synchronize()
                    yield o, e, h
            while pending:
                o, e, h, pev, _keep = pending.popleft()
                pev.synchronize()
                yield o, e, h
    finally:
        model.validate_inputs = prev_validate
        model.train(was_training)


def predict_network_stream(model, loader: Iterable[Batch], return_emb: bool = False, **kw):
    """``predict_network`` (reference utils/utils_model.py:82-111) on the GPU: returns numpy arrays
    ``(y_pred, y_true, idx[, embeddings])`` concatenated over the loader."""
    import numpy as np
    y_pred, y_true, idx, embs = [], [], [], []
    for out, emb, h in predict_stream(model, loader, return_embedding=return_emb, **kw):
        y_pred.append(out.numpy().ravel().copy() if out.shape[1] == 1 else out.numpy().copy())
        if h.y is not None:
            y_true.append(np.asarray(h.y.cpu() if torch.is_tensor(h.y) else h.y))
        if h.idx is not None:
            idx.append(np.asarray(h.idx.cpu() if torch.is_tensor(h.idx) else h.idx))
        if return_emb:
            embs.append(emb.numpy().copy())
    cat = lambda xs: np.concatenate(xs, 0) if xs else np.empty((0,))  # noqa: E731
    if return_emb:
        return cat(y_pred), cat(y_true), cat(idx), cat(embs)
    return cat(y_pred), cat(y_true), cat(idx)


def predict_ensemble_stream(models, batches: Iterable[Batch], device=None, depth: int = 2, validate: bool = False,
                            return_embedding: bool = False, one_pass: bool | None = None):
    """Ensemble screening — the intended ``scripts_experiments/predict.py`` (reference :11-60, dead code
    there): every model of ``models`` (e.g. the 90 ``model_params.pth`` of a nested cross-validation,
    reference scripts_experiments/train_GNN.py:48-137) is run over every batch of the stream.

    A batch crosses PCIe ONCE.  For the default architecture on batches that carry their offsets and size facts
    the whole ensemble then runs in ONE pass over the batch (``hcg_gcn_forward_coo_ensemble``: graphs staged, neighbour
    lists built and the layer-0 aggregation done once per tile, the models' weights streamed through shared memory);
    otherwise the batch is collated once and the network kernel runs per model.  ``one_pass``: ``None`` = when
    possible, ``False`` = never, ``True`` = raise if not possible.  Same bits per model either way.
    Yields ``(pred[M, B, C], mean[B, C], std[B, C], host_batch)`` per batch as pinned HOST tensors (``std`` is the
    population standard deviation over the ensemble); with ``return_embedding`` a fifth item ``emb[M, B, 2H]``."""
    models = list(models)
    if not models:
        raise ValueError("predict_ensemble_stream needs at least one model")
    dims0 = models[0]._dims()
    for m in models[1:]:
        if m._dims() != dims0:
            raise ValueError("all models of an ensemble must share one architecture")
    dev = torch.device(device) if device is not None else ops._require_cuda()
    params = [[ops.to_device(p.detach(), dev, torch.float32) for p in m._flat_params()] for m in models]
    stacked = ops.EnsembleWeights(dims0, params, dev) if one_pass is not False and ops.EnsembleWeights.supported(dims0) else None
    if one_pass is True and stacked is None:
        raise ValueError("one_pass=True: this architecture has no one-pass ensemble kernel")
    pending: deque = deque()

    def finish(item):
        hosts, h, ev, _keep = item
        ev.synchronize()
        return (*hosts, h) if not return_embedding else (*hosts[:3], h, hosts[3])

    with torch.no_grad():
        for d in DevicePrefetcher(batches, device=dev, depth=depth):
            x = ops.to_device(d.x, dev, torch.float32)
            res, plan = None, None
            if stacked is not None and (ops.edge_list_hints_ok(d) or validate or one_pass is True) and \
                    getattr(d, "ptr", None) is not None and getattr(d, "edge_ptr", None) is not None and \
                    d.edge_index.dtype == torch.int64:
                res = ops.forward_edge_list_ensemble(x.contiguous(), d.edge_index.contiguous(), d.ptr, d.edge_ptr, stacked,
                                                     validate=validate or not ops.edge_list_hints_ok(d))
                if res is None and one_pass is True:
                    raise ValueError("one_pass=True: the device refused the batch (oversize graph, in-degree above 16, self-loop)")
            if res is not None:
                pred, emb = res
            else:
                plan = ops.cached_collate(d, d.edge_index, d.batch, d.x.shape[0], d.num_graphs, dev, transpose=False,
                                          validate=validate, max_graph_nodes=int(getattr(d, "max_graph_nodes", 0) or 0),
                                          grouped=getattr(d, "edges_grouped", None))
                pred = torch.empty(len(models), d.num_graphs, dims0.C, dtype=torch.float32, device=dev)
                emb = torch.empty(len(models), d.num_graphs, 2 * dims0.H, dtype=torch.float32, device=dev) if return_embedding else None
                for k, ps in enumerate(params):
                    out, e_k, _ = ops.forward_raw(plan, x, dims0, ps, save=False)
                    pred[k].copy_(out)
                    if emb is not None:
                        emb[k].copy_(e_k)
            mean, std = pred.mean(0), pred.std(0, unbiased=False)
            hosts = []
            for t in (pred, mean, std) + ((emb,) if return_embedding else ()):
                t_h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
                t_h.copy_(t, non_blocking=True)
                hosts.append(t_h)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(dev))
            pending.append((hosts, d._host, ev, (pred, mean, std, emb, plan, x)))
            if len(pending) > 1:
                yield finish(pending.popleft())
        while pending:
            yield finish(pending.popleft())
