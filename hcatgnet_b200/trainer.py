"""Device-resident training step — the body of the reference's ``train_network``
(reference utils/utils_model.py:55-70: zero_grad, forward, ``sqrt(MSELoss)``, backward,
``optimizer.step()``) as six asynchronous C-ABI calls and no autograd graph:

    hcg_collate -> hcg_gcn_forward -> hcg_rmse_seed -> hcg_gcn_backward -> [all-reduce] -> hcg_adam_step

The parameters of the module are re-pointed at views of ONE flat fp32 slab (``state_dict`` keys, shapes
and order are unchanged), the gradients land in one flat slab whose two tail slots carry
``(sum of squared errors, count)``; with several ranks that slab is the single NCCL all-reduce of the
step, and the ``1 / (2 count rmse)`` factor of ``d sqrt(MSE)`` is applied inside the Adam kernel, so
every rank takes exactly the single-device step on the concatenated batch.
Nothing here synchronises with the host: the returned loss is a 0-dim device tensor.

Learning rate, betas and eps are read from ``model.optimizer.param_groups[0]`` at every step, so the
reference's schedulers (``model.scheduler.step(val_loss)``, reference scripts_experiments/train_GNN.py:88)
keep working on the module's own optimizer object; its moment buffers live here.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import _lib, ops


class FusedTrainer:
    def __init__(self, model, group=None, validate: bool = False):
        if not isinstance(model.optimizer, torch.optim.Adam):
            raise NotImplementedError("FusedTrainer implements the reference's default optimizer (Adam) only; "
                                      "use parallel.dp_train_step for SGD / RMSprop")
        pg = model.optimizer.param_groups[0]
        if pg.get("weight_decay", 0) != 0 or pg.get("amsgrad", False) or pg.get("maximize", False):
            raise NotImplementedError("FusedTrainer: Adam with weight_decay / amsgrad / maximize is not implemented")
        if getattr(model.loss, "__class__", None) is not torch.nn.MSELoss:
            raise NotImplementedError("FusedTrainer implements the regression loss sqrt(MSELoss) only")
        self.model, self.group, self.validate = model, group, bool(validate)
        self.params = model._flat_params()
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FusedTrainer: move the model to a CUDA device first (there is no CPU path)")
        self.device, self.dims = dev, model._dims()
        sizes = [p.numel() for p in self.params]
        self.n = sum(sizes)
        f32 = dict(dtype=torch.float32, device=dev)
        self.flat = torch.empty(self.n, **f32)
        self._offsets = [sum(sizes[:i]) for i in range(len(sizes))]
        off = 0
        with torch.no_grad():
            for p, k in zip(self.params, sizes):
                self.flat[off:off + k].copy_(p.detach().reshape(-1))
                p.data = self.flat[off:off + k].view(p.shape)   # same Parameter objects, storage = the slab
                off += k
        self.grad = torch.zeros(self.n + 2, **f32)              # [gradients | sse | count]
        self.exp_avg = torch.zeros(self.n, **f32)
        self.exp_avg_sq = torch.zeros(self.n, **f32)
        self.losses = torch.zeros(16, **f32)                    # ring of loss slots: a returned loss stays valid for 15 more steps
        self.steps = 0
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.timing = None      # bench.py: a list here collects (start, end) CUDA events around every all-reduce
        self._pstruct = self._gstruct = self._ws = None   # hcg_train_step: cached parameter / gradient tables, workspace

    def _bind(self):
        """(Re-)point the module's parameters at views of the flat slab.  ``model.to(...)`` / ``load_state_dict``
        with assignment re-allocate them (the reference's predict_network does ``model.to('cpu')``, reference
        utils/utils_model.py:83): the slab would then be updated while forward reads detached copies."""
        off = 0
        with torch.no_grad():
            for p in self.params:
                k = p.numel()
                view = self.flat[off:off + k].view(p.shape)
                if p.data_ptr() != view.data_ptr():
                    if p.device != self.device:
                        raise RuntimeError("FusedTrainer: the model was moved off the trainer's device; move it back "
                                           "(model.to(device)) before the next step")
                    view.copy_(p.detach())          # keep whatever the caller loaded into the parameter
                    p.data = view
                off += k

    def state_dict(self) -> dict:
        """Optimizer state the module's own ``optimizer.state_dict()`` does not see (the moments live here)."""
        return {"step": int(self.steps), "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone()}

    def load_state_dict(self, sd: dict) -> None:
        self.steps = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"].to(self.device))
        self.exp_avg_sq.copy_(sd["exp_avg_sq"].to(self.device))

    def accumulate(self, batch) -> None:
        """Forward + loss seed + backward of ``batch`` into this rank's flat slab ``self.grad`` =
        ``[parameter gradients of the shard's sum of squared errors | SSE | count]`` (shard-additive, see module doc)."""
        dev, dims = self.device, self.dims
        lib = _lib.load()
        if any(p.data_ptr() != self.flat.data_ptr() + 4 * o for p, o in zip(self.params, self._offsets)):
            self._bind()
        x = ops.to_device(batch.x, dev, torch.float32)
        y = ops.to_device(batch.y, dev, torch.float32)
        B = int(batch.num_graphs)
        plan = ops.cached_collate(batch, batch.edge_index, ops.batch_vector_if_built(batch), x.shape[0], B, dev, transpose=True,
                                  validate=self.validate, max_graph_nodes=int(getattr(batch, "max_graph_nodes", 0) or 0),
                                  grouped=getattr(batch, "edges_grouped", None))
        params = [p.data for p in self.params]
        out, emb, saved = ops.forward_raw(plan, x, dims, params, save=True)
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream(dev).cuda_stream
            g_out = torch.empty_like(out)
            tail = self.grad[self.n:]
            _lib.check(lib.hcg_rmse_seed(out.data_ptr(), y.data_ptr(), B, dims.C, g_out.data_ptr(), tail.data_ptr(), st))
            ops.backward_raw(plan, x, dims, params, saved, emb, g_out, None, need_gx=False, flat_out=self.grad)

    def reduce(self) -> None:
        """Sum the slabs of all ranks (one NCCL all-reduce of 4 (n + 2) bytes); nothing to do with one rank."""
        if self.world > 1:
            if self.timing is not None:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            dist.all_reduce(self.grad, op=dist.ReduceOp.SUM, group=self.group)
            if self.timing is not None:
                e1.record()
                self.timing.append((e0, e1))

    def apply(self) -> torch.Tensor:
        """``hcg_adam_step`` on the (reduced) slab: rescale by 1 / (2 count rmse), Adam update; returns the RMSE."""
        lib = _lib.load()
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream(self.device).cuda_stream
            pg = self.model.optimizer.param_groups[0]
            b1, b2 = pg["betas"]
            self.steps += 1
            loss = self.losses[self.steps % 16]
            tail = self.grad[self.n:]
            _lib.check(lib.hcg_adam_step(self.flat.data_ptr(), self.grad.data_ptr(), self.exp_avg.data_ptr(),
                                         self.exp_avg_sq.data_ptr(), self.n, tail.data_ptr(), float(pg["lr"]), float(b1),
                                         float(b2), float(pg["eps"]), self.steps, loss.data_ptr(), st))
        return loss

    def _step_one_call(self, batch, x, y, ptr, edge_ptr, adam: bool = True):
        """``hcg_train_step``: the whole step in one library call over a persistent workspace; ``adam=False`` stops
        after the backward (data-parallel ranks: ``reduce()`` and ``apply()`` follow)."""
        lib = _lib.load()
        dev = self.device
        ei = batch.edge_index
        N, E, B = int(x.shape[0]), int(ei.shape[1]), int(batch.num_graphs)
        if self._pstruct is None:
            params = [p.data for p in self.params]
            self._pstruct = ops._params_struct(self.dims, params)
            gc, off = _lib.HcgGrads(), 0
            ptrs = []
            for p_ in params:
                ptrs.append(self.grad.data_ptr() + 4 * off)
                off += p_.numel()
            L, R = self.dims.L, self.dims.R
            for l in range(L):
                gc.conv_w[l], gc.conv_b[l] = ptrs[2 * l], ptrs[2 * l + 1]
            for k in range(R):
                gc.ro_w[k], gc.ro_b[k] = ptrs[2 * L + 2 * k], ptrs[2 * L + 2 * k + 1]
            self._gstruct = gc
        import ctypes as C
        need = int(lib.hcg_train_step_workspace_bytes(N, E, B, C.byref(self._pstruct)))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(int(need * 1.25) + 256, dtype=torch.uint8, device=dev)
        pg = self.model.optimizer.param_groups[0]
        b1, b2 = pg["betas"]
        if adam:
            self.steps += 1
        loss = self.losses[self.steps % 16]
        with torch.cuda.device(dev):
            _lib.check(lib.hcg_train_step(
                x.data_ptr(), ei.data_ptr(), E, ptr.data_ptr(), edge_ptr.data_ptr(), y.data_ptr(), N, B,
                int(getattr(batch, "max_graph_nodes", 0) or 0), int(getattr(batch, "max_graph_edges", 0) or 0),
                -1 if getattr(batch, "max_in_degree", None) is None else int(batch.max_in_degree), 1.0,
                C.byref(self._pstruct), C.byref(self._gstruct), self.flat.data_ptr(), self.grad.data_ptr(),
                self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.n, float(pg["lr"]), float(b1), float(b2),
                float(pg["eps"]), self.steps if adam else 0, loss.data_ptr(), self._ws.data_ptr(), self._ws.numel(),
                torch.cuda.current_stream(dev).cuda_stream))
        return loss if adam else None

    def step(self, batch) -> torch.Tensor:
        """One optimisation step on ``batch`` (a device ``Batch``); returns the global RMSE (device scalar)."""
        if not self.validate and getattr(batch, "edges_grouped", None) is True:
            ptr, eptr, x, y, ei = batch.ptr, getattr(batch, "edge_ptr", None), batch.x, batch.y, batch.edge_index
            dev = self.device
            if all(t is not None and t.device == dev and t.is_contiguous() for t in (ptr, eptr, x, y, ei)) and \
                    x.dtype == torch.float32 and y.dtype == torch.float32 and ei.dtype == torch.int64 and \
                    ptr.dtype == torch.int64 and eptr.dtype == torch.int64 and ei.shape[1] > 0 and x.shape[0] > 0:
                if any(p.data_ptr() != self.flat.data_ptr() + 4 * o for p, o in zip(self.params, self._offsets)):
                    self._bind()
                if self.world == 1:
                    return self._step_one_call(batch, x, y, ptr, eptr)
                self._step_one_call(batch, x, y, ptr, eptr, adam=False)     # gradients of this rank's shard
                self.reduce()
                return self.apply()
        self.accumulate(batch)
        self.reduce()
        return self.apply()

    def checksum(self) -> torch.Tensor:
        """Sum of the parameter slab (replicas of a data-parallel job must agree bit for bit)."""
        return self.flat.double().sum()

    # ---- the reference's epoch functions, with the per-batch .item() syncs replaced by device sums ----------
    def train_network(self, train_loader, device=None) -> float:
        """reference utils/utils_model.py:55-70: one epoch; returns sum(loss_b * num_graphs_b) / len(dataset).
        The per-batch RMSE values are accumulated on the device; the host reads one number per epoch."""
        dev = self.device if device is None else torch.device(device)
        self.model.train()
        acc = torch.zeros((), dtype=torch.float32, device=self.device)
        n = 0
        for batch in train_loader:
            batch = batch.to(dev)
            acc.add_(self.step(batch), alpha=float(batch.num_graphs))
            n += int(batch.num_graphs)
        total = len(train_loader.dataset) if hasattr(train_loader, "dataset") else n
        return float(acc.item()) / max(total, 1)

    @torch.no_grad()
    def eval_network(self, loader, device=None) -> float:
        """reference utils/utils_model.py:72-79 (which runs without no_grad and syncs per batch): same number,
        inference kernels only, one host read."""
        dev = self.device if device is None else torch.device(device)
        self.model.eval()
        acc = torch.zeros((), dtype=torch.float32, device=self.device)
        n = 0
        prev_validate = getattr(self.model, "validate_inputs", True)
        self.model.validate_inputs = self.validate   # no per-batch status read-back unless asked for
        try:
            for batch in loader:
                batch = batch.to(dev)
                out = self.model(batch)
                acc += torch.sqrt(self.model.loss(out, torch.unsqueeze(batch.y, dim=1))) * float(batch.num_graphs)
                n += int(batch.num_graphs)
        finally:
            self.model.validate_inputs = prev_validate
        total = len(loader.dataset) if hasattr(loader, "dataset") else n
        return float(acc.item()) / max(total, 1)

    def fit(self, train_loader, val_loader, test_loader=None, epochs: int = 250, early_stopping: int = 6, log=None):
        """The epoch loop of reference scripts_experiments/train_GNN.py:73-111: train every epoch, every 5th
        epoch step the scheduler on the validation loss, keep the best parameters, stop after
        ``early_stopping`` evaluations without improvement.  Returns a dict with the loss curves, the best
        epoch and the best ``state_dict``; the module is left at its last epoch, as in the reference."""
        from copy import deepcopy
        m = self.model
        hist = {"train": [], "val": [], "test": [], "best_epoch": 0, "best_val": float("inf"), "best_state_dict": None}
        strikes = 0
        hist["epochs_run"] = 0
        for epoch in range(epochs):
            if strikes > early_stopping:
                break
            hist["epochs_run"] = epoch + 1
            train_loss = self.train_network(train_loader)
            val_loss = self.eval_network(val_loader)
            test_loss = self.eval_network(test_loader) if test_loader is not None else float("nan")
            if log is not None:
                log(epoch, train_loss, val_loss, test_loss)
            if epoch % 5 == 0:
                m.scheduler.step(val_loss)
                hist["train"].append(train_loss); hist["val"].append(val_loss); hist["test"].append(test_loss)
                if val_loss < hist["best_val"]:
                    hist["best_val"], hist["best_epoch"] = val_loss, epoch
                    strikes = 0
                    hist["best_state_dict"] = deepcopy(m.state_dict())
                else:
                    strikes += 1
        return hist
