"""Data-parallel training step: independent reaction graphs sharded over ranks (one process per
GPU), one NCCL all-reduce of the flat gradient per step.

The reference trains on one device with the loss ``sqrt(MSE)`` over the batch
(reference utils/utils_model.py:60-68).  RMSE is not additive over shards, so averaging per-rank
RMSE gradients would optimise a different objective.  Here every rank back-propagates its local
sum of squared errors; the flat gradient travels together with (SSE, count) in ONE all-reduce and
is then rescaled by d sqrt(S/M)/dS = 1/(2 M rmse): the update equals the single-device step on the
concatenated global batch (DP-correctness note).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def _world(group=None) -> int:
    return dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1


def dp_train_step(model, batch, group=None, validate: bool | None = None) -> torch.Tensor:
    """zero_grad -> forward -> RMSE loss -> backward -> (all-reduce) -> optimizer.step.
    Returns the (global) RMSE as a 0-dim tensor on the compute device; no host sync here."""
    if validate is not None:
        model.validate_inputs = bool(validate)
    params = [p for p in model.parameters() if p.requires_grad]
    model.optimizer.zero_grad(set_to_none=True)
    out = model(batch)
    y = torch.unsqueeze(batch.y, dim=1).to(out.device)
    world = _world(group)
    if world == 1:
        loss = torch.sqrt(model.loss(out, y))
        loss.backward()
        model.optimizer.step()
        return loss.detach()
    sse = torch.sum((out - y) ** 2)
    sse.backward()
    count = torch.tensor([float(out.numel())], device=out.device)
    flat = torch.cat([p.grad.reshape(-1) for p in params] + [sse.detach().reshape(1), count])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    total_sse, total_cnt = flat[-2], flat[-1]
    rmse = torch.sqrt(total_sse / total_cnt)
    scale = 1.0 / (2.0 * total_cnt * rmse)
    off = 0
    for p in params:
        n = p.numel()
        p.grad = (flat[off:off + n] * scale).view_as(p)
        off += n
    model.optimizer.step()
    return rmse.detach()


def shard_range(num_graphs: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block of graph indices owned by ``rank`` (screening inference needs no collective)."""
    per = (num_graphs + world - 1) // world
    lo = min(rank * per, num_graphs)
    return lo, min(lo + per, num_graphs)
