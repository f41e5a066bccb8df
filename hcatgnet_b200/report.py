"""CSV artefacts of the reference's ``network_report`` (reference utils/utils_model.py:115-330) and the screening
driver it never finished (reference scripts_experiments/predict.py:11-60, predict_test.py:47-70), computed by the
CUDA path: ``embeddings.csv`` (the 2H-d graph embedding + ddG_exp, ddG_pred, index, set per reaction — the file the
parity fixtures of this repository were read from) and ``predictions_test_set.csv`` (real_ddG, predicted_ddG, index).
Same column names, order and pandas formatting as the reference; its t-SNE columns and plots are reporting, not the
hot path, and are left out.
"""
from __future__ import annotations

import os
from typing import Iterable, Sequence

import numpy as np
import pandas as pd
import torch

from .screen import predict_ensemble_stream, predict_network_stream


def embeddings_frame(model, loader: Iterable, set_name: str) -> pd.DataFrame:
    """``predict_network(model, loader, True)[3]`` of the reference (utils/utils_model.py:82-111) + the ``set``
    column ``network_report`` adds (``'training'`` / ``'val'`` / ``'test'``)."""
    y_pred, y_true, idx, emb = predict_network_stream(model, loader, return_emb=True)
    df = pd.DataFrame(emb)
    df["ddG_exp"], df["ddG_pred"], df["index"] = y_true, y_pred, idx
    df["set"] = set_name
    return df


def write_report_csvs(log_dir: str, model, loaders: Sequence) -> dict:
    """``embeddings.csv`` and ``predictions_test_set.csv`` of one (outer, inner) run, as ``network_report`` writes
    them (reference utils/utils_model.py:199-205).  ``loaders`` = (train, val, test).  Returns the frames."""
    os.makedirs(log_dir, exist_ok=True)
    train, val, test = loaders
    emb_all = pd.concat([embeddings_frame(model, train, "training"), embeddings_frame(model, val, "val"),
                         embeddings_frame(model, test, "test")], axis=0)
    emb_all.to_csv(os.path.join(log_dir, "embeddings.csv"))
    t = emb_all[emb_all["set"] == "test"]
    pred = pd.DataFrame({"real_ddG": t["ddG_exp"].to_numpy(), "predicted_ddG": t["ddG_pred"].to_numpy(), "index": t["index"].to_numpy()})
    pred.to_csv(os.path.join(log_dir, "predictions_test_set.csv"))
    return {"embeddings": emb_all, "predictions": pred}


def screen_ensemble_to_csv(out_dir: str, models: Sequence, batches: Iterable, names: Sequence[str] | None = None) -> pd.DataFrame:
    """The intended ``predict.py``: every model of a nested cross-validation (``models``: modules of one architecture,
    e.g. loaded from the 90 ``model_params.pth``) over a screening set streamed from host memory.  Writes
    ``predictions_<name>.csv`` per model (``predictions_test_set.csv`` columns) and ``ensemble.csv``
    (index, real_ddG, mean, std over the models); returns the ensemble frame."""
    os.makedirs(out_dir, exist_ok=True)
    names = list(names) if names is not None else [f"model_{k}" for k in range(len(models))]
    preds, ys, idxs = [], [], []
    for p, _mean, _std, h in predict_ensemble_stream(models, batches):
        preds.append(p.numpy()[:, :, 0].copy() if p.shape[2] == 1 else p.numpy().copy())
        ys.append(np.asarray(h.y.cpu() if torch.is_tensor(h.y) else h.y) if h.y is not None else np.full(p.shape[1], np.nan))
        idxs.append(np.asarray(h.idx.cpu() if torch.is_tensor(h.idx) else h.idx) if h.idx is not None else np.arange(p.shape[1]))
    pred = np.concatenate(preds, axis=1)
    y, idx = np.concatenate(ys), np.concatenate(idxs)
    for k, name in enumerate(names):
        pd.DataFrame({"real_ddG": y, "predicted_ddG": pred[k], "index": idx}).to_csv(os.path.join(out_dir, f"predictions_{name}.csv"))
    ens = pd.DataFrame({"index": idx, "real_ddG": y, "mean_predicted_ddG": pred.mean(0), "std_predicted_ddG": pred.std(0)})
    ens.to_csv(os.path.join(out_dir, "ensemble.csv"))
    return ens
