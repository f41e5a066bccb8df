#!/usr/bin/env python
"""End-to-end training-quality parity on the reference's own data and regime: nested
cross-validation runs of rhcaa_learning (668 reactions, batch 40, Adam lr 0.01 eps 1e-9, ReduceLROnPlateau(0.7, 7)
stepped every 5 epochs, early stopping after 6 evaluations without improvement, seed 20232023) through
FusedTrainer.fit over the device-resident packed store — the loop of reference scripts_experiments/train_GNN.py:73-111.
Published single run (results/learning/results_GNN/Fold_1_test_set/Fold_2_val_set/performance.txt): best epoch 110,
test R2 0.930 / MAE 1.986 / RMSE 2.548; over the 10 outer folds: R2 0.914 +- 0.030, MAE 1.937 +- 0.221, RMSE 2.672 +- 0.372.
The shuffle order of the reference's DataLoader (PyG + its RNG stream) cannot be reproduced, so runs are compared
as distributions, not bit for bit.

    python profiles/train_quality.py [runs=9] > gpurun_out/r2_train_quality.json
"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import hcatgnet_b200 as pkg  # noqa: E402
from tests import golden_io  # noqa: E402


def metrics(y, p):
    y, p = np.asarray(y, np.float64), np.asarray(p, np.float64)
    err = p - y
    return {"r2": float(1.0 - (err ** 2).sum() / ((y - y.mean()) ** 2).sum()), "mae": float(np.abs(err).mean()),
            "rmse": float(np.sqrt((err ** 2).mean())), "n": int(y.size)}


def predict(model, loader):
    ys, ps = [], []
    with torch.no_grad():
        for b in loader:
            ps.append(model(b).cpu().numpy().ravel()); ys.append(b.y.cpu().numpy().ravel())
    return np.concatenate(ys), np.concatenate(ps)


def main():
    runs = int(sys.argv[1]) if len(sys.argv) > 1 else 9
    graphs = golden_io.load_graphs("rhcaa_learning")
    ds = pkg.PackedDataset.from_graphs(graphs)
    store = pkg.DeviceGraphStore(ds, "cuda")
    opt = pkg.default_options()
    out = {"dataset": "rhcaa_learning", "graphs": ds.G, "batch_size": opt.batch_size, "runs": []}
    it = pkg.create_loaders(store, opt)
    t_all = time.perf_counter()
    for run in range(runs):
        outer, inner = run // (opt.folds - 1) + 1, run % (opt.folds - 1) + 1
        real_inner = inner + 1 if outer <= inner else inner
        tr, va, te = next(it)
        model = pkg.make_network("GCN", opt, 25).to("cuda")
        trainer = pkg.FusedTrainer(model)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hist = trainer.fit(tr, va, te, epochs=opt.epochs, early_stopping=opt.early_stopping)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        epochs_run = hist.get("epochs_run", None)
        model.load_state_dict(hist["best_state_dict"])
        model.eval()
        rec = {"outer": outer, "inner": real_inner, "best_epoch": hist["best_epoch"], "best_val_loss": hist["best_val"],
               "epochs_run": epochs_run, "wall_s": wall, "train_steps_per_epoch": len(tr),
               "train": metrics(*predict(model, tr)), "val": metrics(*predict(model, va)), "test": metrics(*predict(model, te))}
        out["runs"].append(rec)
        print(f"run {run + 1}/{runs}: outer {outer} inner {real_inner} best epoch {rec['best_epoch']} "
              f"test R2 {rec['test']['r2']:.3f} MAE {rec['test']['mae']:.3f} RMSE {rec['test']['rmse']:.3f} "
              f"({wall:.2f} s, {epochs_run} epochs)", file=sys.stderr)
    out["wall_s_total"] = time.perf_counter() - t_all
    t = [r["test"] for r in out["runs"]]
    out["summary"] = {k: {"mean": float(np.mean([m[k] for m in t])), "sd": float(np.std([m[k] for m in t]))} for k in ("r2", "mae", "rmse")}
    out["summary"]["wall_s_per_run"] = float(np.mean([r["wall_s"] for r in out["runs"]]))
    out["summary"]["ms_per_epoch"] = float(np.mean([r["wall_s"] * 1e3 / max(r["epochs_run"] or 1, 1) for r in out["runs"]]))
    out["published"] = {"single_run_outer1_inner2": {"best_epoch": 110, "r2": 0.930, "mae": 1.986, "rmse": 2.548},
                        "outer_folds": {"r2": [0.914, 0.030], "mae": [1.937, 0.221], "rmse": [2.672, 0.372]}}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
