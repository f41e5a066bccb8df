"""Generate the committed golden fixtures from the reference's shipped artefacts.

Run in the build container only (needs /root/reference, which the GPU box does not have):

    python tests/golden/make_golden.py

Writes, next to this file:
  graphs_<dataset>.npz   packed input graphs of data/datasets/<dataset>/processed/reaction_*.pt
                         (x, edge_index, per-graph node/edge counts, y, idx, fold)
  model_<set>_o<O>_i<I>.npz  state_dict tensors of results/<set>/results_GNN/Fold_O_test_set/
                         Fold_I_val_set/model_params.pth + the matching embeddings.csv rows
                         (128-d graph embedding, ddG_pred, ddG_exp, index, set) — the golden
                         forward vectors written by reference utils/utils_model.py:199-205.
"""
import glob
import os
import re
import sys

import numpy as np
import pandas as pd
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle.pyg_stub import load_reaction  # noqa: E402

REF = "/root/reference"

DATASETS = ["rhcaa_learning", "rhcaa_final_test", "biAryl"]
# (results set, dataset the rows index into, outer, inner)
MODELS = [
    ("learning", 1, 2), ("learning", 8, 10), ("learning", 10, 9),
    ("biaryl", 1, 2), ("biaryl", 5, 6),
    ("final_test", 1, 2), ("final_test", 8, 10),
]
WEIGHTS_FROM = {"final_test": "learning"}  # reference scripts_experiments/predict_test.py:44,55


def pack_dataset(name):
    files = glob.glob(os.path.join(REF, "data/datasets", name, "processed", "reaction_*.pt"))
    files = [f for f in files if re.fullmatch(r"reaction_\d+\.pt", os.path.basename(f))]
    files.sort(key=lambda f: int(re.findall(r"\d+", os.path.basename(f))[0]))
    xs, eis, nn, ne, ys, idxs, folds = [], [], [], [], [], [], []
    for f in files:
        d = load_reaction(f)
        file_idx = int(re.findall(r"\d+", os.path.basename(f))[0])
        assert int(d["idx"]) == file_idx, (f, d["idx"])
        xs.append(d["x"].numpy()); eis.append(d["edge_index"].numpy().astype(np.int32))
        nn.append(d["x"].shape[0]); ne.append(d["edge_index"].shape[1])
        ys.append(float(d["y"][0])); idxs.append(file_idx); folds.append(int(d.get("fold", -1)))
    out = os.path.join(HERE, f"graphs_{name}.npz")
    np.savez_compressed(
        out, x=np.concatenate(xs, 0).astype(np.float32), edge_index=np.concatenate(eis, 1),
        n_nodes=np.array(nn, np.int32), n_edges=np.array(ne, np.int32),
        y=np.array(ys, np.float32), idx=np.array(idxs, np.int64), fold=np.array(folds, np.int32))
    print(out, len(files), "graphs", os.path.getsize(out) // 1024, "KiB")


def pack_model(rset, outer, inner):
    wset = WEIGHTS_FROM.get(rset, rset)
    wdir = os.path.join(REF, "results", wset, "results_GNN", f"Fold_{outer}_test_set", f"Fold_{inner}_val_set")
    rdir = os.path.join(REF, "results", rset, "results_GNN", f"Fold_{outer}_test_set", f"Fold_{inner}_val_set")
    sd = torch.load(os.path.join(wdir, "model_params.pth"), weights_only=True, map_location="cpu")
    df = pd.read_csv(os.path.join(rdir, "embeddings.csv"))
    emb = df[[str(i) for i in range(128)]].to_numpy(np.float64)
    emb32 = emb.astype(np.float32)
    # the CSV holds the shortest round-trip repr of the original float32 values
    assert np.abs(emb32 - emb).max() <= 1e-7 * max(1.0, np.abs(emb).max())
    arrays = {"sd__" + k: v.numpy() for k, v in sd.items()}
    arrays["sd_keys"] = np.array(list(sd.keys()))
    arrays.update(emb=emb32, ddG_pred=df["ddG_pred"].to_numpy(np.float32),
                  ddG_exp=df["ddG_exp"].to_numpy(np.float32), index=df["index"].to_numpy(np.int64),
                  set=df["set"].to_numpy(str))
    out = os.path.join(HERE, f"model_{rset}_o{outer}_i{inner}.npz")
    np.savez_compressed(out, **arrays)
    print(out, len(df), "rows", os.path.getsize(out) // 1024, "KiB")


if __name__ == "__main__":
    for ds in DATASETS:
        pack_dataset(ds)
    for m in MODELS:
        pack_model(*m)
