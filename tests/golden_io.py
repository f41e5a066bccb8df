"""Readers for the committed golden fixtures (tests/golden/*.npz, made by make_golden.py)."""
import functools
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# results set -> (dataset for rows with set != 'test', dataset for rows with set == 'test')
ROW_DATASET = {
    "learning": ("rhcaa_learning", "rhcaa_learning"),
    "biaryl": ("biAryl", "biAryl"),
    "final_test": ("rhcaa_learning", "rhcaa_final_test"),
}


@functools.lru_cache(maxsize=None)
def load_graphs(dataset):
    """List of per-graph dicts {x, edge_index(int64), y, idx} in reaction-index order."""
    z = np.load(os.path.join(GOLDEN_DIR, f"graphs_{dataset}.npz"))
    x, ei = z["x"], z["edge_index"].astype(np.int64)
    no = np.concatenate([[0], np.cumsum(z["n_nodes"])])
    eo = np.concatenate([[0], np.cumsum(z["n_edges"])])
    out = []
    for g in range(len(z["n_nodes"])):
        out.append({
            "x": torch.from_numpy(x[no[g]:no[g + 1]].copy()),
            "edge_index": torch.from_numpy(ei[:, eo[g]:eo[g + 1]].copy()),
            "y": torch.tensor([z["y"][g]]), "idx": int(z["idx"][g]), "fold": int(z["fold"][g]),
        })
    return out


def graphs_by_idx(dataset):
    return {g["idx"]: g for g in load_graphs(dataset)}


@functools.lru_cache(maxsize=None)
def load_model(rset, outer, inner):
    z = np.load(os.path.join(GOLDEN_DIR, f"model_{rset}_o{outer}_i{inner}.npz"))
    sd = {str(k): torch.from_numpy(z["sd__" + str(k)].copy()) for k in z["sd_keys"]}
    rows = {k: z[k] for k in ("emb", "ddG_pred", "ddG_exp", "index", "set")}
    return sd, rows


MODELS = [("learning", 1, 2), ("learning", 8, 10), ("learning", 10, 9), ("biaryl", 1, 2),
          ("biaryl", 5, 6), ("final_test", 1, 2), ("final_test", 8, 10)]


def rows_with_graphs(rset, outer, inner):
    """Yield (graph dict, golden emb[128], golden ddG_pred) for every row of the embeddings.csv."""
    sd, rows = load_model(rset, outer, inner)
    ds_a, ds_b = ROW_DATASET[rset]
    ga, gb = graphs_by_idx(ds_a), graphs_by_idx(ds_b)
    graphs, emb, pred = [], [], []
    for r in range(len(rows["index"])):
        src = gb if rows["set"][r] == "test" else ga
        graphs.append(src[int(rows["index"][r])])
        emb.append(rows["emb"][r]); pred.append(rows["ddG_pred"][r])
    return sd, graphs, np.stack(emb), np.array(pred)
