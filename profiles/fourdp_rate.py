#!/usr/bin/env python
"""north_star: "selectivity predictions on the RhCAA set identical to 4 decimal places".  For every shipped model of
the golden fixtures: how many rows of the reference's own predictions (embeddings.csv, ddG_pred) the CUDA path
reproduces as a 4-decimal string, the largest absolute difference, and the same two figures for the CPU oracle
(a different summation order already flips rows that sit on a rounding boundary).  Runs on the GPU box:

    python profiles/fourdp_rate.py > gpurun_out/r2_fourdp_rate.txt
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import hcatgnet_b200 as pkg  # noqa: E402
from hcatgnet_b200 import _lib  # noqa: E402
from oracle import ref_gcn  # noqa: E402  (checker only)
from tests import golden_io  # noqa: E402


def rate(a, b):
    return float(np.mean([f"{u:.4f}" == f"{v:.4f}" for u, v in zip(a, b)]))


print(f"{'model':22s} {'rows':>5s} {'path':>12s} {'4dp match':>9s} {'max |d|':>10s}   (oracle: {'4dp':>6s} {'max |d|':>10s})")
for rset, outer, inner in golden_io.MODELS:
    sd, graphs, _, gold = golden_io.rows_with_graphs(rset, outer, inner)
    F = graphs[0]["x"].shape[1]
    model = pkg.make_network("GCN", pkg.default_options(), F)
    model.load_state_dict(sd, strict=True)
    model = model.to("cuda").eval()
    ref = []
    for s in range(0, len(graphs), 40):
        b = ref_gcn.collate(graphs[s:s + 40])
        ref.append(ref_gcn.forward(sd, b["x"], b["edge_index"], b["batch"], b["num_graphs"])[0].detach().numpy().ravel())
    ref = np.concatenate(ref)
    for path, name in ((_lib.PATH_AUTO, "auto"), (_lib.PATH_TENSOR_CORE, "tensor-core"), (_lib.PATH_GRAPH_RESIDENT, "fma"),
                       (_lib.PATH_LAYERWISE, "layer-wise")):
        if path == _lib.PATH_TENSOR_CORE and F > 32:
            continue
        prev = _lib.config_set(_lib.HCG_CFG_FORWARD_PATH, path)
        try:
            with torch.no_grad():
                out = torch.cat([model(b.to("cuda")).cpu() for b in pkg.DataLoader(graphs, batch_size=40)]).numpy().ravel()
        finally:
            _lib.config_set(_lib.HCG_CFG_FORWARD_PATH, prev)
        print(f"{rset + f'_o{outer}_i{inner}':22s} {len(gold):5d} {name:>12s} {rate(out, gold):9.4f} {np.abs(out - gold).max():10.2e}"
              f"   (oracle: {rate(ref, gold):6.4f} {np.abs(ref - gold).max():10.2e})")
