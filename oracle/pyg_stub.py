"""TEST INFRASTRUCTURE — fixture loader for the reference's pickled PyG objects, without PyG.

The reference stores one pickled ``torch_geometric.data.Data`` per reaction
(reference: data/datasets.py:74-78, written at data/rhcaa.py:90-92).  torch_geometric
is not installable here, so the four classes the pickles name are replaced by permissive
stubs and the payload is read from ``obj._store._mapping``.

Only ``tests/golden/make_golden.py`` (run in the build container, where /root/reference
exists) uses this module; nothing on the product path imports ``oracle/``.
"""
from __future__ import annotations

import sys
import types

import torch


class _Permissive:
    def __setstate__(self, state):
        if isinstance(state, dict):
            self.__dict__.update(state)
        else:  # (dict, slots) tuples
            for part in state:
                if isinstance(part, dict):
                    self.__dict__.update(part)


_STUBS = {
    "torch_geometric": [],
    "torch_geometric.data": [],
    "torch_geometric.data.data": ["Data", "DataEdgeAttr", "DataTensorAttr"],
    "torch_geometric.data.storage": ["GlobalStorage", "BaseStorage", "NodeStorage", "EdgeStorage"],
    "torch_geometric.loader": [],
    "torch_geometric.loader.dataloader": ["DataLoader", "Collater"],
}


def install_stubs() -> None:
    """Register empty stand-in modules so that ``torch.load`` can resolve the pickled globals."""
    for mod_name, classes in _STUBS.items():
        if mod_name in sys.modules and not getattr(sys.modules[mod_name], "_hcg_stub", False):
            continue  # a real torch_geometric is importable: leave it alone
        mod = sys.modules.get(mod_name) or types.ModuleType(mod_name)
        mod._hcg_stub = True
        for cls in classes:
            if not hasattr(mod, cls):
                setattr(mod, cls, type(cls, (_Permissive,), {"__module__": mod_name}))
        sys.modules[mod_name] = mod


def load_reaction(path: str) -> dict:
    """Return the field dict (x, edge_index, edge_attr, y, idx, fold, ...) of one reaction_*.pt."""
    install_stubs()
    obj = torch.load(path, weights_only=False, map_location="cpu")
    store = obj.__dict__["_store"]
    return dict(store.__dict__["_mapping"])
