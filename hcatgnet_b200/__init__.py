"""hcatgnet_b200 — B200-native (sm_100a) implementation of HCat-GNet's GCN hot path.

Public surface mirrors the reference: ``GCN(opt, n_node_features)`` / ``GCN_explain`` with the
reference's forward signatures and state_dict layout, ``make_network``; plus ``Batch`` /
``DataLoader`` (what the reference takes from torch_geometric) and ``default_options``.
All arithmetic runs in libhcg.so (include/hcg.h); there is no CPU implementation.
"""
from .call_methods import create_loaders, make_network  # noqa: F401
from .data import Batch, DataLoader  # noqa: F401
from .model import GCN, GCN_explain  # noqa: F401
from .options import default_options  # noqa: F401
from .packed import (DeviceGraphStore, PackedDataset, StoreLoader, convert_processed_dir,  # noqa: F401
                     create_store_loaders)
from .screen import DevicePrefetcher, predict_ensemble_stream, predict_network_stream, predict_stream  # noqa: F401
from .trainer import FusedTrainer  # noqa: F401


def __getattr__(name):   # report.py needs pandas: imported on first use only
    if name in ("write_report_csvs", "screen_ensemble_to_csv", "embeddings_frame"):
        from . import report
        return getattr(report, name)
    raise AttributeError(name)

__version__ = "0.1.0"
