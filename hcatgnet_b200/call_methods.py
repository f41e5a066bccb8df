"""Factory with the reference's name and signature (reference call_methods.py:7-12)."""
import argparse


def make_network(network_name: str, opt: argparse.Namespace, n_node_features: int):
    if network_name == "GCN":
        from .model.gcn import GCN
        return GCN(opt=opt, n_node_features=n_node_features)
    raise ValueError(f"Network {network_name} not implemented")


def create_loaders(dataset, opt: argparse.Namespace):
    """reference call_methods.py:15-47 (nested cross-validation loaders).  ``dataset``: a
    ``DeviceGraphStore`` (device-resident packed image -> device batches), a ``PackedDataset`` (uploaded
    first), or a list of per-graph records carrying ``fold`` (host collate, as in the reference)."""
    from .data import DataLoader
    from .packed import DeviceGraphStore, PackedDataset, create_store_loaders
    if isinstance(dataset, PackedDataset):
        dataset = DeviceGraphStore(dataset)
    if isinstance(dataset, DeviceGraphStore):
        yield from create_store_loaders(dataset, opt)
        return
    from .data import _field
    folds = [[] for _ in range(opt.folds)]
    for data in dataset:
        folds[int(_field(data, "fold"))].append(data)
    for outer in range(len(folds)):
        proxy = folds[:outer] + folds[outer + 1:]
        test_loader = DataLoader(folds[outer], batch_size=opt.batch_size, shuffle=False)
        for inner in range(len(proxy)):
            rest = proxy[:inner] + proxy[inner + 1:]
            yield (DataLoader([g for sub in rest for g in sub], batch_size=opt.batch_size, shuffle=True),
                   DataLoader(proxy[inner], batch_size=opt.batch_size, shuffle=False), test_loader)
