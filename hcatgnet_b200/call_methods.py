"""Factory with the reference's name and signature (reference call_methods.py:7-12)."""
import argparse


def make_network(network_name: str, opt: argparse.Namespace, n_node_features: int):
    if network_name == "GCN":
        from .model.gcn import GCN
        return GCN(opt=opt, n_node_features=n_node_features)
    raise ValueError(f"Network {network_name} not implemented")
