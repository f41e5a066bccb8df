from .gcn import GCN, GCN_explain, GCNConv  # noqa: F401
from .networks import BaseNetwork  # noqa: F401
