"""Torch modules over the B200 kernels with the reference's `dgl.nn.pytorch` names and
semantics (python/dgl/nn/pytorch/{softmax.py, conv/graphconv.py, conv/sageconv.py,
conv/gatconv.py, conv/fusedGatConv.py, conv/relgraphconv.py, hetero.py}) -- the callers BASELINE.json's
configs run through. Parameters have the same names / shapes, so a reference `state_dict` loads."""
from .softmax import edge_softmax  # noqa: F401
from .conv import GraphConv, SAGEConv, GATConv, FusedGATConv, RelGraphConv  # noqa: F401
from .hetero import HeteroGraphConv  # noqa: F401

__all__ = ["edge_softmax", "GraphConv", "SAGEConv", "GATConv", "FusedGATConv", "RelGraphConv",
           "HeteroGraphConv"]
