"""hydra_gnn_b200 -- B200-native (sm_100a) message-passing forward of Hydra-GNN's models.

Public surface mirrors ``hydra_gnn.models`` of the reference (models/__init__.py:1-4).
"""
from . import schema  # noqa: F401
from .models import (  # noqa: F401
    HeterogeneousNetwork,
    HeterogeneousNeuralTreeNetwork,
    HomogeneousNetwork,
    HomogeneousNeuralTreeNetwork,
    LeafPool,
)

__version__ = "0.1.0"
