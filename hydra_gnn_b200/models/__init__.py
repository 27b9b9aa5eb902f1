"""Drop-in model classes (same names / ctor kwargs / forward(data) as ``hydra_gnn.models``)."""
from ..nn import LeafPool  # noqa: F401
from .heterogeneous import HeterogeneousNetwork, HeterogeneousNeuralTreeNetwork  # noqa: F401
from .homogeneous import HomogeneousNetwork, HomogeneousNeuralTreeNetwork  # noqa: F401
