"""``oracle.pyg`` -- plain-torch restatement of the PyG 2.3.1 surface the reference uses.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

``install_as_torch_geometric()`` registers this package in ``sys.modules`` under the
names the reference imports (``torch_geometric``, ``.nn``, ``.nn.norm.batch_norm``,
``.data``, ``.utils``, ``.loader``) so that the reference's own, unmodified model and
H-tree files can be imported from ``/root/reference/src`` in the build container.
"""
from __future__ import annotations

import sys
import types

from . import data, nn, utils  # noqa: F401


class DataLoader:
    """Minimal ``torch_geometric.loader.DataLoader`` (no shuffle): batches of
    ``Batch.from_data_list`` (base_training_job.py:164-178)."""

    def __init__(self, dataset, batch_size=1, shuffle=False, **kw):
        self.dataset, self.batch_size = dataset, batch_size

    def __iter__(self):
        for i in range(0, len(self.dataset), self.batch_size):
            yield data.Batch.from_data_list([self.dataset[k] for k in range(i, min(i + self.batch_size, len(self.dataset)))])

    def __len__(self):
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size


def install_as_torch_geometric(force: bool = False) -> str:
    """Returns 'pyg-<version>' if the real package is importable (then nothing is
    stubbed), else 'restatement'."""
    if not force:
        try:
            import torch_geometric  # type: ignore

            if not getattr(torch_geometric, "__oracle_restatement__", False):
                return "pyg-" + torch_geometric.__version__
            return "restatement"
        except ImportError:
            pass
    root = types.ModuleType("torch_geometric")
    root.__oracle_restatement__ = True
    root.__version__ = "2.3.1-restated"
    root.__path__ = []  # mark as package
    norm = types.ModuleType("torch_geometric.nn.norm")
    norm.__path__ = []
    bn = types.ModuleType("torch_geometric.nn.norm.batch_norm")
    bn.BatchNorm = nn.BatchNorm
    norm.batch_norm = bn
    norm.BatchNorm = nn.BatchNorm
    nn.norm = norm
    loader = types.ModuleType("torch_geometric.loader")
    loader.DataLoader = DataLoader
    root.nn, root.data, root.utils, root.loader = nn, data, utils, loader
    sys.modules.update({
        "torch_geometric": root,
        "torch_geometric.nn": nn,
        "torch_geometric.nn.norm": norm,
        "torch_geometric.nn.norm.batch_norm": bn,
        "torch_geometric.data": data,
        "torch_geometric.utils": utils,
        "torch_geometric.loader": loader,
    })
    return "restatement"
