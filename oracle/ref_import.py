"""Import the reference's OWN model / H-tree modules in the build container.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

``/root/reference`` exists only in the build container (never on the GPU box), so
this module is used (a) by ``tests/golden/make_golden.py`` to generate the committed
fixtures and (b) by CPU tests that self-skip when the path is absent.  The reference
files are imported unmodified from where they lie; ``torch_geometric`` resolves to
``oracle.pyg`` (the restatement) unless the real package is importable.
"""
from __future__ import annotations

import os
import sys

REFERENCE_SRC = "/root/reference/src"


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "hydra_gnn"))


def load():
    """Returns (hydra_gnn.models module, construct module, mp3d_dataset module, oracle_kind)."""
    if not available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    from .pyg import install_as_torch_geometric

    kind = install_as_torch_geometric()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import hydra_gnn.models as ref_models  # noqa: E402
    import hydra_gnn.mp3d_dataset as ref_mp3d  # noqa: E402
    import hydra_gnn.neural_tree.construct as ref_construct  # noqa: E402

    return ref_models, ref_construct, ref_mp3d, kind
