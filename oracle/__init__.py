"""CPU oracle for the Hydra-GNN message-passing forward path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``hydra_gnn_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
as the timed CPU baseline -- never as the product path.

What it restates
----------------
The reference (MIT-SPARK/Hydra-GNN) delegates every arithmetic operation of its
forward pass to the third-party package ``torch_geometric`` (PyG), which is
*unpinned* in the reference (``setup.cfg:14``; README.md:53-56 lists 2.3.1 /
2.2.0 / 2.0.4 as tested) and is neither installed nor installable on this image.
``oracle/pyg`` is therefore a plain-``torch`` restatement of the published
algorithm of the PyG 2.3.1 operators the reference calls
(models/utils.py:14,16,18,49,100,132; heterogeneous_neural_tree_network.py:18,95,103;
homogeneous_neural_tree_network.py:63; homogeneous_network.py:9,97), written as
gather -> elementwise -> scatter exactly like PyG's own CPU path.

``oracle/models.py`` restates the control flow of the reference's four model
classes on top of it.  In the build container the *reference's own model files*
are additionally imported from ``/root/reference/src`` with ``oracle.pyg``
registered as ``torch_geometric`` (``oracle/ref_import.py``), which (a) checks
the restated classes against the reference's real control flow and (b) generates
the committed fixtures in ``tests/golden/`` (``tests/golden/make_golden.py``).

Parity status: **parity unpinned** for the floating-point layers.  The reference
ships no test, golden vector or fixture that touches a model
and PyG itself cannot be run here, so the restated layer arithmetic is pinned
only against an independent dense-adjacency formulation (``oracle/dense.py``).
The integer side (H-tree construction) IS pinned against the reference's own
networkx code run in the build container (golden SHA-256, tests/golden).
"""
