"""One GPU test per PyG semantic trap of SURVEY.md section 9 (the items marked with a warning sign), run on the
PRODUCT (CUDA kernels through the C ABI).  The expected values are computed here, in plain Python floats, from edge
lists written out by hand AFTER the PyG rule under test was applied on paper -- nothing in this file imports
``oracle/``.  Weights are chosen so that every intermediate is a short decimal.
"""
import math

import pytest
import torch

from hydra_gnn_b200 import nn as hnn

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _t(v, dtype=torch.float32):
    return torch.tensor(v, dtype=dtype, device=DEV)


def _ei(pairs):
    """[(src, dst), ...] -> int64 [2, E] on the device."""
    if not pairs:
        return torch.zeros(2, 0, dtype=torch.int64, device=DEV)
    return _t(pairs, torch.int64).t().contiguous()


def _leaky(v, slope=0.2):
    return v if v > 0 else slope * v


def _gat_rows(xs, a_src, a_dst, edges, n_dst, bias, a_edge=None):
    """Single-head GAT on paper: ``edges`` = the FINAL (src, dst) list; returns n_dst rows of len(xs[0]) floats.
    softmax with PyG's ``+ 1e-16`` in the denominator; a row with no edge is ``bias``."""
    out = []
    for i in range(n_dst):
        inc = [(k, j) for k, (j, d) in enumerate(edges) if d == i]
        row = [0.0] * len(xs[0])
        if inc:
            sc = [_leaky(a_src[j] + a_dst[i] + (a_edge[k] if a_edge is not None else 0.0)) for k, j in inc]
            m = max(sc)
            ex = [math.exp(s - m) for s in sc]
            den = sum(ex) + 1e-16
            for (k, j), e in zip(inc, ex):
                for c in range(len(row)):
                    row[c] += e / den * xs[j][c]
        out.append([r + b for r, b in zip(row, bias)])
    return out


def _set(p, v):
    with torch.no_grad():
        p.copy_(torch.tensor(v, dtype=torch.float32).reshape(p.shape))


def _close(got, want, what):
    want = torch.tensor(want, dtype=torch.float32)
    torch.testing.assert_close(got.detach().cpu(), want, rtol=1e-5, atol=1e-6, msg=lambda m: f"{what}: {m}")


def _bipartite_gat(add_self_loops, edge_dim=None, fill_value="mean"):
    """GATConv((2, 2) -> 2, one head): lin_src = I, lin_dst = 2 I, att_src = (1, 0), att_dst = (0, 0.5), bias (0.1, -0.2).
    => xs = x_src, a_src[j] = x_src[j][0], a_dst[i] = 0.5 * 2 * x_dst[i][1] = x_dst[i][1]."""
    conv = hnn.GATConv((2, 2), 2, heads=1, concat=True, add_self_loops=add_self_loops, edge_dim=edge_dim,
                       fill_value=fill_value).to(DEV).eval()
    _set(conv.lin_src.weight, [[1, 0], [0, 1]])
    _set(conv.lin_dst.weight, [[2, 0], [0, 2]])
    _set(conv.att_src, [1, 0])
    _set(conv.att_dst, [0, 0.5])
    _set(conv.bias, [0.1, -0.2])
    return conv


XS3 = [[1.0, 2.0], [0.5, -1.0], [-2.0, 0.25]]
XD2 = [[9.0, 0.5], [9.0, -1.5]]


def test_trap_bipartite_gat_self_loops_use_min_of_the_two_node_counts():
    """9.2: ``add_self_loops`` on a bipartite conv with N_src = 3, N_dst = 2 -- PyG takes ``N = min(N_src, N_dst) = 2``,
    drops every input edge whose two INDICES coincide (here 0 -> 0) and appends (0, 0), (1, 1) after the real edges."""
    conv = _bipartite_gat(True)
    with torch.no_grad():
        got = conv((_t(XS3), _t(XD2)), _ei([(0, 0), (2, 1)]))
    final = [(2, 1), (0, 0), (1, 1)]  # 0->0 removed, loops appended last
    want = _gat_rows(XS3, [r[0] for r in XS3], [r[1] for r in XD2], final, 2, [0.1, -0.2])
    _close(got, want, "bipartite loops")
    # row 0 has the single edge 0 -> 0: softmax weight 1 -> xs[0] + bias; had the loop rule used N_src = 3 nothing
    # changes here, had it used N_dst rows of a third loop (2, 2) the conv would have indexed a third target
    assert got.shape == (2, 2)
    _close(got[0], [1.1, 1.8], "row 0 = xs[0] + bias")


def test_trap_input_self_loops_are_removed_then_reappended_last_with_the_fill_attribute():
    """9.2: an input self loop (1 -> 1) carrying edge_attr (5, 5, 5) is REMOVED by ``remove_self_loops`` and the loop
    added back gets ``fill_value`` = zeros(3) as its attribute -- so the (5, 5, 5) must not reach the score."""
    conv = hnn.GATConv(2, 2, heads=1, concat=True, add_self_loops=True, edge_dim=3,
                       fill_value=torch.zeros(3, dtype=torch.float64)).to(DEV).eval()
    _set(conv.lin_src.weight, [[1, 0], [0, 1]])
    _set(conv.att_src, [1, 0])
    _set(conv.att_dst, [0, 1])
    _set(conv.lin_edge.weight, [[1, 0, 0], [0, 0, 0]])  # lin_edge(e) = (e[0], 0)
    _set(conv.att_edge, [0.5, 0])                       # a_edge = 0.5 * e[0]
    _set(conv.bias, [0.0, 0.0])
    x = [[1.0, 0.5], [2.0, -1.0]]
    edges_in = [(0, 1), (1, 1)]
    attr_in = [[2.0, 7.0, 7.0], [5.0, 5.0, 5.0]]
    with torch.no_grad():
        got = conv(_t(x), _ei(edges_in), _t(attr_in))
    final = [(0, 1), (0, 0), (1, 1)]
    a_edge = [0.5 * 2.0, 0.0, 0.0]  # the loops carry the zero fill, NOT the removed loop's (5, 5, 5)
    want = _gat_rows(x, [r[0] for r in x], [r[1] for r in x], final, 2, [0.0, 0.0], a_edge)
    _close(got, want, "removed loop attr")
    wrong = _gat_rows(x, [r[0] for r in x], [r[1] for r in x], final, 2, [0.0, 0.0], [1.0, 0.0, 2.5])
    assert abs(wrong[1][0] - want[1][0]) > 1e-2  # the test discriminates


def test_trap_lin_dst_is_dead_for_same_type_relations_of_a_tuple_in_channels_conv():
    """9.2 / 9.1: the hetero models build every GATConv with a TUPLE in_channels, so ``lin_dst`` exists; for a
    same-type relation HeteroConv passes ONE tensor and PyG computes alpha_dst from ``lin_src(x)``: lin_dst is dead."""
    conv = _bipartite_gat(True)
    _set(conv.lin_dst.weight, [[1e3, -1e3], [1e3, 1e3]])  # poison: must not matter
    het = hnn.HeteroConv({("a", "r", "a"): conv}, aggr="sum").to(DEV).eval()
    x = [[1.0, 2.0], [0.5, -1.0], [-2.0, 0.25]]
    with torch.no_grad():
        got = het({"a": _t(x)}, {("a", "r", "a"): _ei([(0, 1), (2, 1), (1, 2)])})["a"]
    final = [(0, 1), (2, 1), (1, 2), (0, 0), (1, 1), (2, 2)]
    a_dst = [0.5 * r[1] for r in x]  # att_dst . lin_SRC(x) = 0.5 * x[1]
    want = _gat_rows(x, [r[0] for r in x], a_dst, final, 3, [0.1, -0.2])
    _close(got, want, "dead lin_dst")


def test_trap_relation_without_edges_returns_bias_plus_root_term():
    """9.1: a relation with E = 0 still runs its conv: GATConv -> ``bias``; SAGEConv -> ``lin_l.bias + lin_r x_dst``."""
    gat = _bipartite_gat(False)
    with torch.no_grad():
        got = gat((_t(XS3), _t(XD2)), _ei([]))
    _close(got, [[0.1, -0.2], [0.1, -0.2]], "GAT, E = 0")
    sage = hnn.SAGEConv((2, 2), 2).to(DEV).eval()
    _set(sage.lin_l.weight, [[1, 1], [1, -1]])
    _set(sage.lin_l.bias, [0.25, 0.5])
    _set(sage.lin_r.weight, [[0, 1], [2, 0]])
    with torch.no_grad():
        got = sage((_t(XS3), _t(XD2)), _ei([]))
    _close(got, [[0.25 + 0.5, 0.5 + 18.0], [0.25 - 1.5, 0.5 + 18.0]], "SAGE, E = 0")


def test_trap_isolated_target_and_softmax_epsilon():
    """9.2: ``s_i = sum exp + 1e-16``: a target with no in-edge is 0 / (0 + 1e-16)-free -- its row is exactly ``bias``,
    never NaN; a target with one in-edge gets weight 1 / (1 + 1e-16) == 1 in fp32."""
    conv = _bipartite_gat(False)
    with torch.no_grad():
        got = conv((_t(XS3), _t(XD2)), _ei([(2, 1)]))
    assert torch.isfinite(got).all()
    _close(got, [[0.1, -0.2], [-2.0 + 0.1, 0.25 - 0.2]], "isolated target")


def test_trap_gcn_keeps_an_existing_self_loop_once():
    """9.4: ``add_remaining_self_loops``: node 1 already has a loop -- it is kept ONCE (weight 1), node 0 gets one.
    deg = (1, 2): out[0] = xw[0] + b, out[1] = xw[0] / sqrt(1 * 2) + xw[1] / 2 + b."""
    conv = hnn.GCNConv(2, 2).to(DEV).eval()
    _set(conv.lin.weight, [[1, 0], [0, 2]])
    _set(conv.bias, [0.5, -0.5])
    x = [[1.0, 3.0], [-2.0, 0.5]]
    with torch.no_grad():
        got = conv(_t(x), _ei([(0, 1), (1, 1)]))
    xw = [[1.0, 6.0], [-2.0, 1.0]]
    r2 = 1.0 / math.sqrt(2.0)
    want = [[xw[0][0] + 0.5, xw[0][1] - 0.5],
            [xw[0][0] * r2 + xw[1][0] / 2 + 0.5, xw[0][1] * r2 + xw[1][1] / 2 - 0.5]]
    _close(got, want, "GCN existing loop")
    dup = xw[0][0] / math.sqrt(3.0) + 2 * xw[1][0] / 3 + 0.5  # what a double-counted loop would give
    assert abs(dup - want[1][0]) > 1e-2


def test_trap_leafpool_returns_as_many_rows_as_x_and_callers_slice():
    """9.6: ``propagate(edge_index, x=x)`` with ONE tensor infers size = (x.size(0), x.size(0)): the mean pool over
    5 leaves into 2 virtual nodes has 5 rows, rows 2.. are zero (heterogeneous_neural_tree_network.py:181-185 slices
    ``[0:num_virtual]``)."""
    pool = hnn.LeafPool().to(DEV)
    x = [[1.0, 2.0], [3.0, 4.0], [5.0, 6.0], [7.0, 8.0], [-1.0, 0.0]]
    ei = _ei([(0, 0), (1, 0), (2, 1), (3, 1), (4, 1)])
    with torch.no_grad():
        full = pool(_t(x), ei)
        head = pool(_t(x), ei, num_out=2)
    assert full.shape == (5, 2)
    _close(full, [[2.0, 3.0], [11.0 / 3, 14.0 / 3], [0, 0], [0, 0], [0, 0]], "LeafPool rows")
    assert torch.equal(full[:2], head)


def test_trap_heteroconv_skips_relations_without_a_registered_conv():
    """9.1: relations present in ``edge_index_dict`` but not in ``convs`` (the five virtual / init / pool H-tree
    relations in the main layers) are skipped, and the result holds ONLY destination types that received a relation."""
    conv = _bipartite_gat(False)
    het = hnn.HeteroConv({("s", "r", "d"): conv}, aggr="sum").to(DEV).eval()
    x = {"s": _t(XS3), "d": _t(XD2), "v": _t([[1.0, 1.0]])}
    ei = {("s", "r", "d"): _ei([(0, 0), (1, 0), (2, 1)]),
          ("d", "unregistered", "v"): _ei([(0, 0), (1, 0)]),
          ("v", "also_unregistered", "s"): _ei([(0, 2)])}
    with torch.no_grad():
        got = het(x, ei)
    assert list(got.keys()) == ["d"]
    want = _gat_rows(XS3, [r[0] for r in XS3], [r[1] for r in XD2], [(0, 0), (1, 0), (2, 1)], 2, [0.1, -0.2])
    _close(got["d"], want, "registered relation only")


def test_trap_heteroconv_mean_group_divides_by_the_relation_count():
    """9.1 (pre_mp, heterogeneous_neural_tree_network.py:103): ``aggr='mean'`` is ``stack(outs).mean(0)`` over the
    relations that target the type -- biases included in the mean."""
    c1, c2 = _bipartite_gat(False), _bipartite_gat(False)
    _set(c2.bias, [1.0, 1.0])
    het = hnn.HeteroConv({("s", "r1", "d"): c1, ("s", "r2", "d"): c2}, aggr="mean").to(DEV).eval()
    e1, e2 = [(0, 0), (1, 0), (2, 1)], [(2, 0)]
    with torch.no_grad():
        got = het({"s": _t(XS3), "d": _t(XD2)}, {("s", "r1", "d"): _ei(e1), ("s", "r2", "d"): _ei(e2)})["d"]
    a_s, a_d = [r[0] for r in XS3], [r[1] for r in XD2]
    o1 = _gat_rows(XS3, a_s, a_d, e1, 2, [0.1, -0.2])
    o2 = _gat_rows(XS3, a_s, a_d, e2, 2, [1.0, 1.0])
    _close(got, [[(a + b) / 2 for a, b in zip(r1, r2)] for r1, r2 in zip(o1, o2)], "group mean")
