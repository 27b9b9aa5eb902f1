"""GPU parity of the product against the committed outputs of the reference's OWN model files
(tests/golden/logits.npz, made by make_golden.py) and against the oracle, for every model
class / conv block / task of the reference -- calls go through the C ABI."""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import cases  # noqa: E402
import hydra_gnn_b200 as hg  # noqa: E402
from hydra_gnn_b200 import convert, synth  # noqa: E402
from hydra_gnn_b200.data import Batch  # noqa: E402
from oracle import models as om  # noqa: E402
from util import assert_close, clone_data  # noqa: E402

pytestmark = pytest.mark.gpu
GOLD = cases.load_golden_logits()


def _outs(o):
    return list(o) if isinstance(o, (tuple, list)) else [o]


@pytest.mark.parametrize("case", list(cases.CASES))
def test_product_matches_reference_outputs(case):
    """tolerance: north_star's 1e-4 relative in fp32 + identical argmax (util.assert_close)."""
    net = cases.build_model(hg, case).cuda()
    got = _outs(net(clone_data(cases.build_input(cases.CASES[case][2]), "cuda")))
    assert len(got) == len(GOLD[case])
    for k, (g, w) in enumerate(zip(got, GOLD[case])):
        assert_close(g, w, f"{case}[{k}]")


@pytest.mark.parametrize("case", ["nt_het_gat", "nt_hom_gcn", "nt_hom_sage"])
def test_neural_tree_larger_batch_vs_oracle(case):
    """config 3 shape: the fixture H-trees replicated into a 64-tree batch (ragged, 6 node types)."""
    trees = cases.load_htrees("htree_synth")
    order = [k % len(trees) for k in range(64)]
    if case.startswith("nt_hom"):
        homs = [convert.heterogeneous_htree_to_homogeneous(t) for t in trees]
        data = Batch.from_data_list([homs[k] for k in order])
    else:
        data = Batch.from_data_list([trees[k] for k in order])
    ref = cases.build_model(om, case, seed=3)
    net = cases.build_model(hg, case, seed=3).cuda()
    with torch.no_grad():
        want = _outs(ref(clone_data(data)))
    got = _outs(net(clone_data(data, "cuda")))
    for g, w in zip(got, want):
        assert_close(g, w, case)


def test_plan_cache_and_repeat_forward_are_idempotent():
    net = cases.build_model(hg, "mgat").cuda()
    data = clone_data(cases.build_input("synth16"), "cuda")
    a = net(data).clone()
    b = net(data)  # cached plan
    net.plan_cache.enabled = False
    c = net(data)  # plan rebuilt
    assert torch.equal(a, b) and torch.equal(a, c)


def test_empty_relations_and_single_graph():
    """Edge cases of the reference's data: a graph whose rooms_to_rooms is empty [2,0]; one-room graphs."""
    st = cases.load_stanford()
    data = Batch.from_data_list([cases.stanford_hetero(st[k]) for k in (0, 1)])
    assert data["rooms", "rooms_to_rooms", "rooms"].edge_index.size(1) == 0
    for case in ("msage_all",):
        ref, net = cases.build_model(om, case), cases.build_model(hg, case).cuda()
        with torch.no_grad():
            want = _outs(ref(clone_data(data)))
        got = _outs(net(clone_data(data, "cuda")))
        for g, w in zip(got, want):
            assert_close(g, w, case)
    kw = dict(input_dim_dict={"objects": 6, "rooms": 6}, output_dim=15, conv_block="GAT", GAT_hidden_dims=[32],
              GAT_heads=[2, 2], GAT_concats=[True, False])
    ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=5).eval()
    net = hg.HeterogeneousNetwork(**kw)
    net.load_state_dict(ref.state_dict())
    net = net.eval().cuda()
    with torch.no_grad():
        want = ref(clone_data(data))
    assert_close(net(clone_data(data, "cuda")), want, "GAT on graphs without room-room edges")
