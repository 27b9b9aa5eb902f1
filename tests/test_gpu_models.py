"""GPU parity of the four drop-in model classes against the CPU oracle (calls go through the
C ABI: hydra_gnn_b200.models -> hydra_gnn_b200.ops -> libhydra_gnn_b200.so)."""
import pytest
import torch

import hydra_gnn_b200 as hg
from hydra_gnn_b200 import synth
from oracle import models as om
from util import assert_close, clone_data, make_pair

pytestmark = pytest.mark.gpu

M_GAT = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
             GAT_heads=[3, 3, 3], GAT_concats=[False, False, False], dropout=0.4)


def _run_pair(ref, prod, data):
    with torch.no_grad():
        want = ref(clone_data(data))
    got = prod(clone_data(data, "cuda"))
    return got, want


@pytest.mark.parametrize("n_graphs", [1, 7, 256])
def test_hetero_gat_mgat(n_graphs):
    data = synth.synth_mp3d_batch(range(n_graphs))
    ref, prod = make_pair(om.HeterogeneousNetwork, hg.HeterogeneousNetwork, **M_GAT)
    got, want = _run_pair(ref, prod, data)
    assert_close(got, want, f"M-GAT logits ({n_graphs} graphs)")


@pytest.mark.parametrize("heads,concats", [([6, 6], [True, False]), ([5, 5, 5], [False, False, False]),
                                           ([2, 2], [True, True])])
def test_hetero_gat_variants(heads, concats):
    data = synth.synth_mp3d_batch(range(40, 72), with_word2vec=False)
    kw = dict(input_dim_dict={"objects": 6, "rooms": 6}, output_dim_dict={"rooms": 15, "objects": 35},
              conv_block="GAT", GAT_hidden_dims=[64] * (len(heads) - 1), GAT_heads=heads, GAT_concats=concats)
    ref, prod = make_pair(om.HeterogeneousNetwork, hg.HeterogeneousNetwork, **kw)
    got, want = _run_pair(ref, prod, data)
    assert_close(got[0], want[0], "rooms")
    assert_close(got[1], want[1], "objects")


def test_hetero_gat_edge():
    data = synth.synth_mp3d_batch(range(64), relative_pos=True)
    kw = dict(M_GAT, input_dim_dict={"objects": 303, "rooms": 3}, conv_block="GAT_edge")
    ref, prod = make_pair(om.HeterogeneousNetwork, hg.HeterogeneousNetwork, **kw)
    got, want = _run_pair(ref, prod, data)
    assert_close(got, want, "M-GATe logits")


@pytest.mark.parametrize("task", ["room", "all"])
def test_hetero_sage(task):
    data = synth.synth_mp3d_batch(range(100, 164), with_word2vec=False)
    kw = dict(input_dim_dict={"objects": 6, "rooms": 6}, conv_block="GraphSAGE", hidden_dim=128, num_layers=3)
    kw.update(dict(output_dim=15) if task == "room" else dict(output_dim_dict={"rooms": 15, "objects": 35}))
    ref, prod = make_pair(om.HeterogeneousNetwork, hg.HeterogeneousNetwork, **kw)
    got, want = _run_pair(ref, prod, data)
    if task == "room":
        assert_close(got, want, "M-SAGE rooms")
    else:
        assert_close(got[0], want[0], "M-SAGE rooms")
        assert_close(got[1], want[1], "M-SAGE objects")
