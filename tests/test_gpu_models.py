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


# ------------------------------------------------------------------------------ BASELINE.json configs[1] at full size
def test_mgat_full_size_vs_oracle_and_split_invariance():
    """4096 graphs (the bench workload: pair-mode tcgen05 GEMM, 8-CTA/SM GAT build, side-stream plan).
    (1) logits vs the CPU oracle on the same batch; (2) graphs are independent components, so the
    forward of the batch equals the concatenation of the forwards of its two halves."""
    n = 4096
    ref, prod = make_pair(om.HeterogeneousNetwork, hg.HeterogeneousNetwork, **M_GAT)
    data = synth.synth_mp3d_batch(range(n))
    got, want = _run_pair(ref, prod, data)
    assert_close(got, want, "M-GAT logits (4096 graphs)")
    halves = [prod(synth.synth_mp3d_batch(r, device="cuda")) for r in (range(0, n // 2), range(n // 2, n))]
    assert_close(torch.cat(halves), got, "batch == concat(halves)", rtol=1e-5, atol=1e-6)


def test_mgat_full_size_cuda_graph_replay_is_bit_identical():
    """The bench replays the step as a CUDA graph (forked side streams included): same bits as eager."""
    prod = synth.random_state_dict_(hg.HeterogeneousNetwork(**M_GAT), seed=0).eval().cuda()
    prod.plan_cache.enabled = False
    data = synth.synth_mp3d_batch(range(4096), device="cuda")
    eager = prod(data).clone()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        prod(data)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = prod(data)
    torch.cuda.synchronize()
    for _ in range(3):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)


@pytest.mark.parametrize("conv_block,kw", [("GCN", dict(hidden_dim=128, num_layers=4)),
                                           ("GraphSAGE", dict(hidden_dim=128, num_layers=4)),
                                           ("GIN", dict(hidden_dim=128, num_layers=3))])
def test_homogeneous_full_size_vs_oracle(conv_block, kw):
    """seg_reduce tile kernel at scale: homogeneous conversion of 2048 graphs (~130 k nodes)."""
    from hydra_gnn_b200 import convert
    data = convert.dsg_to_homogeneous(synth.synth_mp3d_batch(range(2048)))
    ref, prod = make_pair(om.HomogeneousNetwork, hg.HomogeneousNetwork, input_dim=306, output_dim=26,
                          conv_block=conv_block, **kw)
    got, want = _run_pair(ref, prod, data)
    assert_close(got, want, f"homogeneous {conv_block} logits (2048 graphs)")
