"""Data formats either side of the forward, on the device: relative positions
(``hg_relative_pos`` behind ``convert.compute_relative_pos``), the hetero -> homogeneous conversions on CUDA tensors,
and the Neural-Tree models at the BASELINE.json configs[2] size (4096 H-trees) against the CPU oracle."""
import os
import sys

import numpy as np
import pytest
import torch

import hydra_gnn_b200 as hg
from hydra_gnn_b200 import convert, htree, ops, synth
from hydra_gnn_b200.data import Batch, HeteroData
from oracle import models as om
from util import assert_close, clone_data, make_pair

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import cases  # noqa: E402

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------------------ a13: compute_relative_pos
def test_relative_pos_kernel_is_bit_identical_to_the_indexing_expression():
    """mp3d_dataset.py:316-319: ``edge_attr = pos[target][edge_index[1]] - pos[source][edge_index[0]]``."""
    g = torch.Generator().manual_seed(5)
    for n_src, n_dst, E in [(1, 1, 1), (7, 3, 40), (1000, 1234, 100_003), (5, 5, 0)]:
        ps, pd = torch.randn(n_src, 3, generator=g) * 10, torch.randn(n_dst, 3, generator=g) * 10
        ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), torch.randint(0, n_dst, (E,), generator=g)])
        want = pd[ei[1]] - ps[ei[0]]
        got = ops.relative_pos(ps.cuda(), pd.cuda(), ei.cuda())
        assert got.shape == (E, 3)
        assert torch.equal(got.cpu(), want)


def test_relative_pos_counts_out_of_range_endpoints_without_dereferencing_them():
    ps, pd = torch.ones(4, 3).cuda(), torch.zeros(2, 3).cuda()
    ei = torch.tensor([[0, 4, 3, -1], [1, 0, 2, 0]]).cuda()  # edges 1, 2, 3 are out of range
    status = torch.zeros(1, dtype=torch.int32, device="cuda")
    got = ops.relative_pos(ps, pd, ei, status=status)
    assert int(status) == 3
    assert torch.equal(got[0].cpu(), torch.tensor([-1.0, -1.0, -1.0]))
    assert bool(torch.isnan(got[1:]).all())


def test_compute_relative_pos_on_the_device_equals_the_host_route():
    """The GAT_edge batches the models consume: positions dropped from x, edge_attr attached for every relation."""
    host = synth.synth_mp3d_batch(range(64), relative_pos=True)           # host indexing
    dev = synth.synth_mp3d_batch(range(64), relative_pos=True, device="cuda")  # hg_relative_pos
    assert set(dev.edge_attr_dict) == set(host.edge_attr_dict) == set(hg.schema.EDGE_TYPES)
    for et in hg.schema.EDGE_TYPES:
        assert dev[et].edge_attr.is_cuda
        assert torch.equal(dev[et].edge_attr.cpu(), host[et].edge_attr), et
    for nt in ("objects", "rooms"):
        assert torch.equal(dev[nt].x.cpu(), host[nt].x)
    with pytest.raises(Warning):  # mp3d_dataset.py:304-307
        convert.compute_relative_pos(dev)


# ------------------------------------------------------------------------------ a12 / f2: conversions on CUDA tensors
def _to_cuda(d: HeteroData) -> HeteroData:
    return clone_data(d, "cuda")


def test_convert_on_cuda_tensors_matches_the_reference_generated_fixture():
    """``heterogeneous_htree_to_homogeneous`` / ``heterogeneous_data_to_homogeneous`` (mp3d_dataset.py:29-119) applied
    to device-resident graphs against tests/golden/convert_golden.npz (written by the reference's own functions)."""
    z = np.load(os.path.join(ROOT, "tests", "golden", "convert_golden.npz"))
    for k, t in enumerate(cases.load_htrees("htree_synth")[:3]):
        hom = convert.heterogeneous_htree_to_homogeneous(_to_cuda(t))
        for key in ("edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask", "node_type"):
            assert hom[key].is_cuda
            assert np.array_equal(hom[key].cpu().numpy(), z[f"htree{k}|{key}"]), (k, key)
        assert np.array_equal(hom.x.double().sum(0).cpu().numpy(), z[f"htree{k}|x_sum"])
    for k, gid in enumerate([0, 3]):
        hom, types = convert.heterogeneous_data_to_homogeneous(synth.synth_mp3d_batch([gid], device="cuda"))
        assert types == ["objects", "rooms"]
        assert np.array_equal(hom.edge_index.cpu().numpy(), z[f"dsg{k}|edge_index"])
        assert np.array_equal(hom.node_type.cpu().numpy(), z[f"dsg{k}|node_type"])


def test_batched_htree_conversion_on_the_device_equals_the_per_graph_host_route():
    trees = [htree.generate_htree(synth.synth_mp3d_batch([g])) for g in range(12)]
    want = Batch.from_data_list([convert.heterogeneous_htree_to_homogeneous(clone_data(t)) for t in trees])
    got = convert.batched_htree_to_homogeneous(_to_cuda(Batch.from_data_list([clone_data(t) for t in trees])))
    for key in ("x", "edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask", "node_type"):
        assert got[key].is_cuda
        assert torch.equal(got[key].cpu(), want[key]), key


def test_htree_split_kernel_matches_the_tensor_op_pass_and_the_fixture():
    """hg_htree_split (the device form of mp3d_dataset.py:72-119 + collate): every integer output equals the tensor-op
    pass on a 300-tree batch, the per-graph slices of a fixture batch equal tests/golden/convert_golden.npz (written by
    the reference's own function), and edges out of graph order are counted, not misplaced."""
    het = htree.generate_htree_batch(synth.synth_mp3d_batch(range(300)))
    dev = _to_cuda(het)
    got = convert.batched_htree_to_homogeneous(dev)
    assert int(convert.last_split_status.item()) == 0
    try:
        convert.USE_SPLIT_KERNEL = False
        want = convert.batched_htree_to_homogeneous(_to_cuda(het))
    finally:
        convert.USE_SPLIT_KERNEL = True
    for key in ("x", "edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask", "node_type", "batch", "ptr"):
        assert got[key].dtype == want[key].dtype and torch.equal(got[key], want[key]), key
    assert got.num_graphs == want.num_graphs == 300
    # reference-generated fixture: graph k of the collated batch, shifted back by its node offset
    z = np.load(os.path.join(ROOT, "tests", "golden", "convert_golden.npz"))
    trees = cases.load_htrees("htree_synth")[:3]
    hom = convert.batched_htree_to_homogeneous(_to_cuda(Batch.from_data_list([clone_data(t) for t in trees])))
    ptr = hom.ptr.cpu()
    for k in range(3):
        lo, hi = int(ptr[k]), int(ptr[k + 1])
        for key in ("edge_index", "init_edge_index", "pool_edge_index"):
            ei = hom[key].cpu()
            mine = ei[:, (ei[0] >= lo) & (ei[0] < hi)] - lo
            assert np.array_equal(mine.numpy(), z[f"htree{k}|{key}"]), (k, key)
        for key in ("room_mask", "object_mask", "node_type"):
            assert np.array_equal(hom[key].cpu()[lo:hi].numpy(), z[f"htree{k}|{key}"]), (k, key)
    # a relation whose edges are not grouped by graph: counted in the status word
    bad = _to_cuda(het)
    et = bad.edge_types[0]
    bad[et].edge_index = bad[et].edge_index.flip(1).contiguous()
    convert.batched_htree_to_homogeneous(bad)
    assert int(convert.last_split_status.item()) > 0


# ------------------------------------------------------------------------------ a9 at the configs[2] size
HT_DIMS = {"object": 306, "room": 6, "object-room": 6, "room-room": 6, "object_virtual": 306, "room_virtual": 6}
_NT = {}


def _nt_inputs(n=4096):
    """H-trees of graph ids 0..n-1 from the native builder (bit-identical to the reference's networkx code:
    tests/test_htree.py), in both forms; built once for the three tests below."""
    if n not in _NT:
        het = htree.generate_htree_batch(synth.synth_mp3d_batch(range(n)))
        slim = HeteroData()
        for t in het.node_types:
            slim[t].x = het[t].x
            slim[t].num_nodes = het[t].x.size(0)
        for et in het.edge_types:
            slim[et].edge_index = het[et].edge_index
        _NT[n] = (slim, convert.batched_htree_to_homogeneous(het))
    return _NT[n]


def _hom_view(hom, device=None):
    d = hg.data.Data()
    for k in ("x", "edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask"):
        setattr(d, k, getattr(hom, k).to(device) if device else getattr(hom, k))
    return d


@pytest.mark.parametrize("conv_block", ["GCN", "GraphSAGE"])
def test_homogeneous_neural_tree_at_config3_size_vs_oracle(conv_block):
    """M-NT (i), (ii) of SURVEY.md section 8: 4096 H-trees (~3.2 M tree nodes) as one batch."""
    _, hom = _nt_inputs()
    ref, prod = make_pair(om.HomogeneousNeuralTreeNetwork, hg.HomogeneousNeuralTreeNetwork, input_dim=306, output_dim=26,
                          conv_block=conv_block, hidden_dim=128, num_layers=4, disable_initialization=True)
    with torch.no_grad():
        want = ref(_hom_view(hom))
    got = prod(_hom_view(hom, "cuda"))
    assert got.shape[0] == int(hom.room_mask.sum())
    assert_close(got, want, f"NT homogeneous {conv_block} (4096 trees)")


@pytest.mark.parametrize("pre_mp", [False, True])
def test_heterogeneous_neural_tree_gat_at_config3_size_vs_oracle(pre_mp):
    """M-NT (iii): heterogeneous GAT over the 10 H-tree relations + LeafPool readout, with and without ``pre_mp``."""
    het, _ = _nt_inputs()
    ref, prod = make_pair(om.HeterogeneousNeuralTreeNetwork, hg.HeterogeneousNeuralTreeNetwork, input_dim_dict=HT_DIMS,
                          output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64], GAT_heads=[3, 3, 3],
                          GAT_concats=[False] * 3, disable_initialization=not pre_mp)
    with torch.no_grad():
        want = ref(clone_data(het))
    got = prod(clone_data(het, "cuda"))
    assert_close(got, want, f"NT heterogeneous GAT (4096 trees, pre_mp={pre_mp})")


# ------------------------------------------------------------------------------ device generator (configs[3] streaming)
@pytest.mark.parametrize("first,n", [(0, 1), (0, 64), (12345, 257), (999_000, 1500), (5, 10_000)])
def test_device_generator_reproduces_the_host_generator(first, n):
    """hg_synth_mp3d_counts / _fill vs synth.synth_mp3d_batch on the same graph ids: every integer output and every
    feature that is an affine map of a uniform draw bit-identical; Box-Muller positions to the last float32 bit or two."""
    host = synth.synth_mp3d_batch(range(first, first + n))
    gen = synth.DeviceSynth("cuda")
    sizes = gen.sizes(first, n)
    assert sizes[0] == host["rooms"].x.size(0) and sizes[1] == host["objects"].x.size(0)
    dev = gen.batch(first, n, sizes)
    assert int(dev.synth_status.item()) == 0
    for et in dev.edge_types:
        assert torch.equal(dev[et].edge_index.cpu(), host[et].edge_index), et
    for t in ("rooms", "objects"):
        assert torch.equal(dev[t].y.cpu(), host[t].y) and torch.equal(dev[t].batch.cpu(), host[t].batch), t
    assert torch.equal(dev["rooms"].x.cpu(), host["rooms"].x)
    xo, want = dev["objects"].x.cpu(), host["objects"].x
    assert torch.equal(xo[:, 3:], want[:, 3:])  # sizes + word2vec rows
    # positions: room position + 1.5 * Box-Muller normal, log / cos / sqrt in double on both sides
    diff = (xo[:, :3] - want[:, :3]).abs()
    assert float(diff.max()) <= 4e-6 and float((diff > 0).float().mean()) < 1e-3


def test_device_generator_prefix_sums_and_size_check():
    gen = synth.DeviceSynth("cuda", with_word2vec=False)
    prefix = gen.counts_async(7, 3000).cpu()
    host = synth.synth_mp3d_batch(range(7, 3007), with_word2vec=False)
    rooms = torch.bincount(host["rooms"].batch, minlength=3000)
    objects = torch.bincount(host["objects"].batch, minlength=3000)
    assert torch.equal(prefix[1:, 0] - prefix[:-1, 0], rooms) and torch.equal(prefix[1:, 1] - prefix[:-1, 1], objects)
    assert int(prefix[-1, 2]) * 2 == host["rooms", "rooms_to_rooms", "rooms"].edge_index.size(1)
    assert int(prefix[-1, 3]) * 2 == host["objects", "objects_to_objects", "objects"].edge_index.size(1)
    sizes = gen.sizes(7, 3000)
    bad = gen.batch(7, 3000, (sizes[0], sizes[1] + 1, sizes[2], sizes[3]))  # wrong declared size: flagged, nothing written
    assert int(bad.synth_status.item()) == 1


def test_forward_on_a_device_generated_batch_matches_the_host_generated_one():
    net = synth.random_state_dict_(hg.HeterogeneousNetwork(**cases.CASES["mgat"][1]), seed=0).eval().cuda()
    host = synth.synth_mp3d_batch(range(500, 564), device="cuda")
    dev = synth.DeviceSynth("cuda").batch(500, 64)
    with torch.no_grad():
        a, b = net(host), net(dev)
    assert a.shape == b.shape
    assert_close(b, a.double().cpu(), "forward on the device-generated batch")
