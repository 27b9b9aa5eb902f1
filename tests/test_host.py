"""CPU tests of the host side (no GPU, no compute calls into the library): the C-ABI library
loads and exports every symbol include/hydra_gnn_b200.h declares; containers / collate /
hetero->homogeneous conversions are bit-exact with the reference's (committed golden, made by
the reference's own functions); the synthetic generator is shard-invariant; weight packing
algebra; the product fails loudly without a GPU; world_size-2 gloo run of the sharding path."""
import os
import re
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import cases  # noqa: E402
import hydra_gnn_b200 as hg  # noqa: E402
from hydra_gnn_b200 import _lib, convert, ops, synth  # noqa: E402
from hydra_gnn_b200 import dist as hdist  # noqa: E402
from hydra_gnn_b200 import nn as hnn  # noqa: E402
from hydra_gnn_b200.data import Batch, HeteroData  # noqa: E402
from oracle import models as om  # noqa: E402
from oracle.pyg import data as od  # noqa: E402


# ------------------------------------------------------------------------------ C ABI
def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "hydra_gnn_b200.h")).read()
    declared = set(re.findall(r"HG_API\s+[\w\s\*]+?\b(hg_\w+)\s*\(", header))
    assert len(declared) >= 14
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    h = _lib.load_shared_object()  # dlopen, prototypes, struct-size check
    for name in declared:
        assert hasattr(h, name)
    assert h.hg_abi_version() == 2
    assert h.hg_strerror(0) == b"ok" and b"workspace" in h.hg_strerror(-3)
    for which, st in enumerate(_lib._STRUCTS):
        import ctypes

        assert h.hg_sizeof(which) == ctypes.sizeof(st)


def test_training_entry_points_reject_bad_arguments_on_the_host():
    """Error behaviour of the training ABI (include/hydra_gnn_b200.h): argument checks run before any CUDA call and
    return the documented negative codes (no launch, so this runs without a GPU)."""
    import ctypes as C

    h = _lib.load_shared_object()
    EINVAL, EUNSUPPORTED = -1, -6
    assert h.hg_strerror(EINVAL) == b"invalid argument" and b"unsupported" in h.hg_strerror(EUNSUPPORTED)
    # hg_linear_wgrad: negative sizes, leading dimensions shorter than the row
    assert h.hg_linear_wgrad(None, 0, None, 0, -1, 4, 4, None, 4, 0, None) == EINVAL
    assert h.hg_linear_wgrad(None, 2, None, 4, 8, 4, 4, None, 4, 0, None) == EINVAL
    assert h.hg_linear_wgrad(None, 4, None, 4, 8, 4, 4, None, 2, 0, None) == EINVAL
    assert h.hg_linear_wgrad(None, 4, None, 4, 8, 4, 4, None, 4, 0, None) == EINVAL  # NULL dW
    assert h.hg_linear_wgrad(None, 4, None, 4, 8, 0, 4, None, 4, 0, None) == 0       # empty product: nothing to do
    # hg_adam_step: n < 0; step < 1 without a device counter
    assert h.hg_adam_step(None, None, None, None, -1, 0.1, 0.9, 0.999, 1e-8, 0.0, 1.0, 1, None, None) == EINVAL
    assert h.hg_adam_step(None, None, None, None, 5, 0.1, 0.9, 0.999, 1e-8, 0.0, 1.0, 0, None, None) == EINVAL
    assert h.hg_adam_step(None, None, None, None, 0, 0.1, 0.9, 0.999, 1e-8, 0.0, 1.0, 1, None, None) == 0
    # hg_gat_train_*: NULL, head / width limits of the kernels, dropout probability outside [0, 1)
    assert h.hg_gat_train_fwd(None, None) == EINVAL and h.hg_gat_train_bwd(None, None) == EINVAL
    a = _lib.GatTrainArgs()
    a.n_rel, a.n_dst, a.H, a.C = 1, 0, 5, 8
    assert h.hg_gat_train_fwd(C.byref(a), None) == EUNSUPPORTED   # more than 4 heads
    a.H, a.C = 4, 65
    assert h.hg_gat_train_fwd(C.byref(a), None) == EUNSUPPORTED   # H * C > 256
    a.H, a.C, a.dropout_p = 2, 8, 1.0
    assert h.hg_gat_train_fwd(C.byref(a), None) == EINVAL
    a.dropout_p, a.n_rel = 0.0, 0
    assert h.hg_gat_train_fwd(C.byref(a), None) == EINVAL
    assert h.hg_csr_transpose_map(None, None, None, -1, None, None, None, 0, 0, None, None, None) == EINVAL
    assert h.hg_csr_transpose_map(None, None, None, 0, None, None, None, 0, 0, None, None, None) == 0
    assert h.hg_act_bwd(None, 0, None, 0, -1, 4, 0, None, 0, None, None) == EINVAL
    assert h.hg_act_bwd(None, 0, None, 0, 0, 4, 0, None, 0, None, None) == 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_product_fails_loudly_without_gpu():
    net = hg.HeterogeneousNetwork(**cases.CASES["mgat"][1]).eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        net(synth.synth_mp3d_batch(range(2)))


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "hydra_gnn_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f


# ------------------------------------------------------------------------------ containers
def _to_od(g):
    h = od.HeteroData()
    for nt in g.node_types:
        for k, v in g[nt].items():
            h[nt][k] = v
    for et in g.edge_types:
        for k, v in g[et].items():
            h[et][k] = v
    return h


def test_batch_collate_bit_exact_with_pyg_rule():
    graphs = [synth.synth_mp3d_batch([gid], relative_pos=True) for gid in (3, 0, 7)]
    ours = Batch.from_data_list(graphs)
    want = od.Batch.from_data_list([_to_od(g) for g in graphs])
    for nt in ("objects", "rooms"):
        assert torch.equal(ours[nt].x, want[nt].x) and torch.equal(ours[nt].batch, want[nt].batch)
    for et in ours.edge_types:
        assert ours[et].edge_index.dtype == torch.int64
        assert torch.equal(ours[et].edge_index, want[et].edge_index)
        assert torch.equal(ours[et].edge_attr, want[et].edge_attr)
    # a batch generated in one go is the same disjoint union
    whole = synth.synth_mp3d_batch([3, 0, 7], relative_pos=True)
    for et in ours.edge_types:
        assert torch.equal(ours[et].edge_index, whole[et].edge_index)
    assert torch.equal(ours["objects"].x, whole["objects"].x)


def test_homogeneous_batch_offsets_index_attributes():
    trees = cases.load_htrees("htree_synth")[:3]
    homs = [convert.heterogeneous_htree_to_homogeneous(t) for t in trees]
    b = Batch.from_data_list(homs)
    off = 0
    for h in homs:
        n = h.x.size(0)
        for key in ("edge_index", "init_edge_index", "pool_edge_index"):
            sel = (b[key][0] >= off) & (b[key][0] < off + n)
            assert torch.equal(b[key][:, sel] - off, h[key])
        off += n
    assert b.room_mask.sum() == sum(int(h.room_mask.sum()) for h in homs)


def test_convert_bit_exact_with_reference_functions():
    z = np.load(os.path.join(ROOT, "tests", "golden", "convert_golden.npz"))
    for k, t in enumerate(cases.load_htrees("htree_synth")[:3]):
        hom = convert.heterogeneous_htree_to_homogeneous(t)
        for key in ("edge_index", "init_edge_index", "pool_edge_index", "room_mask", "object_mask", "node_type"):
            assert np.array_equal(hom[key].numpy(), z[f"htree{k}|{key}"]), (k, key)
        assert np.allclose(hom.x.double().sum(0).numpy(), z[f"htree{k}|x_sum"], rtol=0, atol=0)
        assert hom.x.size(1) == 306
    for k, gid in enumerate([0, 3]):
        hom, types = convert.heterogeneous_data_to_homogeneous(synth.synth_mp3d_batch([gid]))
        assert types == ["objects", "rooms"]
        assert np.array_equal(hom.edge_index.numpy(), z[f"dsg{k}|edge_index"])
        assert np.array_equal(hom.node_type.numpy(), z[f"dsg{k}|node_type"])
        assert np.array_equal(hom.x.double().sum(0).numpy(), z[f"dsg{k}|x_sum"])


def test_fill_missing_edge_index_rule_of_the_stanford_loader():
    """tests/test_mp3d_dataset.py:143-180 of the reference: missing rooms_to_rooms -> empty [2,0] int64;
    missing objects_to_rooms -> rooms_to_objects.flip([0])."""
    g = cases.stanford_hetero(cases.load_stanford()[5])
    rr = g["rooms", "rooms_to_rooms", "rooms"].edge_index
    assert rr.shape == (2, 0) and rr.dtype == torch.int64
    assert torch.equal(g["objects", "objects_to_rooms", "rooms"].edge_index,
                       g["rooms", "rooms_to_objects", "objects"].edge_index.flip([0]))
    assert hg.schema.EDGE_TYPES == list(g.edge_types)


def test_htree_fixture_invariants():
    for t in cases.load_htrees("htree_synth") + cases.load_htrees("htree_stanford"):
        n_leaf, n_rv = t["room"].x.size(0), t["room_virtual"].x.size(0)
        pool = t["room", "r_to_rv", "room_virtual"].edge_index
        assert pool.size(1) == n_leaf and n_rv <= n_leaf and int(pool[1].max()) < n_rv
        assert torch.equal(pool[0].sort().values, torch.arange(n_leaf))  # every leaf pools exactly once
        for s, r, d in hg.schema.HTREE_EDGE_TYPES:  # tree relations come in mirrored pairs
            ei = t[s, r, d].edge_index
            assert ei.size(1) == 0 or (int(ei[0].max()) < t[s].x.size(0) and int(ei[1].max()) < t[d].x.size(0))


# ------------------------------------------------------------------------------ generator
def test_synth_is_shard_invariant_and_shaped_like_mp3d():
    whole = synth.synth_mp3d_batch(range(64))
    lo, hi = hdist.shard_range(64, 1, 4)
    part = synth.synth_mp3d_batch(range(lo, hi))
    first = int((whole["objects"].batch < lo).sum())
    n = part["objects"].x.size(0)
    assert torch.equal(whole["objects"].x[first:first + n], part["objects"].x)
    G = 64
    assert 3.5 < whole["rooms"].x.size(0) / G < 6.5 and 40 < whole["objects"].x.size(0) / G < 80
    oo = whole["objects", "objects_to_objects", "objects"].edge_index
    assert not bool((oo[0] == oo[1]).any())
    assert not bool((oo[1][1:] >= oo[1][:-1]).all()), "edges must arrive unsorted by destination"
    assert whole["objects"].x.shape[1] == 306 and whole["rooms"].x.shape[1] == 6


def test_shard_ranges_partition_the_ids():
    for n, w in [(10, 3), (1_000_000, 8), (5, 8), (4096, 1)]:
        spans = [hdist.shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        assert max(h - l for l, h in spans) - min(h - l for l, h in spans) <= 1


# ------------------------------------------------------------------------------ packing algebra / names
def test_attention_fold_equals_projection_then_dot():
    g = torch.Generator().manual_seed(0)
    H, C, K = 3, 8, 11
    W, att, x = torch.randn(H * C, K, generator=g), torch.randn(1, H, C, generator=g), torch.randn(5, K, generator=g)
    want = ((x @ W.t()).view(-1, H, C) * att).sum(-1)
    got = x @ hnn.GATConv.fold(att, W, H, C).t()
    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("case", list(cases.CASES))
def test_product_state_dict_interchanges_with_reference_names(case):
    cls, kw, _ = cases.CASES[case]
    ref_sd = cases.build_model(om, case).state_dict()
    net = getattr(hg, cls)(**kw)
    assert list(net.state_dict().keys()) == list(ref_sd.keys())
    net.load_state_dict(ref_sd, strict=True)
    for k, v in net.state_dict().items():
        assert v.shape == ref_sd[k].shape, k


# ------------------------------------------------------------------------------ world_size 2 (gloo)
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, n_graphs, out):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = hdist.shard_range(n_graphs, rank, world)
        data = synth.synth_mp3d_batch(range(lo, hi), with_word2vec=False)
        rooms = data["rooms"].x
        # stand-in for the per-rank forward (the CUDA path needs a GPU): a fixed linear map of the room rows
        logits = torch.cat([rooms, rooms.sum(1, keepdim=True)], dim=1)
        gather = hdist.LogitGather(7).prepare(logits.size(0), logits.device)
        gather(logits)
        first = gather.unpad().clone()
        # the pipelined form of bench.py's N > 1 loop: two slots, step k staged while step k - 1 is in flight
        piped = hdist.LogitGather(7).prepare(logits.size(0), logits.device, slots=2)
        steps = []
        for k in range(3):
            slot = k % 2
            piped.wait(slot)
            piped.stage(logits * (k + 1), slot)
            piped.launch(slot)
            piped.wait(slot)
            steps.append(piped.unpad(slot).clone())
        if rank == 0:
            out.put((gather.counts, first.numpy(), [t.numpy() for t in steps]))  # by value: no fd passing from a process about to exit
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_gloo_world2_shard_and_gather_equals_single_process():
    n_graphs, world = 13, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, world, port, n_graphs, q)) for r in range(world)]
    for p in procs:
        p.start()
    counts, gathered, steps = q.get(timeout=300)
    gathered, steps = torch.from_numpy(gathered), [torch.from_numpy(t) for t in steps]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rooms = synth.synth_mp3d_batch(range(n_graphs), with_word2vec=False)["rooms"].x
    want = torch.cat([rooms, rooms.sum(1, keepdim=True)], dim=1)
    assert sum(counts) == want.size(0)
    assert torch.equal(gathered, want)
    for k, got in enumerate(steps):  # every step of the two-slot pipeline lands in graph order with its own values
        assert torch.equal(got, want * (k + 1))


def test_checkpoint_keys_of_newer_pyg_are_accepted():
    """accept the PyG >= 2.4 spelling on load (hetero '<a___b___c>' keys, shared GAT 'lin')."""
    net = hg.HeterogeneousNetwork(**cases.CASES["mgat"][1])
    sd = net.state_dict()
    new = {}
    for k, v in sd.items():
        parts = k.split(".")
        if len(parts) > 3 and parts[2] == "convs" and "__" in parts[3]:
            parts[3] = "<" + "___".join(parts[3].split("__")) + ">"
        new[".".join(parts)] = v.clone() + 1.0
    other = hg.HeterogeneousNetwork(**cases.CASES["mgat"][1])
    other.load_state_dict(new, strict=True)
    for k, v in other.state_dict().items():
        assert torch.equal(v, sd[k] + 1.0), k
    # homogeneous GAT: one shared projection stored as lin.weight
    hom = hg.HomogeneousNetwork(**cases.CASES["hom_gat"][1])
    sd = hom.state_dict()
    new = {}
    for k, v in sd.items():
        if k.endswith("lin_dst.weight"):
            continue
        new[k.replace("lin_src.weight", "lin.weight")] = v.clone() * 2.0
    other = hg.HomogeneousNetwork(**cases.CASES["hom_gat"][1])
    other.load_state_dict(new, strict=True)
    for k, v in other.state_dict().items():
        assert torch.equal(v, sd[k] * 2.0), k


# ------------------------------------------------------------------------------ training plumbing (host side)
def test_flat_adam_rehomes_live_parameters_as_views():
    """FlatAdam.flatten(): parameters that received a gradient become views of ONE flat buffer (so a step is one kernel
    and a gradient exchange one all-reduce); dead parameters (no gradient, like PyG's unused lin_dst) stay out."""
    from hydra_gnn_b200.train import FlatAdam
    torch.manual_seed(0)
    a, b, dead = (torch.nn.Parameter(torch.randn(3, 4)), torch.nn.Parameter(torch.randn(5)),
                  torch.nn.Parameter(torch.randn(2)))
    before = [a.detach().clone(), b.detach().clone()]
    opt = FlatAdam([a, b, dead, a], lr=1e-3)  # a listed twice: shared parameters are taken once
    ((a * 2).sum() + (b * 3).sum()).backward()
    opt.flatten()
    assert opt.flat_param.numel() == 17 and [id(p) for p in opt.live] == [id(a), id(b)]
    assert torch.equal(a.detach(), before[0]) and torch.equal(b.detach(), before[1])
    assert a.data_ptr() == opt.flat_param.data_ptr() and b.data_ptr() == opt.flat_param[12:].data_ptr()
    assert torch.equal(opt.flat_grad, torch.cat([torch.full((12,), 2.0), torch.full((5,), 3.0)]))
    opt.zero_grad()
    assert float(a.grad.abs().sum()) == 0.0 and a.grad.data_ptr() == opt.flat_grad.data_ptr()
    ((a * 5).sum()).backward()  # autograd accumulates in place into the view
    assert torch.equal(opt.flat_grad[:12], torch.full((12,), 5.0)) and dead.grad is None


def _gloo_grad_worker(rank, world, port, out):
    import torch.distributed as dist
    from hydra_gnn_b200.train import FlatAdam
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)  # identical replicas
        w = torch.nn.Parameter(torch.randn(4, 3))
        b = torch.nn.Parameter(torch.randn(4))
        opt = FlatAdam([w, b])
        x = torch.arange(6, dtype=torch.float32).reshape(2, 3) + rank  # each rank its own shard
        n_global = 2.0 * world
        ((x @ w.t() + b).sum() / n_global).backward()  # local sum / global count
        opt.flatten()
        dist.all_reduce(opt.flat_grad)  # the one collective of the training step
        out.put((rank, opt.flat_grad.detach().numpy().copy(), w.grad.detach().numpy().copy()))  # by value (see above)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_gradient_all_reduce_gives_the_global_mean():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q, world, port = ctx.Queue(), 2, _free_port()
    procs = [ctx.Process(target=_gloo_grad_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=300) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
    torch.manual_seed(0)
    w = torch.randn(4, 3, requires_grad=True)
    b = torch.randn(4, requires_grad=True)
    x = torch.cat([torch.arange(6, dtype=torch.float32).reshape(2, 3) + r for r in range(world)])
    ((x @ w.t() + b).sum() / x.size(0)).backward()
    want = torch.cat([w.grad.reshape(-1), b.grad.reshape(-1)])
    for _, flat, wg in got:
        flat, wg = torch.from_numpy(flat), torch.from_numpy(wg)
        assert torch.allclose(flat, want, rtol=1e-6, atol=1e-6)
        assert torch.allclose(wg, w.grad, rtol=1e-6, atol=1e-6)  # the parameter's .grad is a view of the reduced buffer


def test_plan_cache_keys_hold_their_tensors_so_recycled_addresses_cannot_alias():
    """A fresh batch's ``room_mask`` / ``edge_index`` is ``_version`` 0 and can land in the allocator block a freed one
    just left: ``(data_ptr, _version)`` alone is not an identity.  The caches keep the keyed tensor alive and compare
    against it."""
    from hydra_gnn_b200.nn import PlanCache, same_tensor

    pc = PlanCache()
    a = torch.tensor([True, False, True, False])
    b = torch.tensor([False, True, True, True])
    for k in range(20):  # alternating fresh clones: temporaries are freed and their blocks reused
        m = (a if k % 2 == 0 else b).clone()
        rows = pc.mask_rows("room_rows", m)
        assert rows.tolist() == m.nonzero().view(-1).tolist(), k
        del m
    m = a.clone()
    r1 = pc.mask_rows("room_rows", m)
    assert pc.mask_rows("room_rows", m) is r1            # same tensor: cached
    assert pc.mask_rows("room_rows", m[:]) is r1         # a full view of it too
    m[1] = True                                          # in-place edit bumps the version
    assert pc.mask_rows("room_rows", m).tolist() == [0, 1, 2]
    e = torch.zeros(2, 3, dtype=torch.long)
    assert same_tensor(e, e) and same_tensor(e, e.view(2, 3)) and not same_tensor(e, e.clone())


# ------------------------------------------------------------------------------ device generator (host-side checks)
def test_synth_entry_points_reject_bad_arguments_on_the_host():
    import ctypes as C

    h = _lib.load_shared_object()
    EINVAL = -1
    assert h.hg_synth_mp3d_counts(None, None) == EINVAL and h.hg_synth_mp3d_fill(None, None) == EINVAL
    a = _lib.SynthArgs()
    a.n_graphs = -1
    assert h.hg_synth_mp3d_counts(C.byref(a), None) == EINVAL
    a.n_graphs, a.graph_id0 = 4, -3
    assert h.hg_synth_mp3d_counts(C.byref(a), None) == EINVAL
    a.graph_id0 = 0  # tables / prefix missing
    assert h.hg_synth_mp3d_counts(C.byref(a), None) == EINVAL and h.hg_synth_mp3d_fill(C.byref(a), None) == EINVAL


def test_generator_enumeration_order_restated_per_graph():
    """The device generator (csrc/synth.cu) walks one graph at a time: rooms, then per room the object pairs in (a, b)
    order, both directions next to each other.  The same walk written as scalar Python over the same counter hash must
    reproduce the vectorised host generator -- this pins the ORDER the kernel relies on without a GPU."""
    import math

    M = (1 << 64) - 1

    def mix(z):
        z = (z + 0x9E3779B97F4A7C15) & M
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
        return z ^ (z >> 31)

    def u01(k, stream, idx):
        k = mix(k ^ ((stream * 0xD1B54A32D192ED03) & M))
        return float(mix(k ^ mix(idx)) >> 11) * (1.0 / 9007199254740992.0)

    cr, co = synth._poisson_cdf(5.0, 64), synth._poisson_cdf(12.0, 128)
    first, n = 4242, 48
    host = synth.synth_mp3d_batch(range(first, first + n))
    rr, oo, o2r, labels = [[], []], [[], []], [], []
    room0 = obj0 = 0
    for g in range(n):
        k = mix(synth.SEED ^ mix(first + g))
        R = min(max(int(np.searchsorted(cr, u01(k, 1, 0), side="right")), 1), synth.MAX_ROOMS)
        nobj = [min(max(int(np.searchsorted(co, u01(k, 2, r), side="right")), 1), synth.MAX_OBJ) for r in range(R)]
        for a in range(R):
            for b in range(a + 1, R):
                tree = u01(k, 4, b) < 0.8 and a == int(math.floor(u01(k, 3, b) * b))
                if tree or u01(k, 5, a * synth.MAX_ROOMS + b) < 0.05:
                    rr[0] += [room0 + a, room0 + b]
                    rr[1] += [room0 + b, room0 + a]
        s = 0
        for r in range(R):
            p = min(1.0, 4.0 / max(nobj[r] - 1, 1))
            for a in range(nobj[r]):
                for b in range(a + 1, nobj[r]):
                    if u01(k, 6, (r * synth.MAX_OBJ + a) * synth.MAX_OBJ + b) < p:
                        oo[0] += [obj0 + s + a, obj0 + s + b]
                        oo[1] += [obj0 + s + b, obj0 + s + a]
            o2r += [room0 + r] * nobj[r]
            labels += [int(math.floor(u01(k, 11, s + o) * synth.N_OBJECT_LABELS)) for o in range(nobj[r])]
            s += nobj[r]
        room0, obj0 = room0 + R, obj0 + s
    assert torch.equal(torch.tensor(rr), host["rooms", "rooms_to_rooms", "rooms"].edge_index)
    assert torch.equal(torch.tensor(oo), host["objects", "objects_to_objects", "objects"].edge_index)
    assert torch.equal(torch.tensor(o2r), host["objects", "objects_to_rooms", "rooms"].edge_index[1])
    assert torch.equal(torch.tensor(labels), host["objects"].y)


# ------------------------------------------------------------------------------ shape heuristics (host-known sizes only)
def test_gat_mode_choice_from_shapes():
    """nn._gat_modes: aggregate-first only for far fewer destinations than sources; wide raw rows (306 vs 192) only for
    batches, even widths and <= 4 heads; never under a HeteroConv mean group."""
    from types import SimpleNamespace as NS

    o2r, o2o = ("objects", "objects_to_rooms", "rooms"), ("objects", "objects_to_objects", "objects")
    conv = NS(heads=3, out_channels=64)
    convs = {o2r: conv, o2o: conv}

    def modes(n_obj, n_room, width, group="sum", c=conv):
        x = {"objects": torch.empty(n_obj, width), "rooms": torch.empty(n_room, 6)}
        return hnn._gat_modes({o2r: c, o2o: c}, [o2r, o2o], x, group)

    assert modes(242968, 20277, 306) == (True, False)      # the bench's input layer: wide rows, a batch
    assert modes(242968, 20277, 64) == (True, False)       # hidden layers: raw rows narrower than the projection
    assert modes(59, 5, 306) == (False, False)             # batch 1 (serving): the wide form costs latency
    assert modes(59, 5, 64) == (True, False)
    assert modes(242968, 20277, 303) == (False, False)     # odd width (GAT_edge input): float2 gathers impossible
    assert modes(242968, 100000, 64) == (False, False)     # not 4x fewer destinations
    assert modes(242968, 20277, 306, group="mean") == (False, False)
    assert modes(242968, 20277, 306, c=NS(heads=6, out_channels=64)) == (True, False)  # 306 <= 384: narrower than the projection
    assert modes(242968, 20277, 400) == (False, False)     # wider than 1.6 x 192 and than 320
    assert convs  # (silence linters: both relations share the conv stand-in)


def test_window_hint_is_given_only_for_short_windows():
    """ops.window_applies: every relation has a window, <= 9 edges per destination row, <= 1.5 source rows per destination."""
    from types import SimpleNamespace as NS

    def rel(n_src, cap, win=True):
        return NS(csr=NS(tile_win=object() if win else None, cap=cap, n_src=n_src))

    n_o, n_r = 242968, 20277
    assert ops.window_applies([rel(n_o, 971518 + n_o), rel(n_r, n_o)], n_o)            # objects <- objects (+ loops), <- rooms
    assert ops.window_applies([rel(n_r, 29294 + n_r)], n_r)                          # rooms <- rooms
    assert not ops.window_applies([rel(n_o, n_o)], n_r)                              # rooms <- objects: 12 source rows per row
    assert not ops.window_applies([rel(n_o, 971518 + n_o), rel(n_r, n_o, win=False)], n_o)
    assert not ops.window_applies([rel(n_o, 10 * n_o)], n_o)                         # too many edges per row for the pool
    assert not ops.window_applies([], 0)
