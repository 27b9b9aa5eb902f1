"""hg_plan_build_static -- the allocation-free variant of hg_plan_build (scratch in a file-scope
device array inside the library; include/hydra_gnn_b200.h).  It has no counterpart in the reference.
Collected last on
purpose.

The GPU test calls both entry points through the C ABI (ctypes) on the same edge lists and requires
bit-identical integer outputs (the bar of tests/test_gpu_ops.py::test_plan_bit_exact) on two calls
in a row -- the second call sees the first call's leftovers in the shared array."""
import ctypes as C

import pytest
import torch

from hydra_gnn_b200 import _lib, synth

ECAPACITY = -7


def _plan_rels(edge_sets, device):
    """PlanRel array + output tensors for [(edge_index [2,E] int64, n_src, n_dst, self_loops)]."""
    rels = (_lib.PlanRel * len(edge_sets))()
    outs = []
    for k, (ei, n_src, n_dst, loops) in enumerate(edge_sets):
        E = ei.size(1)
        cap = E + (min(n_src, n_dst) if loops else 0)
        rowptr = torch.full((n_dst + 1,), -7, dtype=torch.int32, device=device)
        col = torch.full((max(cap, 1),), -7, dtype=torch.int32, device=device)
        perm = torch.full((max(cap, 1),), -7, dtype=torch.int32, device=device)
        norm = torch.full((max(n_dst, 1),), -7.0, dtype=torch.float32, device=device)
        r = rels[k]
        r.src, r.dst = ei.data_ptr(), ei.data_ptr() + E * 8
        r.n_edges, r.n_src, r.n_dst, r.self_loops = E, n_src, n_dst, int(loops)
        r.rowptr, r.col, r.perm, r.inv_sqrt_deg = rowptr.data_ptr(), col.data_ptr(), perm.data_ptr(), norm.data_ptr()
        outs.append((rowptr, col, perm, norm, cap))
    return rels, outs


def _edge_sets(graph_ids, device):
    b = synth.synth_mp3d_batch(graph_ids, device=device)
    n = {t: b[t].x.size(0) for t in b.node_types}
    sets = []
    for (s, rel, d) in b.edge_types:
        ei = b[(s, rel, d)].edge_index.contiguous()
        sets.append((ei, n[s], n[d], s == d))
    return sets


def test_static_plan_rejects_sizes_beyond_its_capacity_before_any_cuda_call():
    h = _lib.load_shared_object()
    assert b"capacity" in h.hg_strerror(ECAPACITY)
    r = (_lib.PlanRel * 1)()
    r[0].n_edges, r[0].n_src, r[0].n_dst = (_lib.HG_PLAN_STATIC_SCRATCH_BYTES // 4) + 1, 8, 8  # ids alone overflow
    assert h.hg_plan_workspace_bytes(r, 1) - 256 > _lib.HG_PLAN_STATIC_SCRATCH_BYTES
    assert h.hg_plan_build_static(r, 1, None, None) == ECAPACITY
    r[0].n_edges = -1
    assert h.hg_plan_build_static(r, 1, None, None) == -1  # same argument checks as hg_plan_build


@pytest.mark.gpu
def test_static_plan_matches_hg_plan_build_two_calls_in_a_row():
    h = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    # second batch is smaller and different: stale counters / ids of the first call lie beyond and under it
    for ids in (range(0, 512), range(700, 900)):
        sets = _edge_sets(ids, dev)
        rels_a, outs_a = _plan_rels(sets, dev)
        rels_b, outs_b = _plan_rels(sets, dev)
        nbytes = h.hg_plan_workspace_bytes(rels_a, len(sets))
        assert 0 < nbytes - 256 <= _lib.HG_PLAN_STATIC_SCRATCH_BYTES
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        st_a = torch.full((1,), 99, dtype=torch.int32, device=dev)
        st_b = torch.full((1,), 99, dtype=torch.int32, device=dev)
        _lib.check(h.hg_plan_build(rels_a, len(sets), st_a.data_ptr(), ws.data_ptr(), ws.numel(), stream), "hg_plan_build")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
        _lib.check(h.hg_plan_build_static(rels_b, len(sets), st_b.data_ptr(), stream), "hg_plan_build_static")
        torch.cuda.synchronize()
        assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before
        assert int(st_a) == 0 and int(st_b) == 0
        for (rp_a, col_a, perm_a, nrm_a, cap), (rp_b, col_b, perm_b, nrm_b, _) in zip(outs_a, outs_b):
            assert torch.equal(rp_a, rp_b)
            n_live = int(rp_a[-1])
            assert 0 < n_live <= cap
            assert torch.equal(col_a[:n_live], col_b[:n_live])
            assert torch.equal(perm_a[:n_live], perm_b[:n_live])
            assert torch.equal(nrm_a, nrm_b)  # deg^-1/2: same kernel, same bits


@pytest.mark.gpu
def test_static_plan_over_capacity_on_device_leaves_outputs_untouched():
    h = _lib.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    E = _lib.HG_PLAN_STATIC_SCRATCH_BYTES // 4 + 1024
    ei = torch.zeros(2, E, dtype=torch.int64, device=dev)
    rels, outs = _plan_rels([(ei, 16, 16, False)], dev)
    st = torch.full((1,), 99, dtype=torch.int32, device=dev)
    rc = h.hg_plan_build_static(rels, 1, st.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == ECAPACITY
    with pytest.raises(RuntimeError, match="capacity"):
        _lib.check(rc, "hg_plan_build_static")
    torch.cuda.synchronize()
    assert int(st) == 99 and int(outs[0][0][0]) == -7
