"""GPU parity of the individual C-ABI entry points (through hydra_gnn_b200.ops) against the
oracle / exact references: K1 bit-exact, K2/K5, K6 (fp32 and tcgen05 TF32x3)."""
import numpy as np
import pytest
import torch

from hydra_gnn_b200 import ops, synth
from hydra_gnn_b200.ops import CSRSpec
from oracle.pyg import utils as ou
from util import assert_close

pytestmark = pytest.mark.gpu


# ------------------------------------------------------------------------------ K1 (integer, bit-exact)
def csr_oracle(ei, n_src, n_dst, self_loops):
    """Stable sort by destination of PyG's (remove_self_loops -> add_self_loops) edge list."""
    ei = ei.cpu()
    E = ei.size(1)
    ids = torch.arange(E)
    if self_loops:
        keep = ei[0] != ei[1]
        n = min(n_src, n_dst)
        loops = torch.arange(n)
        src = torch.cat([ei[0][keep], loops])
        dst = torch.cat([ei[1][keep], loops])
        ids = torch.cat([ids[keep], torch.full((n,), -1, dtype=torch.long)])
    else:
        src, dst = ei[0], ei[1]
    order = torch.sort(dst, stable=True).indices
    rowptr = torch.zeros(n_dst + 1, dtype=torch.long)
    rowptr[1:] = torch.bincount(dst, minlength=n_dst).cumsum(0)
    return rowptr.int(), src[order].int(), ids[order].int()


@pytest.mark.parametrize("n_src,n_dst,E,loops", [(50, 50, 400, True), (50, 50, 400, False), (1000, 37, 5000, False),
                                                 (37, 1000, 5000, True), (7, 7, 0, True), (0, 5, 0, False),
                                                 (20000, 20000, 300000, True), (3, 3, 5000, True)])
def test_plan_bit_exact(n_src, n_dst, E, loops):
    g = torch.Generator().manual_seed(E + n_src)
    ei = torch.stack([torch.randint(0, max(n_src, 1), (E,), generator=g),
                      torch.randint(0, max(n_dst, 1), (E,), generator=g)]) if E else torch.zeros(2, 0, dtype=torch.long)
    (csr,), status = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, loops, want_perm=True, want_norm=True)])
    rowptr, col, perm = csr_oracle(ei, n_src, n_dst, loops)
    n = int(rowptr[-1])
    assert int(status.item()) == 0
    assert torch.equal(csr.rowptr.cpu(), rowptr)
    assert torch.equal(csr.col.cpu()[:n], col)
    assert torch.equal(csr.perm.cpu()[:n], perm)
    deg = (rowptr[1:] - rowptr[:-1]).float()
    want = torch.where(deg > 0, deg.pow(-0.5), torch.zeros_like(deg))
    assert torch.allclose(csr.inv_sqrt_deg.cpu()[:n_dst], want, rtol=1e-6, atol=0)


def test_plan_many_relations_and_transpose():
    g = torch.Generator().manual_seed(1)
    specs, eis = [], []
    for r in range(19):  # > HG_MAX_PLAN_RELATIONS: exercises chunking
        ns, nd, E = 10 + 13 * r, 200 - 7 * r, 50 * r
        ei = torch.stack([torch.randint(0, ns, (E,), generator=g), torch.randint(0, nd, (E,), generator=g)])
        eis.append((ei, ns, nd))
        specs.append(CSRSpec(ei.cuda(), ns, nd, False, want_perm=True, transpose=(r % 2 == 1)))
    csrs, status = ops.build_csr(specs)
    assert int(status.item()) == 0
    for r, (csr, (ei, ns, nd)) in enumerate(zip(csrs, eis)):
        if r % 2 == 1:
            ei, ns, nd = ei.flip(0), nd, ns
        rowptr, col, perm = csr_oracle(ei, ns, nd, False)
        n = int(rowptr[-1])
        assert torch.equal(csr.rowptr.cpu(), rowptr) and torch.equal(csr.col.cpu()[:n], col)
        assert torch.equal(csr.perm.cpu()[:n], perm)


def test_plan_out_of_range_edges_are_skipped_and_counted():
    ei = torch.tensor([[0, 1, 99, 2, -1], [1, 2, 0, 77, 0]])
    (csr,), status = ops.build_csr([CSRSpec(ei.cuda(), 3, 3, False, want_perm=True)])
    assert int(status.item()) == 3
    assert csr.rowptr.cpu().tolist() == [0, 0, 1, 2]
    assert csr.col.cpu()[:2].tolist() == [0, 1]


# ------------------------------------------------------------------------------ K2 / K5
@pytest.mark.parametrize("F", [1, 3, 6, 26, 64, 78, 128, 306, 768, 1100])
@pytest.mark.parametrize("op", ["sum", "mean", "max"])
def test_seg_reduce(F, op):
    g = torch.Generator().manual_seed(F)
    n_src, n_dst, E = 301, 257, 2000
    ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), torch.randint(0, n_dst - 40, (E,), generator=g)])
    x = torch.randn(n_src, F, generator=g)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst)])
    got = ops.seg_reduce(x.cuda(), csr, op)
    want = ou.scatter(x.double()[ei[0]], ei[1], 0, dim_size=n_dst, reduce=op)
    assert_close(got, want, f"seg_reduce {op} F={F}", check_argmax=False)


def test_seg_reduce_epilogue_and_row_map():
    g = torch.Generator().manual_seed(3)
    n, E, F = 500, 3000, 96
    ei = torch.stack([torch.randint(0, n, (E,), generator=g), torch.randint(0, n, (E,), generator=g)])
    x, xs = torch.randn(n, F, generator=g), torch.randn(n, F, generator=g)
    ss, ds = torch.rand(n, generator=g), torch.rand(n, generator=g)
    bias, eps = torch.randn(F, generator=g), torch.tensor([0.3])
    rows = torch.randperm(n, generator=g)[:77].int()
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n, n)])
    got = ops.seg_reduce(x.cuda(), csr, "sum", src_scale=ss.cuda(), dst_scale=ds.cuda(), row_map=rows.cuda(),
                         x_self=xs.cuda(), self_coef=eps.cuda(), self_coef_add=1.0, bias=bias.cuda(), act="relu")
    agg = ou.scatter((ss[ei[0]].unsqueeze(1) * x[ei[0]]).double(), ei[1], 0, dim_size=n, reduce="sum")
    want = torch.relu(ds.unsqueeze(1).double() * agg + 1.3 * xs.double() + bias.double())[rows.long()]
    assert_close(got, want, "seg_reduce epilogue", check_argmax=False)


def test_seg_reduce_strided_views():
    g = torch.Generator().manual_seed(4)
    n, E = 300, 1500
    ei = torch.stack([torch.randint(0, n, (E,), generator=g), torch.randint(0, n, (E,), generator=g)])
    Y = torch.randn(n, 256, generator=g)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n, n)])
    Yc = Y.cuda()
    got = ops.seg_reduce(Yc[:, :128], csr, "mean", x_self=Yc[:, 128:], self_coef_add=1.0)
    want = ou.scatter(Y[:, :128].double()[ei[0]], ei[1], 0, dim_size=n, reduce="mean") + Y[:, 128:].double()
    assert_close(got, want, "strided", check_argmax=False)


# ------------------------------------------------------------------------------ K6
SHAPES = [(1000, 396, [306]), (777, 84, [64]), (129, 16, [6]), (5000, 396, [64]), (2048, 128, [128, 128]),
          (600, 35, [6, 6, 6]), (513, 300, [306]), (4096, 768, [768]),
          (40001, 204, [306]), (38000, 396, [64])]  # the last two: >= 2 m-tiles per SM -> 2-CTA clusters sharing W


@pytest.mark.parametrize("precision", ["fp32", "tf32x3"])
@pytest.mark.parametrize("M,N,Ks", SHAPES)
def test_linear(M, N, Ks, precision):
    g = torch.Generator().manual_seed(M + N)
    segs = [(torch.randn(M, K, generator=g), torch.randn(N, K, generator=g) / K ** 0.5) for K in Ks]
    bias = torch.randn(N, generator=g)
    got = ops.linear([(x.cuda(), w.cuda()) for x, w in segs], N, bias=bias.cuda(), precision=precision)
    want = sum(x.double() @ w.double().t() for x, w in segs) + bias.double()
    torch.cuda.synchronize()
    # tf32x3: the tensor core truncates on accumulate (~1 ulp of the accumulator per K=8 MMA)
    tol = 2e-5 if precision == "fp32" else 2e-5 + 2e-8 * sum(Ks)
    assert_close(got, want, f"linear {precision} M={M} N={N} K={Ks}", rtol=tol, atol=tol, check_argmax=False)
    err = (got.double().cpu() - want).abs().max().item()
    print(f"linear {precision} M={M} N={N} K={Ks}: max abs err {err:.2e}")


@pytest.mark.parametrize("precision", ["fp32", "tf32x3"])
def test_linear_epilogue_strided(precision):
    g = torch.Generator().manual_seed(9)
    M, N, K = 700, 100, 306
    xfull = torch.randn(M, 400, generator=g)
    x = xfull[:, 3:3 + K]  # odd offset: rows only 4-byte aligned -> scalar load path
    w, bias = torch.randn(N, K, generator=g) / K ** 0.5, torch.randn(N, generator=g)
    scale, shift = torch.rand(N, generator=g) + 0.5, torch.randn(N, generator=g)
    out = torch.zeros(M, 128, device="cuda")
    ops.linear([(xfull.cuda()[:, 3:3 + K], w.cuda())], N, bias=bias.cuda(), scale=scale.cuda(), shift=shift.cuda(),
               act="elu", precision=precision, out=out[:, 4:4 + N])
    want = torch.nn.functional.elu((x.double() @ w.double().t() + bias.double()) * scale.double() + shift.double())
    assert_close(out[:, 4:4 + N], want, "linear epilogue", rtol=2e-5, atol=2e-5, check_argmax=False)
    assert float(out[:, :4].abs().max()) == 0 and float(out[:, 4 + N:].abs().max()) == 0


def test_gather_rows():
    x = torch.randn(100, 37)
    idx = torch.tensor([5, 99, 0, 5], dtype=torch.int32)
    got = ops.gather_rows(x.cuda(), idx.cuda(), act="relu")
    assert torch.equal(got.cpu(), torch.relu(x[idx.long()]))


# ------------------------------------------------------------------------------ K3 + K4
def gat_reference(xs, a_src, a_dst, ei, n_dst, H, C, concat, src_shared=False, slope=0.2):
    """fp64 gather -> leaky_relu -> segment softmax (+1e-16) -> weighted scatter (PyG GATConv, section 9.2)."""
    j, i = ei[0], ei[1]
    s = torch.nn.functional.leaky_relu(a_src.double()[j] + a_dst.double()[i], slope)  # [E, H]
    w = ou.softmax(s, i, None, n_dst)
    src = xs.double()[j]
    msg = w.unsqueeze(-1) * (src.unsqueeze(1) if src_shared else src.view(-1, H, C))
    out = ou.scatter(msg, i, 0, dim_size=n_dst, reduce="sum")
    return (out.reshape(n_dst, H * C) if (concat or src_shared) else out.mean(1)), w


GAT_SHAPES = [(3, 64, False), (3, 26, False), (6, 64, True), (1, 6, False), (1, 306, False), (5, 35, True),
              (8, 16, True), (2, 640, True), (3, 7, False)]


@pytest.mark.parametrize("H,C,concat", GAT_SHAPES)
def test_gat_fused_two_relations(H, C, concat):
    g = torch.Generator().manual_seed(H * 1000 + C)
    n_dst, n_a, n_b = 333, 333, 57
    rels, want, biases = [], 0, []
    for n_src, E in ((n_a, 1700), (n_b, 400)):
        ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), torch.randint(0, n_dst - 20, (E,), generator=g)])
        xs = torch.randn(n_src, H * C + 8, generator=g)  # strided view: leading dimension != width
        a_s, a_d = torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
        bias = torch.randn(H * C if concat else C, generator=g)
        (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, want_perm=True)])
        alpha = torch.zeros(E, H, device="cuda")
        rels.append(ops.GatRelation(xs.cuda()[:, :H * C], a_s.cuda(), a_d.cuda(), csr, bias.cuda(), None, alpha))
        o, w = gat_reference(xs[:, :H * C], a_s, a_d, ei, n_dst, H, C, concat)
        want = want + o + bias.double()
        biases.append((csr, w, alpha))
    addend = torch.randn(n_dst, H * C if concat else C, generator=g)
    got = ops.gat_fused(rels, n_dst, H, C, concat, act="elu", addend=addend.cuda())
    assert_close(got, torch.nn.functional.elu(want + addend.double()), f"gat H={H} C={C}", check_argmax=False)
    for csr, w, alpha in biases:  # softmax weights saved in CSR order
        perm = csr.perm.cpu().long()
        assert_close(alpha.cpu(), w[perm].float(), "alpha", check_argmax=False)


def test_gat_fused_group_mean_and_aggregate_first():
    g = torch.Generator().manual_seed(11)
    n_src, n_dst, E, H, C = 400, 90, 1500, 3, 64
    ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), torch.randint(0, n_dst, (E,), generator=g)])
    x = torch.randn(n_src, C, generator=g)
    a_s, a_d = torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst)])
    rel = ops.GatRelation(x.cuda(), a_s.cuda(), a_d.cuda(), csr)
    got = ops.gat_fused([rel], n_dst, H, C, True, src_shared=True)
    want, _ = gat_reference(x, a_s, a_d, ei, n_dst, H, C, True, src_shared=True)
    assert_close(got, want, "aggregate-first", check_argmax=False)
    xs = torch.randn(n_src, H * C, generator=g)
    rel2 = ops.GatRelation(xs.cuda(), a_s.cuda(), a_d.cuda(), csr, torch.ones(C).cuda())
    got = ops.gat_fused([rel2, rel2], n_dst, H, C, False, group="mean", act="relu")
    want, _ = gat_reference(xs, a_s, a_d, ei, n_dst, H, C, False)
    assert_close(got, torch.relu(want + 1.0), "group mean", check_argmax=False)


@pytest.mark.parametrize("H,C,pad", [(3, 306, 0), (3, 306, 2), (1, 306, 0), (4, 256, 0), (2, 130, 0), (3, 384, 4)])
def test_gat_aggregate_first_wide_rows(H, C, pad):
    """Aggregate-first over wide source rows (objects -> rooms at the 306-d input layer; float2 / float4 chunks, up to 5
    per lane): rows longer than a tile's share, empty rows, strided sources, a bias, edge logits."""
    g = torch.Generator().manual_seed(7 * H + C)
    n_src, n_dst = 3000, 257
    deg = torch.randint(0, 25, (n_dst,), generator=g)
    deg[5], deg[6], deg[100] = 0, 77, 33
    dst = torch.repeat_interleave(torch.arange(n_dst), deg)
    E = int(dst.numel())
    ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), dst])[:, torch.randperm(E, generator=g)]
    xb = torch.randn(n_src, C + pad, generator=g)
    x = xb[:, :C]
    a_s, a_d = torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, want_perm=True)])
    want, _ = gat_reference(x, a_s, a_d, ei, n_dst, H, C, True, src_shared=True)
    rel = ops.GatRelation(xb.cuda()[:, :C], a_s.cuda(), a_d.cuda(), csr)
    got = ops.gat_fused([rel], n_dst, H, C, True, src_shared=True)
    assert_close(got, want, f"aggregate-first H={H} C={C}", check_argmax=False)
    assert float(got[5].abs().max()) == 0.0  # a row without edges
    bias = torch.randn(H * C, generator=g)
    rel_b = ops.GatRelation(xb.cuda()[:, :C], a_s.cuda(), a_d.cuda(), csr, bias.cuda())
    assert_close(ops.gat_fused([rel_b], n_dst, H, C, True, src_shared=True), want + bias.double(), "with bias", check_argmax=False)
    # GAT_edge: per-edge logits in CSR order
    a_e = torch.randn(csr.cap, H, generator=g)
    perm = csr.perm.cpu().long()[:E]
    j, i = ei[0][perm], ei[1][perm]
    sc = torch.nn.functional.leaky_relu(a_s.double()[j] + a_d.double()[i] + a_e.double()[:E], 0.2)
    w = ou.softmax(sc, i, None, n_dst)
    want_e = ou.scatter(w.unsqueeze(-1) * x.double()[j].unsqueeze(1), i, 0, dim_size=n_dst, reduce="sum").reshape(n_dst, H * C)
    rel_e = ops.GatRelation(xb.cuda()[:, :C], a_s.cuda(), a_d.cuda(), csr, None, a_e.cuda())
    assert_close(ops.gat_fused([rel_e], n_dst, H, C, True, src_shared=True), want_e, "aggregate-first + edge logits",
                 check_argmax=False)


def test_gat_fused_hub_rows_take_the_streaming_path():
    g = torch.Generator().manual_seed(12)
    n_src, n_dst, H, C = 5000, 70, 3, 64
    hub = torch.stack([torch.randint(0, n_src, (6000,), generator=g), torch.full((6000,), 3)])  # one row of degree 6000
    rest = torch.stack([torch.randint(0, n_src, (500,), generator=g), torch.randint(0, n_dst, (500,), generator=g)])
    ei = torch.cat([hub, rest], dim=1)
    xs = torch.randn(n_src, H * C, generator=g)
    a_s, a_d = torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst)])
    for concat in (False, True):
        got = ops.gat_fused([ops.GatRelation(xs.cuda(), a_s.cuda(), a_d.cuda(), csr)], n_dst, H, C, concat)
        want, _ = gat_reference(xs, a_s, a_d, ei, n_dst, H, C, concat)
        assert_close(got, want, f"hub concat={concat}", check_argmax=False)


# ------------------------------------------------------------------------------ tile windows + the shared-memory window kernel
def _tile_windows(rowptr, col, n_dst):
    """{first source row, rows spanned} per tile of 32 destination rows, from the CSR (integer, exact)."""
    rowptr, col = rowptr.cpu().long(), col.cpu().long()
    out = torch.zeros((n_dst + 31) // 32, 2, dtype=torch.int32)
    for t in range(out.size(0)):
        c = col[rowptr[32 * t]: rowptr[min(32 * t + 32, n_dst)]]
        if c.numel():
            out[t, 0], out[t, 1] = int(c.min()), int(c.max() - c.min() + 1)
    return out


@pytest.mark.parametrize("n_src,n_dst,E,loops", [(50, 50, 400, True), (1000, 37, 5000, False), (37, 1000, 5000, True),
                                                 (7, 7, 0, True), (0, 5, 0, False), (3000, 2999, 9000, False)])
def test_plan_tile_windows_bit_exact(n_src, n_dst, E, loops):
    g = torch.Generator().manual_seed(E + n_dst)
    ei = torch.stack([torch.randint(0, max(n_src, 1), (E,), generator=g),
                      torch.randint(0, max(n_dst, 1), (E,), generator=g)]) if E else torch.zeros(2, 0, dtype=torch.long)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, loops, want_window=True)])
    assert torch.equal(csr.tile_win.cpu(), _tile_windows(csr.rowptr, csr.col, n_dst))


@pytest.mark.parametrize("loops", [True, False])
def test_plan_single_pass_scan_many_tiles(loops):
    """147 scan tiles of 2048 rows chained by look-back, two relations of different size in one call, windows + perm +
    degree^-1/2; one relation's edges grouped by destination block (the batch layout), the other's random."""
    g = torch.Generator().manual_seed(5)
    n, E = 300001, 1500000
    blk = torch.randint(0, n // 50, (E,), generator=g).sort().values * 50
    local = torch.stack([blk + torch.randint(0, 50, (E,), generator=g), blk + torch.randint(0, 50, (E,), generator=g)])
    rand = torch.stack([torch.randint(0, 70000, (E // 3,), generator=g), torch.randint(0, 4097, (E // 3,), generator=g)])
    specs = [CSRSpec(local.cuda(), n, n, loops, want_perm=True, want_norm=True, want_window=True),
             CSRSpec(rand.cuda(), 70000, 4097, False, want_perm=True, want_window=True)]
    for _ in range(2):  # the second call reuses a warm workspace
        csrs, status = ops.build_csr(specs)
        assert int(status.item()) == 0
        for csr, (ei, ns, nd, lp) in zip(csrs, [(local, n, n, loops), (rand, 70000, 4097, False)]):
            rowptr, col, perm = csr_oracle(ei, ns, nd, lp)
            m = int(rowptr[-1])
            assert torch.equal(csr.rowptr.cpu(), rowptr)
            assert torch.equal(csr.col.cpu()[:m], col) and torch.equal(csr.perm.cpu()[:m], perm)
            assert torch.equal(csr.tile_win.cpu(), _tile_windows(csr.rowptr, csr.col, nd))
    deg = (csrs[0].rowptr[1:] - csrs[0].rowptr[:-1]).float()
    assert torch.equal(csrs[0].inv_sqrt_deg[:n], torch.where(deg > 0, deg.pow(-0.5), torch.zeros_like(deg))) or \
        torch.allclose(csrs[0].inv_sqrt_deg[:n], torch.where(deg > 0, deg.pow(-0.5), torch.zeros_like(deg)), rtol=1e-6, atol=0)


def _scene_batch_relations(graph_ids, H, g, a_edge=False):
    """o->o (self loops) and r->o relations of a synthetic scene-graph batch with random projected rows / logits."""
    b = synth.synth_mp3d_batch(graph_ids, device="cuda")
    n_o, n_r = b["objects"].x.size(0), b["rooms"].x.size(0)
    specs = [CSRSpec(b[("objects", "objects_to_objects", "objects")].edge_index, n_o, n_o, True, want_perm=True, want_window=True),
             CSRSpec(b[("rooms", "rooms_to_objects", "objects")].edge_index, n_r, n_o, False, want_perm=True, want_window=True)]
    csrs, _ = ops.build_csr(specs)
    rels = []
    for csr, n_src in zip(csrs, (n_o, n_r)):
        Y = torch.randn(n_src, H * 64 + 12, generator=g).cuda()  # xs is a column slice of a wider projection output
        ae = torch.randn(csr.cap, H, generator=g).cuda() if a_edge else None
        rels.append(ops.GatRelation(Y[:, :H * 64], Y[:, H * 64: H * 64 + H], torch.randn(n_o, H, generator=g).cuda(), csr,
                                    torch.randn(64, generator=g).cuda(), ae))
    return rels, n_o


def _fp64_gat(rels, n_dst, H, C):
    """fp64 head-mean GAT over CSR relations (rows in CSR order), biases included."""
    want = 0
    for r in rels:
        rowptr = r.csr.rowptr.cpu().long()
        n_live = int(rowptr[-1])
        col = r.csr.col.cpu().long()[:n_live]
        dst = torch.repeat_interleave(torch.arange(n_dst), rowptr[1:] - rowptr[:-1])
        a_src, a_dst = r.a_src.cpu(), r.a_dst.cpu()
        if r.a_edge is not None:  # fold the edge term into a per-edge source logit
            s = a_src.double()[col] + r.a_edge.cpu().double()[:n_live] + a_dst.double()[dst]
            s = torch.nn.functional.leaky_relu(s, 0.2)
            w = ou.softmax(s, dst, None, n_dst)
            msg = w.unsqueeze(-1) * r.xs.cpu().double()[col].view(-1, H, C)
            o = ou.scatter(msg, dst, 0, dim_size=n_dst, reduce="sum").mean(1)
        else:
            o, _ = gat_reference(r.xs.cpu(), a_src, a_dst, torch.stack([col, dst]), n_dst, H, C, False)
        want = want + o + r.bias.cpu().double()
    return want


@pytest.mark.parametrize("H", [1, 2, 3, 4])
@pytest.mark.parametrize("n_graphs,a_edge", [(37, False), (300, False), (64, True), (1200, False), (1150, True)])
def test_gat_window_kernel_matches_fp64_and_the_global_gather_kernel_bitwise(H, n_graphs, a_edge):
    g = torch.Generator().manual_seed(100 * H + n_graphs)
    rels, n_o = _scene_batch_relations(range(n_graphs), H, g, a_edge)
    addend = torch.randn(n_o, 64, generator=g).cuda()
    assert ops.window_applies(rels, n_o)
    got = ops.gat_fused(rels, n_o, H, 64, False, act="elu", addend=addend)
    assert_close(got, torch.nn.functional.elu(_fp64_gat(rels, n_o, H, 64) + addend.cpu().double()), f"window H={H}",
                 check_argmax=False)
    # <= 65,536 rows: the small-destination row kernel; more: the window kernel (n_graphs >= 1150).  Either must equal
    # the v6 tile kernel bit for bit (same accumulation order, same rounding); asking for the saved softmax weights
    # forces v6, and so does dropping the window hint above 65,536 rows
    for r in rels:
        r.alpha_out = torch.empty(r.csr.cap, H, device="cuda")
    base = ops.gat_fused(rels, n_o, H, 64, False, act="elu", addend=addend)
    for r in rels:
        r.alpha_out = None
    assert torch.equal(got, base)
    try:
        ops.USE_TILE_WINDOW = False
        assert torch.equal(ops.gat_fused(rels, n_o, H, 64, False, act="elu", addend=addend), base)
    finally:
        ops.USE_TILE_WINDOW = True
    # one relation only, group mean over a repeated relation, ReLU
    got1 = ops.gat_fused([rels[0], rels[0], rels[1]], n_o, H, 64, False, group="mean", act="relu")
    want1 = torch.relu((2 * _fp64_gat(rels[:1], n_o, H, 64) + _fp64_gat(rels[1:], n_o, H, 64)) / 3)
    assert_close(got1, want1, "window, 3 relations, group mean", check_argmax=False)


@pytest.mark.parametrize("H,C", [(3, 26), (3, 64), (1, 32), (4, 12), (2, 60)])
def test_gat_row_kernel_small_destination_sets(H, C):
    """<= 65,536 destination rows, head-mean, C <= 64: the row kernel (8 threads own a row; the output layer's C = 26 takes
    float2 chunks).  Against fp64, and bit for bit against the tile kernel (asking for the saved weights forces it): two
    relations, bias, addend, ELU, an empty row and a 100-edge row (warp-per-row path)."""
    g = torch.Generator().manual_seed(31 * H + C)
    n_dst, rels, want = 3001, [], 0
    for n_src, E in ((3001, 9000), (411, 3001)):
        dst = torch.randint(0, n_dst, (E,), generator=g)
        dst[dst == 7] = 8          # row 7 has no edges
        dst[:100] = 1500           # row 1500: 100 edges from this relation
        ei = torch.stack([torch.randint(0, n_src, (E,), generator=g), dst])
        xs = torch.randn(n_src, H * C + 4, generator=g)
        a_s, a_d = torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
        bias = torch.randn(C, generator=g)
        (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, want_perm=True)])
        rels.append(ops.GatRelation(xs.cuda()[:, :H * C], a_s.cuda(), a_d.cuda(), csr, bias.cuda()))
        o, _ = gat_reference(xs[:, :H * C], a_s, a_d, ei, n_dst, H, C, False)
        want = want + o + bias.double()
    addend = torch.randn(n_dst, C, generator=g)
    got = ops.gat_fused(rels, n_dst, H, C, False, act="elu", addend=addend.cuda())
    assert_close(got, torch.nn.functional.elu(want + addend.double()), f"row kernel H={H} C={C}", check_argmax=False)
    for r in rels:
        r.alpha_out = torch.empty(r.csr.cap, H, device="cuda")
    tile = ops.gat_fused(rels, n_dst, H, C, False, act="elu", addend=addend.cuda())
    rows_1500 = torch.zeros(n_dst, dtype=torch.bool)
    rows_1500[1500 // 4 * 4: 1500 // 4 * 4 + 4] = True  # the warp holding the long row takes the warp-per-row path
    assert torch.equal(got.cpu()[~rows_1500], tile.cpu()[~rows_1500])
    assert_close(got.cpu()[rows_1500], tile.cpu()[rows_1500], "long row", check_argmax=False)


def test_gat_window_kernel_any_graph_and_any_hint():
    """Sources outside the staged window are gathered from global memory: random graphs (windows span everything),
    windows that do not cover the tile's sources at all, and hub tiles (edge pool overflow) all give the same bits."""
    g = torch.Generator().manual_seed(5)
    H, C, n_src, n_dst = 3, 64, 100000, 70000 + 17
    ei = torch.stack([torch.randint(0, n_src, (400000,), generator=g), torch.randint(0, n_dst, (400000,), generator=g)])
    hub = torch.stack([torch.randint(0, n_src, (700,), generator=g), torch.full((700,), 40)])  # tile 1: 700 edges > pool
    ei = torch.cat([ei, hub], dim=1)
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), n_src, n_dst, False, want_window=True)])
    xs, a_s, a_d = torch.randn(n_src, H * C, generator=g), torch.randn(n_src, H, generator=g), torch.randn(n_dst, H, generator=g)
    rel = ops.GatRelation(xs.cuda(), a_s.cuda(), a_d.cuda(), csr)
    want, _ = gat_reference(xs, a_s, a_d, ei, n_dst, H, C, False)
    assert ops.window_applies([rel], n_dst)
    got = ops.gat_fused([rel], n_dst, H, C, False)
    assert_close(got, want, "window kernel, random graph", check_argmax=False)
    win = csr.tile_win.clone()
    for hint in (torch.zeros_like(win), torch.stack([torch.full_like(win[:, 0], n_src - 5), torch.full_like(win[:, 1], 5)], 1),
                 torch.stack([win[:, 0] + 3, (win[:, 1] - 3).clamp(min=0)], 1)):
        csr.tile_win = hint.contiguous()
        assert torch.equal(ops.gat_fused([rel], n_dst, H, C, False), got)
    csr.tile_win = None
    base = ops.gat_fused([rel], n_dst, H, C, False)  # the global-gather kernel
    hub_rows = torch.zeros(n_dst, dtype=torch.bool)
    hub_rows[32:64] = True  # its staging pool holds this tile; the window kernel's does not (warp-per-row path: other rounding)
    assert torch.equal(base[~hub_rows], got[~hub_rows])
    assert_close(base[hub_rows], got[hub_rows], "hub tile", check_argmax=False)


# ------------------------------------------------------------------------------ GPU collate (bit-exact)
def _graph_list(n, hetero=True):
    from hydra_gnn_b200 import convert
    out = []
    for i in range(n):
        g = synth.synth_mp3d_batch([i], relative_pos=True)
        out.append(g if hetero else convert.dsg_to_homogeneous(g))
    return out


@pytest.mark.parametrize("hetero", [True, False])
def test_gpu_collate_is_bit_identical_to_cpu_collate(hetero):
    from hydra_gnn_b200.data import Batch
    graphs = _graph_list(37, hetero)
    want = Batch.from_data_list(graphs)
    got = Batch.from_data_list(graphs, device="cuda")
    torch.cuda.synchronize()
    stores_w = list(want._stores()) if hetero else [want]
    stores_g = list(got._stores()) if hetero else [got]
    assert len(stores_w) == len(stores_g)
    n = 0
    for sw, sg in zip(stores_w, stores_g):
        assert set(sw.keys()) == set(sg.keys())
        for k, v in sw.items():
            if torch.is_tensor(v):
                assert sg[k].is_cuda and sg[k].dtype == v.dtype and torch.equal(sg[k].cpu(), v), k
                n += 1
            else:
                assert sg[k] == v, k
    assert n >= 8


def test_collate_offsets_with_empty_graphs():
    # graphs 1 and 3 have no edges; graph 2 has no nodes of the source type
    edge_counts, src_counts, dst_counts = [3, 0, 2, 0, 4], [2, 5, 0, 1, 3], [4, 1, 2, 2, 1]
    def prefix(c):
        p = torch.zeros(len(c) + 1, dtype=torch.long)
        p[1:] = torch.tensor(c).cumsum(0)
        return p
    ep, sp, dp = prefix(edge_counts), prefix(src_counts), prefix(dst_counts)
    g = torch.Generator().manual_seed(3)
    idx = torch.randint(0, 5, (2, int(ep[-1])), generator=g)
    want = idx.clone()
    for k in range(len(edge_counts)):
        want[0, ep[k]:ep[k + 1]] += sp[k]
        want[1, ep[k]:ep[k + 1]] += dp[k]
    got = ops.collate_offsets_(idx.cuda(), ep.cuda(), sp.cuda(), dp.cuda())
    assert torch.equal(got.cpu(), want)
    bv = ops.batch_vector(sp.cuda(), int(sp[-1]))
    assert torch.equal(bv.cpu(), torch.repeat_interleave(torch.arange(len(src_counts)), torch.tensor(src_counts)))


def test_nvtx_ranges_around_abi_calls_do_not_change_results():
    g = torch.Generator().manual_seed(2)
    ei = torch.stack([torch.randint(0, 300, (2000,), generator=g), torch.randint(0, 200, (2000,), generator=g)])
    x = torch.randn(300, 64, generator=g).cuda()
    (csr,), _ = ops.build_csr([CSRSpec(ei.cuda(), 300, 200)])
    want = ops.seg_reduce(x, csr, "sum")
    ops.set_nvtx(True)
    try:
        (csr2,), _ = ops.build_csr([CSRSpec(ei.cuda(), 300, 200)])
        got = ops.seg_reduce(x, csr2, "sum")
    finally:
        ops.set_nvtx(False)
    assert torch.equal(got, want)
