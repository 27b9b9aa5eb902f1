"""TEST INFRASTRUCTURE ONLY: plain-torch CPU stand-ins for the handful of ``hydra_gnn_b200.ops`` entry points the
training glue (``hydra_gnn_b200/train.py``) composes, so that the HOST logic of the autograd Functions -- what is
saved, which operand gets which gradient, how scales / degrees / row maps enter the backward -- can be checked on a
machine without a GPU against the oracle's autograd.  The stand-ins restate the documented contract of each
operator (ops.py docstrings, include/hydra_gnn_b200.h); the kernels themselves are checked on the GPU
(tests/test_gpu_ops.py, tests/test_gpu_train.py).  Nothing under ``hydra_gnn_b200/`` imports this file.
"""
from __future__ import annotations

import contextlib
from dataclasses import dataclass
from typing import Optional

import torch
from torch import Tensor


@dataclass
class CpuCSR:
    """Destination-sorted adjacency with the fields ``ops.CSR`` exposes to the glue."""
    rowptr: Tensor
    col: Tensor
    perm: Optional[Tensor]
    inv_sqrt_deg: Optional[Tensor]
    n_src: int
    n_dst: int
    cap: int
    self_loops: bool
    edge_index: Optional[Tensor] = None
    train: object = None
    join: object = None

    def ready(self):
        return self


def build_csr(edge_index: Tensor, n_src: int, n_dst: int, self_loops: bool = False, norm: bool = False) -> CpuCSR:
    """hg_plan_build's contract: stable sort by destination; ``self_loops`` appends one (i, i) entry per node that has
    none (the test graphs carry no self loops); ``norm``: inv_sqrt_deg = in-degree^-1/2 counted over the final rows."""
    src, dst = edge_index[0].long(), edge_index[1].long()
    eid = torch.arange(src.numel())  # original edge id of every entry, -1 for an added self loop
    if self_loops:
        assert n_src == n_dst and not bool((src == dst).any())
        loop = torch.arange(n_dst)
        src, dst = torch.cat([src, loop]), torch.cat([dst, loop])
        eid = torch.cat([eid, torch.full((n_dst,), -1, dtype=torch.long)])
    order = torch.sort(dst, stable=True).indices
    col = src[order].to(torch.int32)
    perm = eid[order].to(torch.int32)
    counts = torch.bincount(dst, minlength=n_dst)
    rowptr = torch.zeros(n_dst + 1, dtype=torch.int32)
    rowptr[1:] = counts.cumsum(0).to(torch.int32)
    isd = counts.clamp(min=1).float().rsqrt() if norm else None
    return CpuCSR(rowptr, col, perm, isd, n_src, n_dst, int(col.numel()), self_loops,
                  edge_index=torch.stack([src, dst]))


def _act(v: Tensor, act):
    if act is None:
        return v
    if act == "relu":
        return torch.relu(v)
    if act == "elu":
        return torch.nn.functional.elu(v)
    raise ValueError(act)


def seg_reduce(x, csr, op="sum", *, src_scale=None, dst_scale=None, row_map=None, x_self=None, self_coef=None,
               self_coef_add=0.0, bias=None, act=None, out=None):
    """out[i] = act(dst_scale[i] * reduce_{p in row i}(src_scale[col[p]] * x[col[p]]) + (self_coef + add) * x_self[i]
    + bias); rows = row_map (int32) or all."""
    assert not (torch.is_grad_enabled() and x.requires_grad), "raw op called with a gradient pending"
    n_dst, F = csr.n_dst, x.size(1)
    col = csr.col.long()
    deg = (csr.rowptr[1:] - csr.rowptr[:-1]).long()
    row_of = torch.repeat_interleave(torch.arange(n_dst), deg)
    rows = x.index_select(0, col)
    if src_scale is not None:
        rows = rows * src_scale.index_select(0, col)[:, None]
    full = torch.zeros(n_dst, F).index_add_(0, row_of, rows)
    if op == "mean":
        full = full / deg.clamp(min=1).float()[:, None]
    elif op not in ("sum", "add"):
        raise NotImplementedError(op)
    if dst_scale is not None:
        full = full * dst_scale[:, None]
    if x_self is not None:
        coef = self_coef_add + (float(self_coef) if self_coef is not None else 0.0)
        full = full + coef * x_self
    if row_map is not None:
        assert row_map.dtype == torch.int32
        full = full.index_select(0, row_map.long())
    if bias is not None:
        full = full + bias
    return _act(full, act)


def linear(segments, N, *, bias=None, scale=None, shift=None, act=None, precision=None, out=None, pack_cache=None):
    """y = act((sum_s x_s @ W_s^T + bias) * scale + shift)"""
    y = sum(x @ W.t() for x, W in segments)
    assert y.size(1) == N
    if bias is not None:
        y = y + bias
    if scale is not None:
        y = y * scale
    if shift is not None:
        y = y + shift
    return _act(y, act)


def linear_wgrad(dy, x, out=None, accumulate=False):
    dw = dy.t() @ x
    if out is not None:
        out.copy_(out + dw if accumulate else dw)
        return out
    return dw


def act_bwd(dy, y, act=None, want_g=True, want_dbias=False):
    """g = dy * act'(y) with y the activation's OUTPUT; dbias = g.sum(0)"""
    if act is None:
        g = dy
    elif act == "relu":
        g = dy * (y > 0).float()
    elif act == "elu":
        g = dy * torch.where(y > 0, torch.ones_like(y), y + 1.0)
    else:
        raise ValueError(act)
    return (g if want_g else None), (g.sum(0) if want_dbias else None)


def reversed_csr(csr: CpuCSR) -> CpuCSR:
    ei = csr.edge_index
    return build_csr(torch.stack([ei[1], ei[0]]), csr.n_dst, csr.n_src)


def gather_rows(x, rows, act=None, out=None):
    return _act(x.index_select(0, rows.long()), act)


def _gat_group(views, n_dst, H, C, concat, group, act, slope):
    """PyG GATConv per relation (section 9.2 of SURVEY.md) summed / averaged over the relations of one destination
    type.  views: (xs [n_src, H*C], a_src [n_src, H], a_dst [n_dst, H], csr, bias).  Returns (out, [alpha [E', H]])."""
    total, alphas = 0, []
    for xs, a_s, a_d, csr, bias, *rest in views:
        col = csr.col.long()
        deg = (csr.rowptr[1:] - csr.rowptr[:-1]).long()
        row_of = torch.repeat_interleave(torch.arange(n_dst), deg)
        raw = a_s.index_select(0, col) + a_d.index_select(0, row_of)
        if rest and rest[0] is not None:
            raw = raw + rest[0][: col.numel()]  # edge term, CSR order
        e = torch.nn.functional.leaky_relu(raw, slope)
        m = torch.full((n_dst, H), float("-inf")).scatter_reduce(0, row_of[:, None].expand(-1, H), e.detach(), "amax")
        p = torch.exp(e - m.index_select(0, row_of))
        denom = torch.zeros(n_dst, H).index_add(0, row_of, p)
        alpha = p / (denom.index_select(0, row_of) + 1e-16)
        msg = xs.index_select(0, col).view(-1, H, C) * alpha[:, :, None]
        out = torch.zeros(n_dst, H, C).index_add(0, row_of, msg)
        out = out.reshape(n_dst, H * C) if concat else out.mean(1)
        if bias is not None:
            out = out + bias
        total = total + out
        alphas.append(alpha)
    if group == "mean":
        total = total / len(views)
    return _act(total, act), alphas


def gat_fused(rels, n_dst, H, Cc, concat, *, group="sum", act=None, neg_slope=0.2, out=None, addend=None,
              src_shared=False, dropout_p=0.0, seed=0, seed_dev=None):
    """``ops.gat_fused`` (no attention dropout here).  ``src_shared`` (the aggregate-first form of the header): xs is
    [n_src, C], every head weighs the SAME row, output [n_dst, H*C]; ``addend`` joins after the relation group, before
    the activation."""
    assert dropout_p == 0.0
    views = []
    for r in rels:
        xs = r.xs.repeat(1, H) if src_shared else r.xs  # [n_src, H*C] with the shared row under every head
        views.append((xs, r.a_src, r.a_dst, r.csr, r.bias, r.a_edge))
    res, alphas = _gat_group(views, n_dst, H, Cc, True if src_shared else concat, group, None, neg_slope)
    if addend is not None:
        res = res + addend
    res = _act(res, act)
    for r, a in zip(rels, alphas):
        if r.alpha_out is not None:
            r.alpha_out[: a.size(0)].copy_(a)
    return res


def gat_train_bwd(rels, dys, n_dst, H, Cc, concat, out, dout, *, group="sum", act=None, neg_slope=0.2, dropout_p=0.0,
                  seed=0, dbias=None):
    """``ops.gat_train_bwd``: overwrites each relation's column ranges of (dY_src, dY_dst); ``dbias`` accumulates."""
    assert dropout_p == 0.0
    HC = H * Cc
    with torch.enable_grad():
        leaves, views = [], []
        for r in rels:
            xs = r.y_src[:, r.xs_off: r.xs_off + HC].detach().clone().requires_grad_(True)
            a_s = r.y_src[:, r.as_off: r.as_off + H].detach().clone().requires_grad_(True)
            a_d = r.y_dst[:, r.ad_off: r.ad_off + H].detach().clone().requires_grad_(True)
            # the gradient of the raw scores (dlogit) through a zero edge term when the relation has none
            a_e = (r.a_edge if r.a_edge is not None else torch.zeros(r.csr.cap, H)).detach().clone().requires_grad_(True)
            leaves += [xs, a_s, a_d, a_e]
            views.append((xs, a_s, a_d, r.csr, None, a_e))
        pre, _ = _gat_group(views, n_dst, H, Cc, concat, group, None, neg_slope)
        g = act_bwd(dout, out, act)[0] if act is not None else dout
        grads = torch.autograd.grad(pre, leaves, g)
    for k, (r, (dys_, dyd_)) in enumerate(zip(rels, dys)):
        dys_[:, r.xs_off: r.xs_off + HC] = grads[4 * k]
        dys_[:, r.as_off: r.as_off + H] = grads[4 * k + 1]
        dyd_[:, r.ad_off: r.ad_off + H] = grads[4 * k + 2]
    if dbias is not None:
        dbias += g.sum(0) * (1.0 / len(rels) if group == "mean" else 1.0)
    return [grads[4 * k + 3] for k in range(len(rels))]


def edge_logits(edge_attr, csr, v, H, fill=None):
    """a_edge[p, h] = <edge_attr[perm[p]], v[h]> in CSR order; added self loops: <fill, v[h]> or 0."""
    perm = csr.perm.long()
    rows = edge_attr.index_select(0, perm.clamp(min=0)) * (perm >= 0)[:, None].float()
    if fill is not None:
        rows = rows + (perm < 0)[:, None].float() * fill
    return rows @ v.t()


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1, beta2, eps, weight_decay, grad_scale=1.0, step_dev=None):
    """hg_adam_step = torch.optim.Adam's single-tensor update (L2 weight decay added to the gradient), in place."""
    t = int(step_dev) if step_dev is not None else step
    g = grad * grad_scale + weight_decay * param
    exp_avg.mul_(beta1).add_(g, alpha=1 - beta1)
    exp_avg_sq.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    denom = exp_avg_sq.sqrt() / (1 - beta2 ** t) ** 0.5 + eps
    param.addcdiv_(exp_avg, denom, value=-lr / (1 - beta1 ** t))


def build_csr_specs(specs, overlap=False):
    """``ops.build_csr``: one CSR per CSRSpec + the status word (edges with endpoints out of range: none here)."""
    out = []
    for s in specs:
        ei = s.edge_index
        if s.transpose:
            ei = torch.stack([ei[1], ei[0]])
        out.append(build_csr(ei, s.n_dst if s.transpose else s.n_src, s.n_src if s.transpose else s.n_dst,
                             self_loops=s.self_loops, norm=s.want_norm))
    return out, torch.zeros(1, dtype=torch.int32)


@contextlib.contextmanager
def patched_ops():
    """Swap the stand-ins into ``hydra_gnn_b200.ops`` for the duration of a test."""
    from hydra_gnn_b200 import ops

    names = ("seg_reduce", "linear", "linear_wgrad", "act_bwd", "reversed_csr", "gather_rows", "build_csr", "gat_fused",
             "gat_train_bwd", "edge_logits", "adam_step")
    standin = dict(globals(), build_csr=build_csr_specs)
    saved = {n: getattr(ops, n) for n in names}
    try:
        for n in names:
            setattr(ops, n, standin[n])
        yield
    finally:
        for n, f in saved.items():
            setattr(ops, n, f)
