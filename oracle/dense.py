"""Independent dense-adjacency formulation of the layer arithmetic.

ORACLE -- test infrastructure only (see ``oracle/__init__.py``).

The gather/scatter restatement in ``oracle/pyg/nn.py`` cannot be diffed against
real PyG here, so every layer is cross-checked against a second formulation that
shares no code with it: an ``[N_dst, N_src]`` edge-multiplicity matrix ``A`` and
masked dense softmax / matmul.  float64 throughout.
Only usable on tiny graphs (O(N^2) memory).
"""
from __future__ import annotations

import torch


def adjacency(edge_index, n_src, n_dst, dtype=torch.float64):
    """A[i, j] = number of edges j -> i."""
    A = torch.zeros(n_dst, n_src, dtype=dtype)
    if edge_index.numel():
        A.index_put_((edge_index[1], edge_index[0]), torch.ones(edge_index.size(1), dtype=dtype), accumulate=True)
    return A


def with_gat_self_loops(edge_index, n_src, n_dst):
    """GATConv(add_self_loops=True): drop existing loops, add (k,k), k < min(n_src, n_dst)."""
    keep = edge_index[0] != edge_index[1]
    n = min(n_src, n_dst)
    loops = torch.arange(n).repeat(2, 1)
    return torch.cat([edge_index[:, keep], loops], dim=1)


def gat(x_src, x_dst, edge_index, W_src, W_dst, att_src, att_dst, bias, heads, concat,
        add_self_loops, negative_slope=0.2, same_type=False, edge_score=None):
    """Dense GAT.  ``same_type``: single-Tensor call => x_dst projections use W_src.
    ``edge_score``: optional dense [N_dst, N_src, H] additive score (only valid when
    there are no duplicate edges)."""
    f64 = torch.float64
    x_src, x_dst = x_src.to(f64), x_dst.to(f64)
    H = heads
    C = W_src.size(0) // H
    hs = (x_src @ W_src.to(f64).t()).view(-1, H, C)
    hd = hs if same_type else (x_dst @ W_dst.to(f64).t()).view(-1, H, C)
    a_s = (hs * att_src.to(f64).view(1, H, C)).sum(-1)  # [Ns, H]
    a_d = (hd * att_dst.to(f64).view(1, H, C)).sum(-1)  # [Nd, H]
    n_src, n_dst = x_src.size(0), x_dst.size(0)
    if add_self_loops:
        edge_index = with_gat_self_loops(edge_index, n_src, n_dst)
    A = adjacency(edge_index, n_src, n_dst)  # [Nd, Ns]
    s = a_d.unsqueeze(1) + a_s.unsqueeze(0)  # [Nd, Ns, H]
    if edge_score is not None:
        s = s + edge_score.to(f64)
    s = torch.nn.functional.leaky_relu(s, negative_slope)
    mask = (A > 0).unsqueeze(-1)
    s_masked = torch.where(mask, s, torch.full_like(s, -float("inf")))
    m = s_masked.max(dim=1, keepdim=True).values
    m = torch.where(torch.isfinite(m), m, torch.zeros_like(m))
    e = torch.where(mask, (s - m).exp(), torch.zeros_like(s)) * A.unsqueeze(-1)
    w = e / (e.sum(dim=1, keepdim=True) + 1e-16)  # [Nd, Ns, H]
    out = torch.einsum("ijh,jhc->ihc", w, hs)
    out = out.reshape(n_dst, H * C) if concat else out.mean(dim=1)
    if bias is not None:
        out = out + bias.to(f64)
    return out


def sage(x_src, x_dst, edge_index, W_l, b_l, W_r):
    f64 = torch.float64
    A = adjacency(edge_index, x_src.size(0), x_dst.size(0))
    deg = A.sum(1, keepdim=True).clamp(min=1)
    mean = (A @ x_src.to(f64)) / deg
    out = mean @ W_l.to(f64).t() + x_dst.to(f64) @ W_r.to(f64).t()
    if b_l is not None:
        out = out + b_l.to(f64)
    return out


def gcn(x, edge_index, W, bias):
    f64 = torch.float64
    n = x.size(0)
    A = adjacency(edge_index, n, n)
    A = A - torch.diag(torch.diag(A)) + torch.eye(n, dtype=f64)  # loops removed, one unit loop per node
    deg = A.sum(1)
    dis = deg.pow(-0.5)
    dis[torch.isinf(dis)] = 0
    Ahat = dis.unsqueeze(1) * A * dis.unsqueeze(0)
    out = Ahat @ (x.to(f64) @ W.to(f64).t())
    return out + bias.to(f64) if bias is not None else out


def gin(x, edge_index, eps, W0, b0, W1, b1):
    f64 = torch.float64
    n = x.size(0)
    A = adjacency(edge_index, n, n)
    h = (1 + float(eps)) * x.to(f64) + A @ x.to(f64)
    h = torch.relu(h @ W0.to(f64).t() + b0.to(f64))
    return h @ W1.to(f64).t() + b1.to(f64)


def leaf_pool(x, edge_index):
    """LeafPool(mean) with a single-Tensor x: out has x.size(0) rows."""
    n = x.size(0)
    A = adjacency(edge_index, n, n)
    return (A @ x.to(torch.float64)) / A.sum(1, keepdim=True).clamp(min=1)
