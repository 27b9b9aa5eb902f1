"""Training step of the GAT models (BASELINE.json configs[4]; reference: base_training_job.py:203-219 --
``loss = model.loss(model(data), label, mask); loss.backward(); optimizer.step()``).

``torch.autograd`` is the tape only: every gradient of a hot operator is produced by a hand-written kernel
(``hg_gat_train_bwd``, ``hg_linear_wgrad``, ``hg_linear`` with the transposed weight) and the optimizer update by
``hg_adam_step`` over one flat parameter buffer.  Parameter-space bookkeeping on tensors of a few hundred rows
(stacking ``lin_src`` blocks and the folded attention rows into the per-node-type projection matrix) is left to
torch's own differentiable ops, which also carries the gradient of the fold back to ``att_src`` / ``att_dst``.

The training form keeps every relation project-first (the aggregate-first rewrite of the inference path would
need the gradient of the re-arranged weights; not worth it while the weight gradient dominates the step).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from . import ops
from .ops import CSR, GatTrainRelation

EdgeType = Tuple[str, str, str]


def _pad4(n: int) -> int:
    return (n + 3) // 4 * 4


# =============================================================================== projection
class _LinearFn(torch.autograd.Function):
    """``y = x @ W^T`` on hg_linear; backward: ``dx = dy @ W`` (hg_linear), ``dW = dy^T @ x`` (hg_linear_wgrad)."""

    @staticmethod
    def forward(ctx, x: Tensor, W: Tensor):
        ctx.save_for_backward(x, W)
        return ops.linear([(x, W)], W.size(0))

    @staticmethod
    def backward(ctx, dy: Tensor):
        x, W = ctx.saved_tensors
        dx = dW = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear([(dy, W.t().contiguous())], W.size(1))
        if ctx.needs_input_grad[1]:
            dW = ops.linear_wgrad(dy, x)
        return dx, dW


def linear(x: Tensor, W: Tensor) -> Tensor:
    return _LinearFn.apply(x, W)


# =============================================================================== GAT layer
class _EdgeLogitsFn(torch.autograd.Function):
    """``a_edge[p, h] = <edge_attr[perm[p]], v[h]>`` in CSR order (``hg_edge_logits``; added self loops take ``fill`` or 0).
    ``v = W_edge^T att_edge`` is folded in parameter space (torch autograd carries its gradient to ``lin_edge.weight`` and
    ``att_edge``).  Backward: ``dv[h] = sum_p da_edge[p, h] * edge_attr[perm[p]]`` -- a gather of the 3-wide attribute rows
    into CSR order and one ``hg_linear_wgrad``; the incoming gradient is zero past the live entry count."""

    @staticmethod
    def forward(ctx, v, edge_attr, csr, H, fill):
        ctx.csr, ctx.fill = csr, fill
        ctx.save_for_backward(edge_attr)
        return ops.edge_logits(edge_attr, csr, v.contiguous(), H, fill=fill)

    @staticmethod
    def backward(ctx, da):
        (edge_attr,) = ctx.saved_tensors
        csr = ctx.csr
        n = min(da.size(0), csr.perm.numel())  # an empty relation still has a 1-row buffer
        if n == 0:
            return da.new_zeros(da.size(1), edge_attr.size(1)), None, None, None, None
        da = da[:n]
        perm = csr.perm[:n].long()
        live = torch.arange(n, device=da.device) < csr.rowptr[-1]
        real = live & (perm >= 0)
        rows = edge_attr.index_select(0, torch.where(real, perm, torch.zeros_like(perm))) * real[:, None].to(da.dtype)
        if ctx.fill is not None:
            rows = rows + (live & (perm < 0))[:, None].to(da.dtype) * ctx.fill
        return ops.linear_wgrad(da.contiguous(), rows.contiguous()), None, None, None, None


class _GatLayerFn(torch.autograd.Function):
    """All destination types of one GAT layer.  Inputs: ``cfg`` (python description), then the projection outputs
    ``Y_T`` (one per node type, in ``cfg.types`` order), then the relation biases.  Outputs: one tensor per
    destination type.  The backward fills ONE gradient buffer per ``Y_T`` (each column range is owned by exactly one
    relation role, so no accumulation across relations is needed)."""

    @staticmethod
    def forward(ctx, cfg, *tensors):
        nT, nB = len(cfg.types), len(cfg.bias_rels)
        Y = dict(zip(cfg.types, tensors[:nT]))
        biases = dict(zip(cfg.bias_rels, tensors[nT:nT + nB]))
        edges = dict(zip(cfg.edge_rels, tensors[nT + nB:]))  # [cap, H] edge terms of the raw scores (GAT_edge)
        outs, saved_rels = [], []
        for T in cfg.dsts:
            rels = [GatTrainRelation(Y[et[0]], Y[T], cfg.xs_off[et], cfg.as_off[et], cfg.ad_off[et], cfg.csr[et],
                                     biases.get(et), None, edges.get(et)) for et in cfg.rels if et[2] == T]
            out = _forward_group(rels, Y[T].size(0), cfg.H[T], cfg.C[T], cfg.concat[T], cfg.group, cfg.act,
                                 cfg.slope[T], cfg.dropout[T], cfg.seed[T])
            outs.append(out)
            saved_rels.append(rels)
        ctx.cfg, ctx.saved_rels, ctx.Y = cfg, saved_rels, Y
        ctx.save_for_backward(*outs)
        return tuple(outs)

    @staticmethod
    def backward(ctx, *douts):
        cfg, Y = ctx.cfg, ctx.Y
        outs = ctx.saved_tensors
        # every xs / a_src / a_dst column is written by exactly one relation role; only the pad columns need zeros
        dY = {T: torch.empty_like(y) for T, y in Y.items()}
        for T, spans in cfg.pad_cols.items():
            for a, b in spans:
                dY[T][:, a:b].zero_()
        dbias: Dict[str, Tensor] = {}
        dedge: Dict[EdgeType, Tensor] = {}
        for T, rels, out, dout in zip(cfg.dsts, ctx.saved_rels, outs, douts):
            has_bias = any(r.bias is not None for r in rels)
            db = torch.zeros(out.size(1), dtype=torch.float32, device=out.device) if has_bias else None
            dbias[T] = db
            mine = [et for et in cfg.rels if et[2] == T]
            dys = [(dY[et[0]], dY[T]) for et in mine]
            dlogits = ops.gat_train_bwd(rels, dys, out.size(0), cfg.H[T], cfg.C[T], cfg.concat[T], out, dout.contiguous(),
                                        group=cfg.group, act=cfg.act, neg_slope=cfg.slope[T], dropout_p=cfg.dropout[T],
                                        seed=cfg.seed[T], dbias=db)
            for et, dl in zip(mine, dlogits):
                dedge[et] = dl  # the gradient of the raw scores IS the gradient of their edge term
        grads: List[Optional[Tensor]] = [None]
        grads += [dY[T] for T in cfg.types]
        used = set()
        for et in cfg.bias_rels:  # the bias gradient is the same vector for every relation of a group
            g = dbias[et[2]]
            grads.append(g.clone() if et[2] in used else g)
            used.add(et[2])
        grads += [dedge[et] for et in cfg.edge_rels]
        ctx.saved_rels = ctx.Y = None
        return tuple(grads)


# Device-resident dropout seed (CUDA-graph training): when set, every attention-dropout kernel adds this scalar to its
# (static) per-call salt, so a replayed graph draws a new mask each step; ``advance_seed()`` belongs inside the graph.
_DEVICE_SEED: Optional[Tensor] = None
_salt = 0


def enable_device_seed(device, seed: int = 0) -> Tensor:
    global _DEVICE_SEED
    _DEVICE_SEED = torch.full((1,), int(seed), dtype=torch.int64, device=device)
    return _DEVICE_SEED


def disable_device_seed() -> None:
    global _DEVICE_SEED
    _DEVICE_SEED = None


def advance_seed() -> None:
    if _DEVICE_SEED is not None:
        _DEVICE_SEED.add_(0x632BE5AB)  # odd stride; the kernels hash (seed, relation, edge, head)


TILE_FORWARD = os.environ.get("HG_TRAIN_TILE_FWD", "1") != "0"  # training forward through the inference tile kernel (hg_gat_fused + dropout + saved weights)


def _forward_group(rels: Sequence[GatTrainRelation], n_dst: int, H: int, C: int, concat: bool, group: str, act: str,
                   slope: float, dropout_p: float, seed: int) -> Tensor:
    """Forward of one destination type.  Default: ``hg_gat_fused`` (the tile kernel of the inference path) with the
    softmax weights saved for the backward and attention dropout applied in its softmax stage; ``TILE_FORWARD =
    False`` selects ``hg_gat_train_fwd`` (row-per-half-warp kernel, same results up to the fast exp)."""
    if not TILE_FORWARD:
        if any(r.a_edge is not None for r in rels):
            raise NotImplementedError("edge attributes need the tile forward (HG_TRAIN_TILE_FWD=1)")
        if _DEVICE_SEED is not None and dropout_p > 0:
            raise NotImplementedError("the device-resident dropout seed needs the tile forward (HG_TRAIN_TILE_FWD=1)")
        return ops.gat_train_fwd(rels, n_dst, H, C, concat, group=group, act=act, neg_slope=slope, dropout_p=dropout_p,
                                 seed=seed)
    HC = H * C
    fused = []
    for r in rels:
        r.alpha = torch.empty(max(r.csr.cap, 1), H, dtype=torch.float32, device=r.y_src.device)
        fused.append(ops.GatRelation(r.y_src[:, r.xs_off: r.xs_off + HC], r.y_src[:, r.as_off: r.as_off + H],
                                     r.y_dst[:, r.ad_off: r.ad_off + H], r.csr, r.bias, r.a_edge, r.alpha))
    return ops.gat_fused(fused, n_dst, H, C, concat, group=group, act=act, neg_slope=slope, dropout_p=dropout_p,
                         seed=seed, seed_dev=_DEVICE_SEED if dropout_p > 0 else None)


class _LayerCfg:
    pass


def run_gat_layer_train(convs: Dict[EdgeType, nn.Module], rels: Sequence[EdgeType], x: Dict[str, Tensor],
                        csr: Dict[EdgeType, CSR], *, act: str = None, group: str = "sum",
                        edge_attr=None) -> Dict[str, Tensor]:
    """Differentiable form of ``nn.run_gat_layer`` (same HeteroConv semantics)."""
    from .nn import GATConv
    if any(convs[et].heads == 1 and _pad4(convs[et].out_channels) > 256 for et in rels):
        return _run_wide_single_head(convs, rels, x, csr, act=act, group=group, edge_attr=edge_attr)
    if any(convs[et].heads > 4 or convs[et].heads * _kernel_channels(convs[et]) > 256 for et in rels):
        return _run_head_groups(convs, rels, x, csr, act=act, group=group, edge_attr=edge_attr)

    cfg = _LayerCfg()
    cfg.rels = list(rels)
    cfg.types = []
    for s, _, d in rels:
        for t in (s, d):
            if t not in cfg.types:
                cfg.types.append(t)
    cfg.dsts = list(dict.fromkeys(et[2] for et in rels))
    cfg.xs_off, cfg.as_off, cfg.ad_off, cfg.csr = {}, {}, {}, dict(csr)
    cfg.H, cfg.C, cfg.concat, cfg.slope, cfg.dropout, cfg.seed = {}, {}, {}, {}, {}, {}
    cfg.group, cfg.act, cfg.C_real = group, act, {}
    for et in rels:
        c = convs[et]
        T = et[2]
        # head mean with C % 4 != 0 (the 26-class layer): the per-head blocks of the projection are padded to a multiple
        # of 4 channels (zero weight rows, zero bias) so the float4 kernels apply; the pad columns of the output are 0
        # and sliced away below
        Cp = _pad4(c.out_channels) if (not c.concat or c.heads == 1) else c.out_channels
        if c.heads > 4 or c.heads * Cp > 256:  # the limits hg_gat_train_fwd / _bwd validate (HG_EUNSUPPORTED)
            raise NotImplementedError(f"GAT training kernels take heads <= 4 and heads * channels <= 256 (got {c.heads} x "
                                      f"{c.out_channels}: e.g. the 306-channel pre_mp of the homogeneous Neural Tree)")
        cfg.C_real[T] = c.out_channels
        sig = (c.heads, Cp, c.concat, c.negative_slope, c.dropout if c.training else 0.0)
        if T in cfg.H and (cfg.H[T], cfg.C[T], cfg.concat[T], cfg.slope[T], cfg.dropout[T]) != sig:
            raise NotImplementedError("relations targeting one node type must share heads/channels/concat/dropout")
        cfg.H[T], cfg.C[T], cfg.concat[T], cfg.slope[T], cfg.dropout[T] = sig
    global _salt
    for T in cfg.dsts:
        if cfg.dropout[T] <= 0:
            cfg.seed[T] = 0
        elif _DEVICE_SEED is not None:  # static salt per call site; the device scalar makes it fresh per step
            _salt += 1
            cfg.seed[T] = (_salt * 0x9E3779B97F4A7C15) & 0x3FFFFFFFFFFFFFFF
        else:  # host-side draw from torch's CPU generator (torch.manual_seed makes a run repeatable)
            cfg.seed[T] = int(torch.randint(0, 2 ** 62, (1,)).item())

    # per node type: [lin_src blocks of the relations leaving it | folded att_src rows | folded att_dst rows]
    Y: Dict[str, Tensor] = {}
    cfg.pad_cols = {}
    for T in cfg.types:
        blocks, col = [], 0
        pads = cfg.pad_cols.setdefault(T, [])
        for et in rels:
            if et[0] != T:
                continue
            c = convs[et]
            W = c.lin_src.weight
            Cp = cfg.C[et[2]]
            if Cp != c.out_channels:
                W = F.pad(W.view(c.heads, c.out_channels, -1), (0, 0, 0, Cp - c.out_channels)).reshape(c.heads * Cp, -1)
            rows = c.heads * Cp
            blocks.append(W)
            if _pad4(rows) != rows:
                blocks.append(W.new_zeros(_pad4(rows) - rows, W.size(1)))
                pads.append((col + rows, col + _pad4(rows)))
            cfg.xs_off[et] = col
            col += _pad4(rows)
        for et in rels:
            if et[0] != T:
                continue
            c = convs[et]
            fs = getattr(c, "fold_src", None)  # a channel chunk of a wide conv carries the whole conv's folded rows
            blocks.append(fs if fs is not None else GATConv.fold(c.att_src, c.lin_src.weight, c.heads, c.out_channels))
            cfg.as_off[et] = col
            col += c.heads
        for et in rels:
            if et[2] != T:
                continue
            c = convs[et]
            Wd = c.lin_src.weight if et[0] == et[2] else c.lin_dst.weight
            fd = getattr(c, "fold_dst", None)
            blocks.append(fd if fd is not None else GATConv.fold(c.att_dst, Wd, c.heads, c.out_channels))
            cfg.ad_off[et] = col
            col += c.heads
        if _pad4(col) != col:
            blocks.append(blocks[0].new_zeros(_pad4(col) - col, x[T].size(1)))
            pads.append((col, _pad4(col)))
        Y[T] = linear(x[T], torch.cat(blocks, dim=0))

    # GAT_edge: the edge term of the raw scores, a_edge[p, h] = <edge_attr[perm[p]], (W_edge^T att_edge)[h]>, per relation
    cfg.edge_rels = [et for et in rels if convs[et].lin_edge is not None and edge_attr is not None
                     and edge_attr.get(et) is not None]
    for T in cfg.dsts:
        mine = [et for et in rels if et[2] == T]
        if 0 < sum(et in cfg.edge_rels for et in mine) < len(mine):
            raise NotImplementedError("relations targeting one node type must all carry edge attributes, or none")
    a_edges = []
    for et in cfg.edge_rels:
        c = convs[et]
        v = GATConv.fold(c.att_edge, c.lin_edge.weight, c.heads, c.out_channels)  # [H, edge_dim], parameter space
        a_edges.append(_EdgeLogitsFn.apply(v, edge_attr[et], csr[et], c.heads, c.fill_vector(x[et[2]].device)))
    cfg.bias_rels = [et for et in rels if convs[et].bias is not None]
    biases = []
    for et in cfg.bias_rels:
        b, width = convs[et].bias, (cfg.H[et[2]] * cfg.C[et[2]] if cfg.concat[et[2]] else cfg.C[et[2]])
        biases.append(b if b.numel() == width else F.pad(b, (0, width - b.numel())))
    outs = _GatLayerFn.apply(cfg, *[Y[T] for T in cfg.types], *biases, *a_edges)
    res = {}
    for T, o in zip(cfg.dsts, outs):
        res[T] = o if cfg.C[T] == cfg.C_real[T] else o[:, :cfg.C_real[T]]
    return res


def _kernel_channels(c) -> int:
    """Channels per head as the kernels see them (head mean with C % 4 != 0 is padded, see run_gat_layer_train)."""
    return _pad4(c.out_channels) if (not c.concat or c.heads == 1) else c.out_channels


class _HeadGroup:
    """Heads [h0, h1) of a GATConv as a conv of its own (heads attend independently: own rows of the projections, of the
    attention vectors and -- concat -- of the bias).  Head mean: the caller weights the group by (h1 - h0) / H, and the
    bias rides on the first group, pre-scaled so that the weighting restores it."""

    def __init__(self, conv, h0: int, h1: int, first: bool):
        H, C = conv.heads, conv.out_channels
        self.heads, self.out_channels, self.concat = h1 - h0, C, conv.concat
        self.negative_slope, self.dropout, self.training = conv.negative_slope, conv.dropout, conv.training
        self.lin_edge = None if conv.lin_edge is None else _Rows(conv.lin_edge.weight[h0 * C: h1 * C])
        self.att_edge = None if conv.lin_edge is None else conv.att_edge[:, h0:h1]
        self.fill_vector = conv.fill_vector
        self.lin_src = _Rows(conv.lin_src.weight[h0 * C: h1 * C])
        self.lin_dst = self.lin_src if conv.lin_dst is conv.lin_src else _Rows(conv.lin_dst.weight[h0 * C: h1 * C])
        self.att_src, self.att_dst = conv.att_src[:, h0:h1], conv.att_dst[:, h0:h1]
        if conv.bias is None:
            self.bias = None
        elif conv.concat:
            self.bias = conv.bias[h0 * C: h1 * C]
        else:
            self.bias = conv.bias * (H / (h1 - h0)) if first else None


def _run_head_groups(convs, rels, x, csr, *, act, group, edge_attr):
    """A GAT layer with more heads (or more heads x channels) than one launch of the training kernels takes
    (heads <= 4, heads x channels <= 256; the reference's configs go to 6 heads): the heads are split into groups, each
    group is a GAT layer of its own over its rows of the parameters, and the groups are concatenated (concat) or
    averaged with weights |group| / H (head mean) before the activation.  Exact: heads do not interact before that."""
    heads = {convs[et].heads for et in rels}
    concat = {bool(convs[et].concat) for et in rels}
    if len(heads) != 1 or len(concat) != 1:
        raise NotImplementedError("splitting heads into groups needs every relation of the layer to share heads / concat")
    H, is_concat = heads.pop(), concat.pop()
    per = min(4, 256 // max(_kernel_channels(convs[et]) for et in rels))
    if per < 1:
        raise NotImplementedError("GAT training kernels take heads x channels <= 256 with more than one head")
    n_groups = -(-H // per)
    size = -(-H // n_groups)
    parts: Dict[str, List[Tensor]] = {}
    for k, h0 in enumerate(range(0, H, size)):
        h1 = min(h0 + size, H)
        res = run_gat_layer_train({et: _HeadGroup(convs[et], h0, h1, k == 0) for et in rels}, rels, x, csr, act=None,
                                  group=group, edge_attr=edge_attr)
        for T, o in res.items():
            parts.setdefault(T, []).append(o if is_concat else o * ((h1 - h0) / H))
    out = {}
    for T, ps in parts.items():
        o = torch.cat(ps, dim=1) if is_concat else sum(ps)
        out[T] = o if act is None else (torch.relu(o) if act == "relu" else F.elu(o))
    return out


class _Rows:
    """Stands where ``run_gat_layer_train`` reads ``conv.lin_src.weight``."""

    def __init__(self, weight: Tensor):
        self.weight = weight


class _ChannelChunk:
    """Output channels [a, b) of a single-head GATConv as a conv of its own: its rows of the projection and of the
    bias, and the attention logits of the WHOLE conv (``fold_src`` / ``fold_dst``, read by the packing above)."""

    def __init__(self, conv, a: int, b: int, same_type: bool):
        from .nn import GATConv
        C = conv.out_channels
        self.heads, self.out_channels, self.concat = 1, b - a, conv.concat
        self.negative_slope, self.dropout, self.training, self.lin_edge = conv.negative_slope, 0.0, conv.training, None
        self.lin_src = _Rows(conv.lin_src.weight[a:b])
        self.lin_dst = self.lin_src
        self.att_src = self.att_dst = None
        self.bias = conv.bias[a:b] if conv.bias is not None else None
        self.fold_src = GATConv.fold(conv.att_src, conv.lin_src.weight, 1, C)
        self.fold_dst = GATConv.fold(conv.att_dst, conv.lin_src.weight if same_type else conv.lin_dst.weight, 1, C)


def _run_wide_single_head(convs, rels, x, csr, *, act, group, edge_attr):
    """A single-head GAT layer wider than the training kernels' 256 columns (the 306-channel ``pre_mp`` of the
    homogeneous Neural Tree, homogeneous_neural_tree_network.py:63-66,86): with one head the softmax weights are shared
    by every channel, so the layer is computed in channel chunks -- each chunk a GAT layer over its own columns of the
    projection with the attention logits of the whole conv -- and the chunks are concatenated.  Exact: the aggregation
    is per channel, and the softmax backward is linear in ``d alpha = sum_c g_c xs_c``, so the per-chunk contributions to
    the attention gradients add up in parameter space (torch autograd sums them through the shared folded rows)."""
    if len(rels) != 1:
        raise NotImplementedError("a single-head GAT layer wider than 256 channels must be the only relation of its layer")
    et = rels[0]
    c = convs[et]
    if c.training and c.dropout > 0:
        raise NotImplementedError("attention dropout on a channel-chunked GAT layer (every chunk must draw the same mask)")
    if c.lin_edge is not None and edge_attr is not None and edge_attr.get(et) is not None:
        raise NotImplementedError("training with edge attributes (GAT_edge) is not implemented")
    C = c.out_channels
    n_chunks = -(-_pad4(C) // 256)
    step = _pad4(-(-C // n_chunks))
    outs = []
    for a in range(0, C, step):
        chunk = _ChannelChunk(c, a, min(a + step, C), et[0] == et[2])
        outs.append(run_gat_layer_train({et: chunk}, [et], x, csr, act=act, group=group)[et[2]])
    return {et[2]: torch.cat(outs, dim=1)}


# =============================================================================== optimizer
class FlatAdam:
    """``torch.optim.Adam(params, lr, weight_decay=...)`` (base_training_job.py:77-79) over ONE flat buffer:
    after the first backward the live parameters (those that received a gradient -- PyG's dead ``lin_dst`` of
    same-type relations never does, and torch's Adam skips ``grad is None``) are re-homed as views of
    ``flat_param`` / ``flat_grad``, so a step is one ``hg_adam_step`` launch and a data-parallel gradient
    exchange is one all-reduce of ``flat_grad``."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0):
        seen, self.params = set(), []
        for p in params:
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                self.params.append(p)
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.t = 0
        self.flat_param = self.flat_grad = self.exp_avg = self.exp_avg_sq = None
        self.live: List[nn.Parameter] = []

    def flatten(self):
        """Idempotent; call after the first backward (``step`` does) and before a gradient all-reduce."""
        if self.flat_param is not None:
            return
        self.live = [p for p in self.params if p.grad is not None]
        if not self.live:
            raise RuntimeError("FlatAdam.step() before any backward")
        dev = self.live[0].device
        n = sum(p.numel() for p in self.live)
        self.flat_param = torch.empty(n, dtype=torch.float32, device=dev)
        self.flat_grad = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.t_dev = torch.full((1,), self.t, dtype=torch.int32, device=dev)  # step count t, device-resident so a
        off = 0                                                               # captured step advances it on replay
        with torch.no_grad():
            for p in self.live:
                k = p.numel()
                self.flat_param[off:off + k].copy_(p.reshape(-1))
                self.flat_grad[off:off + k].copy_(p.grad.reshape(-1))
                p.data = self.flat_param[off:off + k].view(p.shape)
                p.grad = self.flat_grad[off:off + k].view(p.shape)
                off += k

    @property
    def param_groups(self):
        """One group, torch.optim style (``lr`` is read from here at every eager step, so a ``StepLR``-like scheduler that
        rewrites ``param_groups[0]['lr']`` works; a captured CUDA graph bakes the value it was captured with)."""
        if "_groups" not in self.__dict__:
            self._groups = [{"params": self.params, "lr": self.lr, "betas": self.betas, "eps": self.eps,
                             "weight_decay": self.weight_decay}]
        return self._groups

    def zero_grad(self, set_to_none: bool = False):
        if self.flat_grad is None:
            for p in self.params:
                p.grad = None
        else:
            self.flat_grad.zero_()

    def _rebind_grads(self):
        """``net.zero_grad()`` (set_to_none) or ``p.grad = None`` detaches a parameter from ``flat_grad``; the next
        backward then writes a fresh ``.grad`` tensor.  Re-home it (copy + rebind) instead of stepping on stale data."""
        off = 0
        for p in self.live:
            k = p.numel()
            view = self.flat_grad[off:off + k]
            if p.grad is None:
                view.zero_()
                p.grad = view.view(p.shape)
            elif p.grad.data_ptr() != view.data_ptr():
                with torch.no_grad():
                    view.copy_(p.grad.reshape(-1))
                p.grad = view.view(p.shape)
            off += k

    def step(self, grad_scale: float = 1.0):
        self.flatten()
        if not (self.flat_param.is_cuda and torch.cuda.is_current_stream_capturing()):
            self._rebind_grads()
        g = self.param_groups[0]
        self.lr, self.weight_decay = g["lr"], g["weight_decay"]
        self.t += 1  # host mirror (not advanced by graph replays); the kernel reads the device counter
        self.t_dev.add_(1)
        ops.adam_step(self.flat_param, self.flat_grad, self.exp_avg, self.exp_avg_sq, self.t, self.lr, self.betas[0],
                      self.betas[1], self.eps, self.weight_decay, grad_scale, step_dev=self.t_dev)
        ops.bump_param_generation()  # weights changed through a raw pointer: every packed-weight cache key moves on


# =============================================================================== SAGE / GCN / LeafPool (a3, a6, a8)
def wants_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in tensors)


class _LinearMultiFn(torch.autograd.Function):
    """``y = act(sum_s x_s @ W_s^T + bias)`` (the K-segmented hg_linear).  Backward: ``g = dy * act'(y)`` and the bias
    gradient in one ``hg_act_bwd`` pass, ``dx_s = g @ W_s`` (hg_linear), ``dW_s = g^T @ x_s`` (hg_linear_wgrad)."""

    @staticmethod
    def forward(ctx, act, n_seg, *tensors):
        xs, Ws, bias = tensors[:n_seg], tensors[n_seg:2 * n_seg], tensors[2 * n_seg]
        y = ops.linear(list(zip(xs, Ws)), Ws[0].size(0), bias=bias, act=act)
        ctx.act, ctx.n_seg, ctx.has_bias = act, n_seg, bias is not None
        ctx.save_for_backward(y, *xs, *Ws)
        return y

    @staticmethod
    def backward(ctx, dy):
        n = ctx.n_seg
        y, xs, Ws = ctx.saved_tensors[0], ctx.saved_tensors[1:1 + n], ctx.saved_tensors[1 + n:]
        need = ctx.needs_input_grad[2:]
        g, dbias = ops.act_bwd(dy.contiguous(), y, ctx.act, want_dbias=ctx.has_bias and need[2 * n])
        dxs = [ops.linear([(g, W.t().contiguous())], W.size(1)) if need[k] else None for k, W in enumerate(Ws)]
        dWs = [ops.linear_wgrad(g, x) if need[n + k] else None for k, x in enumerate(xs)]
        return (None, None, *dxs, *dWs, dbias)


def linear_multi(segments, bias: Optional[Tensor] = None, act: str = None) -> Tensor:
    xs, Ws = [x for x, _ in segments], [w for _, w in segments]
    return _LinearMultiFn.apply(act, len(segments), *xs, *Ws, bias)


class _SegReduceFn(torch.autograd.Function):
    """``out = act(dst_scale * reduce(src_scale * x[col]) + bias)`` (hg_seg_reduce, sum | mean, optional ``row_map``).
    Backward: ``g = dout * act'(out)`` (hg_act_bwd, + bias gradient), then the SAME kernel over the reversed CSR:
    ``dx[j] = src_scale[j] * sum_{i <- j} (dst_scale[i] / deg[i]) g[i]`` -- a gather, no atomics."""

    @staticmethod
    def forward(ctx, x, bias, csr, op, src_scale, dst_scale, row_map, act):
        out = ops.seg_reduce(x, csr, op, src_scale=src_scale, dst_scale=dst_scale, row_map=row_map, bias=bias, act=act)
        ctx.csr, ctx.op, ctx.src_scale, ctx.dst_scale, ctx.row_map, ctx.act = csr, op, src_scale, dst_scale, row_map, act
        ctx.has_bias = bias is not None
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, dout):
        (out,) = ctx.saved_tensors
        csr = ctx.csr
        g, dbias = ops.act_bwd(dout.contiguous(), out, ctx.act, want_dbias=ctx.has_bias and ctx.needs_input_grad[1])
        dx = None
        if ctx.needs_input_grad[0]:
            if ctx.row_map is not None:  # rows the forward did not compute carry no gradient
                full = torch.zeros(csr.n_dst, g.size(1), dtype=torch.float32, device=g.device)
                full.index_copy_(0, ctx.row_map.long(), g)
                g = full
            scale = ctx.dst_scale
            if ctx.op == "mean":
                csr.ready()
                inv_deg = (csr.rowptr[1:] - csr.rowptr[:-1]).clamp(min=1).to(torch.float32).reciprocal()
                scale = inv_deg if scale is None else scale * inv_deg
            dx = ops.seg_reduce(g, ops.reversed_csr(csr), "sum", src_scale=scale, dst_scale=ctx.src_scale)
        return dx, dbias, None, None, None, None, None, None


def seg_reduce(x: Tensor, csr: CSR, op: str = "sum", *, src_scale: Tensor = None, dst_scale: Tensor = None,
               row_map: Tensor = None, bias: Tensor = None, act: str = None) -> Tensor:
    """``ops.seg_reduce`` that records its backward when a gradient is wanted (sum / mean)."""
    if not wants_grad(x, bias):
        return ops.seg_reduce(x, csr, op, src_scale=src_scale, dst_scale=dst_scale, row_map=row_map, bias=bias, act=act)
    if op not in ("sum", "add", "mean"):
        raise NotImplementedError("backward of the max aggregation is not implemented")
    return _SegReduceFn.apply(x, bias, csr, op, src_scale, dst_scale, row_map, act)


def run_sage_train(conv, x_src: Tensor, x_dst: Tensor, csr: CSR, act: str = None) -> Tensor:
    """SAGEConv, differentiable: ``lin_l(mean_j x_j) + lin_r(x_i)`` as one K-segmented projection."""
    mean = seg_reduce(x_src, csr, "mean")
    return linear_multi([(mean, conv.lin_l.weight), (x_dst, conv.lin_r.weight)], conv.lin_l.bias, act)


def run_hetero_sage_train(convs, rels: Sequence[EdgeType], x: Dict[str, Tensor], csr: Dict[EdgeType, CSR], *,
                          act: str = None, group: str = "sum") -> Dict[str, Tensor]:
    out: Dict[str, Tensor] = {}
    for T in dict.fromkeys(et[2] for et in rels):
        mine = [et for et in rels if et[2] == T]
        scale = 1.0 / len(mine) if group == "mean" else 1.0
        w_root = sum(convs[et].lin_r.weight for et in mine) * scale  # parameter-space ops: torch autograd
        biases = [convs[et].lin_l.bias for et in mine if convs[et].lin_l.bias is not None]
        bias = (sum(biases) * scale) if biases else None
        segs = [(seg_reduce(x[et[0]], csr[et], "mean"),
                 convs[et].lin_l.weight * scale if scale != 1.0 else convs[et].lin_l.weight) for et in mine]
        segs.append((x[T], w_root))
        out[T] = linear_multi(segs, bias, act)
    return out


def run_gcn_train(conv, x: Tensor, csr: CSR, act: str = None) -> Tensor:
    y = linear_multi([(x, conv.lin.weight)])
    return seg_reduce(y, csr, "sum", src_scale=csr.inv_sqrt_deg, dst_scale=csr.inv_sqrt_deg, bias=conv.bias, act=act)


def run_gin_train(conv, x: Tensor, csr: CSR, act: str = None, bn=None) -> Tensor:
    """GINConv, differentiable: ``nn((1 + eps) x_i + sum_j x_j)`` -- the neighbour sum is
    ``hg_seg_reduce`` (backward: the same kernel over the reversed CSR), the self term is a torch elementwise op
    (it carries the gradient of a trainable ``eps``), every Linear of the MLP is ``hg_linear`` with its ReLU fused.
    ``bn`` (HomogeneousNetwork, homogeneous_network.py:94-97,133-134): PyG ``BatchNorm`` between the last Linear and
    the layer's activation -- batch statistics in training mode, torch's ``BatchNorm1d`` (library call; GIN is on
    none of the benchmark configurations)."""
    h = seg_reduce(x, csr, "sum") + (1.0 + conv.eps) * x
    layers = list(conv.nn)
    lin_idx = [k for k, m in enumerate(layers) if isinstance(m, nn.Linear)]
    for k, m in enumerate(layers):
        if isinstance(m, nn.Linear):
            last = k == lin_idx[-1]
            fuse_relu = (not last) and k + 1 < len(layers) and isinstance(layers[k + 1], nn.ReLU)
            fused = "relu" if fuse_relu else (act if (last and bn is None) else None)
            h = linear_multi([(h, m.weight)], m.bias, fused)
        elif isinstance(m, nn.ReLU):
            continue  # fused into the preceding projection
        else:
            raise NotImplementedError(f"GINConv.nn may hold Linear / ReLU only, got {type(m).__name__}")
    if bn is not None:
        h = bn.module(h)
        if act == "relu":
            h = torch.relu(h)
        elif act == "elu":
            h = torch.nn.functional.elu(h)
        elif act is not None:
            raise NotImplementedError(f"activation {act!r}")
    return h
