"""Host logic of the training glue (hydra_gnn_b200/train.py) on CPU: the autograd Functions composed over plain-torch
stand-ins of the CUDA operators (tests/cpu_backend.py) must reproduce the outputs AND the gradients of the oracle's
layers (oracle/pyg/nn.py, PyG 2.3.1 semantics) -- what each Function saves, which operand receives which gradient, how
degrees / normalisation / row maps / a trainable eps enter the backward.  The kernels are checked on the GPU
(tests/test_gpu_train.py); this file needs none.
"""
import os
import sys

import pytest
import torch
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import cpu_backend as cb  # noqa: E402
from hydra_gnn_b200 import nn as hnn, train  # noqa: E402
from oracle.pyg import nn as onn  # noqa: E402  (checker)


def _graph(n_src, n_dst, n_edges, seed, bipartite=False):
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n_src, (n_edges,), generator=g)
    dst = torch.randint(0, max(n_dst - 3, 1), (n_edges,), generator=g)  # the last rows stay isolated
    if not bipartite:
        keep = src != dst
        src, dst = src[keep], dst[keep]
    ei = torch.unique(torch.stack([src, dst]), dim=1)  # simple graph
    return ei[:, torch.randperm(ei.size(1), generator=g)]  # unsorted, as the loaders emit it


def _pair(oracle_conv, product_conv):
    missing = product_conv.load_state_dict(oracle_conv.state_dict(), strict=False)
    assert not missing.missing_keys, missing
    return oracle_conv, product_conv


def _check(out_ref, out, params_ref, params, inputs_ref, inputs, seed=0):
    torch.testing.assert_close(out, out_ref, rtol=1e-5, atol=1e-6)
    w = torch.randn(out_ref.shape, generator=torch.Generator().manual_seed(seed))
    (out_ref * w).sum().backward()
    (out * w).sum().backward()
    for a, b in zip(inputs_ref, inputs):
        torch.testing.assert_close(b.grad, a.grad, rtol=1e-4, atol=1e-6)
    ref = dict(params_ref)
    n = 0
    for name, p in params:
        if name in ref and ref[name].grad is not None:
            assert p.grad is not None, name
            torch.testing.assert_close(p.grad, ref[name].grad, rtol=1e-4, atol=1e-6, msg=lambda m, name=name: f"{name}: {m}")
            n += 1
    assert n > 0


@pytest.mark.parametrize("act", [None, "relu", "elu"])
def test_sage_layer_gradients(act):
    torch.manual_seed(1)
    ei = _graph(40, 40, 160, 1)
    o, p = _pair(onn.SAGEConv(12, 20), hnn.SAGEConv(12, 20))
    x_ref = torch.randn(40, 12, requires_grad=True)
    x = x_ref.detach().clone().requires_grad_(True)
    with cb.patched_ops():
        out = train.run_sage_train(p, x, x, cb.build_csr(ei, 40, 40), act=act)
        _check(cb._act(o(x_ref, ei), act), out, o.named_parameters(), p.named_parameters(), [x_ref], [x])


def test_bipartite_sage_layer_gradients():
    torch.manual_seed(2)
    ei = _graph(30, 17, 90, 2, bipartite=True)
    o, p = _pair(onn.SAGEConv((9, 5), 11), hnn.SAGEConv((9, 5), 11))
    xs_ref, xd_ref = torch.randn(30, 9, requires_grad=True), torch.randn(17, 5, requires_grad=True)
    xs, xd = xs_ref.detach().clone().requires_grad_(True), xd_ref.detach().clone().requires_grad_(True)
    with cb.patched_ops():
        out = train.run_sage_train(p, xs, xd, cb.build_csr(ei, 30, 17), act="relu")
        _check(torch.relu(o((xs_ref, xd_ref), ei)), out, o.named_parameters(), p.named_parameters(),
               [xs_ref, xd_ref], [xs, xd])


@pytest.mark.parametrize("group", ["sum", "mean"])
def test_hetero_sage_group_gradients(group):
    """Two relations into one node type: the root weights / biases are combined in parameter space (torch autograd),
    the neighbour means go through one K-segmented projection."""
    torch.manual_seed(3)
    ets = [("a", "r1", "b"), ("b", "r2", "b")]
    ei = {ets[0]: _graph(25, 18, 70, 3, bipartite=True), ets[1]: _graph(18, 18, 60, 4)}
    dims = {"a": 7, "b": 6}
    oc = {et: onn.SAGEConv((dims[et[0]], dims[et[2]]), 10) for et in ets}
    pc = {et: hnn.SAGEConv((dims[et[0]], dims[et[2]]), 10) for et in ets}
    for et in ets:
        _pair(oc[et], pc[et])
    x_ref = {"a": torch.randn(25, 7, requires_grad=True), "b": torch.randn(18, 6, requires_grad=True)}
    x = {k: v.detach().clone().requires_grad_(True) for k, v in x_ref.items()}
    outs = [oc[et]((x_ref[et[0]], x_ref[et[2]]), ei[et]) for et in ets]
    ref = torch.nn.functional.elu(sum(outs) / (len(outs) if group == "mean" else 1))
    with cb.patched_ops():
        csr = {et: cb.build_csr(ei[et], x[et[0]].size(0), 18) for et in ets}
        out = train.run_hetero_sage_train(pc, ets, x, csr, act="elu", group=group)["b"]
        named_ref = [(f"{et[1]}.{n}", q) for et in ets for n, q in oc[et].named_parameters()]
        named = [(f"{et[1]}.{n}", q) for et in ets for n, q in pc[et].named_parameters()]
        _check(ref, out, named_ref, named, [x_ref["a"], x_ref["b"]], [x["a"], x["b"]])


@pytest.mark.parametrize("act", [None, "relu"])
def test_gcn_layer_gradients(act):
    torch.manual_seed(5)
    ei = _graph(36, 36, 150, 5)
    o, p = _pair(onn.GCNConv(8, 14), hnn.GCNConv(8, 14))
    x_ref = torch.randn(36, 8, requires_grad=True)
    x = x_ref.detach().clone().requires_grad_(True)
    with cb.patched_ops():
        csr = cb.build_csr(ei, 36, 36, self_loops=True, norm=True)
        out = train.run_gcn_train(p, x, csr, act=act)
        _check(cb._act(o(x_ref, ei), act), out, o.named_parameters(), p.named_parameters(), [x_ref], [x])


def _mlp(i, h, o):
    return nn.Sequential(nn.Linear(i, h), nn.ReLU(), nn.Linear(h, o))


@pytest.mark.parametrize("train_eps", [False, True])
@pytest.mark.parametrize("act", [None, "elu"])
def test_gin_layer_gradients(train_eps, act):
    torch.manual_seed(6)
    ei = _graph(33, 33, 120, 6)
    o, p = _pair(onn.GINConv(_mlp(7, 16, 9), eps=0.3, train_eps=train_eps),
                 hnn.GINConv(_mlp(7, 16, 9), eps=0.3, train_eps=train_eps))
    x_ref = torch.randn(33, 7, requires_grad=True)
    x = x_ref.detach().clone().requires_grad_(True)
    with cb.patched_ops():
        out = train.run_gin_train(p, x, cb.build_csr(ei, 33, 33), act=act)
        _check(cb._act(o(x_ref, ei), act), out, o.named_parameters(), p.named_parameters(), [x_ref], [x])
    if train_eps:
        assert p.eps.grad is not None and p.eps.grad.abs().item() > 0


def test_gin_layer_with_batch_norm_in_training_mode():
    """HomogeneousNetwork's GIN layers: conv -> BatchNorm (batch statistics) -> activation
    (homogeneous_network.py:94-97,133-134); running statistics advance exactly as the oracle's do."""
    torch.manual_seed(7)
    ei = _graph(50, 50, 200, 7)
    o, p = _pair(onn.GINConv(_mlp(6, 12, 12), eps=0.0, train_eps=True), hnn.GINConv(_mlp(6, 12, 12), eps=0.0, train_eps=True))
    obn, pbn = onn.BatchNorm(12), hnn.BatchNorm(12)
    pbn.load_state_dict(obn.state_dict())
    obn.train(), pbn.train()
    x_ref = torch.randn(50, 6, requires_grad=True)
    x = x_ref.detach().clone().requires_grad_(True)
    with cb.patched_ops():
        out = train.run_gin_train(p, x, cb.build_csr(ei, 50, 50), act="relu", bn=pbn)
        ref = torch.relu(obn(o(x_ref, ei)))
        named_ref = list(o.named_parameters()) + [("bn." + n, q) for n, q in obn.named_parameters()]
        named = list(p.named_parameters()) + [("bn." + n, q) for n, q in pbn.named_parameters()]
        _check(ref, out, named_ref, named, [x_ref], [x])
    torch.testing.assert_close(pbn.module.running_mean, obn.module.running_mean)
    torch.testing.assert_close(pbn.module.running_var, obn.module.running_var)


def test_run_gin_dispatches_to_the_training_form():
    """nn.run_gin takes the differentiable route when a gradient is wanted or the BatchNorm is in training mode
    (the inference route folds running statistics into the projection's epilogue and refuses batch statistics)."""
    torch.manual_seed(8)
    ei = _graph(20, 20, 60, 8)
    conv, bn = hnn.GINConv(_mlp(5, 8, 8), train_eps=True), hnn.BatchNorm(8)
    x = torch.randn(20, 5)
    with cb.patched_ops():
        csr = cb.build_csr(ei, 20, 20)
        y = hnn.run_gin(conv, x, csr, act="relu", bn=bn.train())
        assert y.requires_grad and y.shape == (20, 8)
        with torch.no_grad():
            y2 = hnn.run_gin(conv, x, csr, act="relu", bn=bn.train())  # batch statistics, no tape
        assert not y2.requires_grad
        torch.testing.assert_close(y2, y.detach())


def test_seg_reduce_row_map_gradient_reaches_only_mapped_rows():
    """LeafPool / last-layer row selection: rows the forward did not compute carry no gradient."""
    torch.manual_seed(9)
    ei = _graph(28, 28, 100, 9)
    x_ref = torch.randn(28, 5, requires_grad=True)
    x = x_ref.detach().clone().requires_grad_(True)
    rows = torch.tensor([3, 0, 11, 12, 20], dtype=torch.int32)
    bias_ref = torch.randn(5, requires_grad=True)
    bias = bias_ref.detach().clone().requires_grad_(True)
    src, dst = ei[0], ei[1]
    deg = torch.bincount(dst, minlength=28).clamp(min=1).float()
    full = torch.zeros(28, 5).index_add(0, dst, x_ref.index_select(0, src)) / deg[:, None]
    ref = torch.relu(full.index_select(0, rows.long()) + bias_ref)
    with cb.patched_ops():
        out = train.seg_reduce(x, cb.build_csr(ei, 28, 28), "mean", row_map=rows, bias=bias, act="relu")
        _check(ref, out, [("bias", bias_ref)], [("bias", bias)], [x_ref], [x])


# ------------------------------------------------------------------------------ whole models against the REFERENCE's gradients
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np  # noqa: E402

import cases  # noqa: E402
import hydra_gnn_b200 as hg  # noqa: E402

# every golden case of tests/golden/cases.py.  mgat_edge / hom_gat_edge: the edge term of the raw scores is a
# differentiable input of the layer (train._EdgeLogitsFn), its gradient is the kernels' dlogit.  nt_hom_gcn_init
# carries the 306-channel single-head pre_mp, computed in channel chunks (train._run_wide_single_head); hom_gat has 6
# heads, computed in two groups of 3 (train._run_head_groups)
_TRAINABLE_CASES = ["mgat", "msage_all", "nt_het_gat", "nt_het_gat_init", "nt_het_sage_all", "nt_hom_gcn", "nt_hom_sage",
                    "nt_hom_gcn_init", "nt_hom_gat_all", "hom_gcn", "hom_gin", "hom_sage_all", "hom_gat", "mgat_edge",
                    "hom_gat_edge"]


@pytest.mark.parametrize("case", _TRAINABLE_CASES)
def test_model_glue_reproduces_reference_gradient_fingerprints(case):
    """The product's model classes (plan requests, layer wiring, parameter packing of the GAT projections, channel
    padding, row selection, heads, the autograd Functions), run over the CPU stand-ins, against tests/golden/grads.npz
    -- gradient fingerprints computed by the REFERENCE's own model classes (make_golden.py step 6).  The kernels
    themselves meet the oracle's gradients on the GPU (tests/test_gpu_train.py)."""
    golden = cases.load_golden_grads()[case]
    with cb.patched_ops():
        model = cases.build_model(hg, case).train()  # train(): the forward records a tape; dropout off as in the golden run
        model.dropout = 0.0
        for m in model.modules():
            if isinstance(m, hnn.BatchNorm):
                m.eval()  # running statistics, as in the golden (eval-mode) run
            if isinstance(m, hnn.GATConv):
                m.dropout = 0.0
        fp = cases.grad_fingerprint(model, cases.build_input(cases.CASES[case][2]))
    scale = max(abs(v[1]) for v in golden.values() if not np.isnan(v).all())
    checked = 0
    for name, want in golden.items():
        got = fp[name]
        if np.isnan(want).all():
            assert np.isnan(got).all() or np.allclose(got, 0), name
            continue
        # relative to the gradient's own norm, with a floor at the scale of the model's gradients: attention vectors of
        # relations whose rows have one in-edge have a mathematically zero gradient, both sides hold round-off there
        np.testing.assert_allclose(got, want, rtol=0, atol=5e-4 * abs(want[1]) + 1e-6 * scale, err_msg=name)
        checked += 1
    assert checked >= 6


def test_all_golden_cases_are_trainable():
    assert sorted(_TRAINABLE_CASES) == sorted(cases.CASES)


def test_six_head_gat_edge_runs_as_head_groups():
    """GAT_edge with 6 heads (config/mp3d/*_gt*.yaml pairs GAT_edge with 5 and 6 heads): head groups carry their rows of
    lin_edge / att_edge; against the oracle's autograd."""
    from oracle import models as om
    from hydra_gnn_b200 import synth

    kw = dict(input_dim_dict={"objects": 303, "rooms": 3}, output_dim=26, conv_block="GAT_edge", GAT_hidden_dims=[64],
              GAT_heads=[6, 6], GAT_concats=[True, False], dropout=0.0)
    ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).train()
    net = hg.HeterogeneousNetwork(**kw)
    net.load_state_dict(ref.state_dict())
    net.train()
    with cb.patched_ops():
        want = ref(synth.synth_mp3d_batch(range(6), relative_pos=True))
        got = net(synth.synth_mp3d_batch(range(6), relative_pos=True))
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
        probe = torch.randn(want.shape, generator=torch.Generator().manual_seed(3))
        (want * probe).sum().backward()
        (got * probe).sum().backward()
    ref_p, got_p = dict(ref.named_parameters()), dict(net.named_parameters())
    scale = max(float(p.grad.abs().max()) for p in ref_p.values() if p.grad is not None)
    checked = 0
    for name, p in ref_p.items():
        if p.grad is None:
            continue
        torch.testing.assert_close(got_p[name].grad, p.grad, rtol=5e-4, atol=1e-5 * scale, msg=lambda m, name=name: f"{name}: {m}")
        checked += "edge" in name
    assert checked >= 8  # lin_edge.weight and att_edge of the live relations


@pytest.mark.parametrize("heads,concats", [([6, 6, 6], [False, False, False]), ([5, 5, 5], [True, True, False]),
                                           ([6, 6], [True, False])])
def test_hetero_gat_with_more_heads_than_one_launch_takes(heads, concats):
    """The reference's configs go to 5 and 6 heads (config/mp3d/*.yaml) while one launch of the GAT training kernels
    takes 4: the layer is split into head groups (train._run_head_groups).  Heterogeneous GAT, head mean with the
    26 -> 28 channel padding and concat, against the oracle's autograd."""
    from oracle import models as om
    from hydra_gnn_b200 import synth

    kw = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT",
              GAT_hidden_dims=[64] * (len(heads) - 1), GAT_heads=heads, GAT_concats=concats, dropout=0.0)
    ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).train()
    net = hg.HeterogeneousNetwork(**kw)
    net.load_state_dict(ref.state_dict())
    net.train()
    data = synth.synth_mp3d_batch(range(6))
    with cb.patched_ops():
        want, got = ref(synth.synth_mp3d_batch(range(6))), net(data)
        torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
        probe = torch.randn(want.shape, generator=torch.Generator().manual_seed(3))
        (want * probe).sum().backward()
        (got * probe).sum().backward()
    ref_p, got_p = dict(ref.named_parameters()), dict(net.named_parameters())
    scale = max(float(p.grad.abs().max()) for p in ref_p.values() if p.grad is not None)
    checked = 0
    for name, p in ref_p.items():
        if p.grad is None:
            assert got_p[name].grad is None or float(got_p[name].grad.abs().max()) == 0.0, name
            continue
        torch.testing.assert_close(got_p[name].grad, p.grad, rtol=5e-4, atol=1e-5 * scale, msg=lambda m, name=name: f"{name}: {m}")
        checked += 1
    assert checked >= 20


@pytest.mark.parametrize("conv_block,heads", [("GAT", [3, 3, 3]), ("GAT_edge", [6, 6, 6])])
def test_training_steps_track_torch_adam(conv_block, heads):
    """Whole steps -- forward, masked cross-entropy (label 25 ignored, base_training_job.py:63), backward, FlatAdam over the
    re-homed flat buffers (weight decay as in base_training_job.py:77-79) -- against the oracle model under
    torch.optim.Adam: the parameters after 4 steps."""
    from oracle import models as om
    from hydra_gnn_b200 import synth
    from hydra_gnn_b200.train import FlatAdam

    edge = conv_block == "GAT_edge"
    kw = dict(input_dim_dict={"objects": 303, "rooms": 3} if edge else {"objects": 306, "rooms": 6}, output_dim=26,
              conv_block=conv_block, GAT_hidden_dims=[64, 64], GAT_heads=heads, GAT_concats=[False] * 3, dropout=0.0)
    ref = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).train()
    net = hg.HeterogeneousNetwork(**kw)
    net.load_state_dict(ref.state_dict())
    net.train()
    data = lambda: synth.synth_mp3d_batch(range(8), relative_pos=edge)  # noqa: E731
    y = data()["rooms"].y.long()
    mask = y != 25
    lr, wd, steps = 1e-2, 1e-3, 4
    opt_ref = torch.optim.Adam(ref.parameters(), lr=lr, weight_decay=wd)
    with cb.patched_ops():
        opt = FlatAdam(net.parameters(), lr=lr, weight_decay=wd)
        for _ in range(steps):
            opt_ref.zero_grad()
            ref.loss(ref(data()), y, mask).backward()
            opt_ref.step()
            opt.zero_grad()
            net.loss(net(data()), y, mask).backward()
            opt.step()
    got = dict(net.named_parameters())
    start = synth.random_state_dict_(om.HeterogeneousNetwork(**kw), seed=0).state_dict()
    moved = 0
    for name, p in ref.named_parameters():
        err = float((got[name].detach() - p.detach()).abs().max())
        assert err <= 0.05 * lr * steps, f"{name}: {err:.3e}"  # Adam normalises: compare in units of the step size
        moved += float((p.detach() - start[name]).abs().max()) > 0
    assert moved >= 20


def test_eval_after_flat_adam_steps_uses_the_updated_weights():
    """The reference loop trains an epoch, then ``net.eval()`` + validation, every epoch (base_training_job.py:197-232).
    FlatAdam updates the weights through a raw pointer (neither ``data_ptr`` nor ``_version`` of a parameter moves), so
    every packed-weight cache must be keyed on more than those: step, eval, step, eval -- each eval must equal the eval of a
    fresh model holding the same weights (cold caches)."""
    from hydra_gnn_b200 import synth
    from hydra_gnn_b200.train import FlatAdam

    kw = dict(input_dim_dict={"objects": 306, "rooms": 6}, output_dim=26, conv_block="GAT", GAT_hidden_dims=[64, 64],
              GAT_heads=[3, 3, 3], GAT_concats=[False] * 3, dropout=0.0)
    net = synth.random_state_dict_(hg.HeterogeneousNetwork(**kw), seed=0)
    data = lambda: synth.synth_mp3d_batch(range(8))  # noqa: E731
    y = data()["rooms"].y.long()
    mask = y != 25

    def cold_eval():
        fresh = hg.HeterogeneousNetwork(**kw)
        fresh.load_state_dict({k: v.clone() for k, v in net.state_dict().items()})
        return fresh.eval()(data())

    from hydra_gnn_b200 import ops
    saved = ops.PARALLEL_BRANCHES
    ops.PARALLEL_BRANCHES = False  # the inference path forks side streams for independent branches: none on a CPU
    try:
        _eval_after_steps(net, data, y, mask, cold_eval, FlatAdam)
    finally:
        ops.PARALLEL_BRANCHES = saved


def _eval_after_steps(net, data, y, mask, cold_eval, FlatAdam):
    with cb.patched_ops():
        opt = FlatAdam(net.parameters(), lr=1e-2)
        seen = []
        for steps in (1, 3):
            net.train()
            for _ in range(steps):
                opt.zero_grad()
                net.loss(net(data()), y, mask).backward()
                opt.step()
            net.eval()
            got = net(data())
            assert torch.equal(got, net(data()))  # warm caches agree with themselves ...
            torch.testing.assert_close(got, cold_eval(), rtol=0, atol=0)  # ... and with a cold model
            seen.append(got)
        assert float((seen[0] - seen[1]).abs().max()) > 1e-3  # the weights did move between the two evals
        # same without the train()/eval() switch in between (generation counter in the cache keys)
        opt.zero_grad()
        with torch.enable_grad():
            net.train()
            net.loss(net(data()), y, mask).backward()
        net.eval()
        before = net(data())
        opt.step()  # eval mode, no mode switch after the step
        after = net(data())
        assert float((before - after).abs().max()) > 1e-4
        torch.testing.assert_close(after, cold_eval(), rtol=0, atol=0)


def test_flat_adam_survives_zero_grad_set_to_none_and_exposes_param_groups():
    """``net.zero_grad()`` (set_to_none) detaches the parameters from the flat gradient buffer; the next step must use the
    fresh gradients, not stale flat contents.  ``param_groups[0]['lr']`` drives the step size (StepLR in the reference)."""
    from hydra_gnn_b200.train import FlatAdam

    lin_a, lin_b = torch.nn.Linear(3, 2), torch.nn.Linear(3, 2)
    lin_b.load_state_dict(lin_a.state_dict())
    x = torch.randn(5, 3)
    with cb.patched_ops():
        opt = FlatAdam(lin_a.parameters(), lr=1e-2)
        ref = torch.optim.Adam(lin_b.parameters(), lr=1e-2)
        for k in range(4):
            lin_a.zero_grad()  # set_to_none=True: p.grad rebound by the next backward
            ref.zero_grad()
            (lin_a(x) ** 2).sum().backward()
            (lin_b(x) ** 2).sum().backward()
            if k == 2:
                opt.param_groups[0]["lr"] = 1e-3
                ref.param_groups[0]["lr"] = 1e-3
            opt.step()
            ref.step()
    torch.testing.assert_close(lin_a.weight, lin_b.weight, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lin_a.bias, lin_b.bias, rtol=1e-5, atol=1e-6)
