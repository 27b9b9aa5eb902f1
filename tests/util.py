"""Shared helpers for the parity tests: the same model built twice (oracle on CPU, product on
CUDA) with identical non-trivial weights, and comparison at the tolerance north_star states
(1e-4 relative in fp32, identical argmax)."""
import copy

import torch

from hydra_gnn_b200 import synth

RTOL, ATOL = 1e-4, 1e-5  # BASELINE.json north_star: logits within 1e-4 relative in fp32


def clone_data(data, device=None):
    out = copy.copy(data)
    from hydra_gnn_b200.data import Data, HeteroData

    if isinstance(data, HeteroData):
        new = HeteroData()
        for t in data.node_types:
            for k, v in data[t].items():
                new[t][k] = v.clone() if torch.is_tensor(v) else v
        for t in data.edge_types:
            for k, v in data[t].items():
                new[t][k] = v.clone() if torch.is_tensor(v) else v
        for k, v in data._g.items():
            new._g[k] = v
    else:
        new = Data(**{k: (v.clone() if torch.is_tensor(v) else v) for k, v in data.items()})
    return new.to(device) if device is not None else new


def make_pair(oracle_cls, product_cls, seed=0, **kwargs):
    """(oracle model on CPU, product model on CUDA) with the same seeded weights, eval mode."""
    ref = oracle_cls(**kwargs)
    synth.random_state_dict_(ref, seed=seed)
    ref.eval()
    prod = product_cls(**kwargs)
    missing = prod.load_state_dict(ref.state_dict(), strict=True)
    prod.eval()
    if torch.cuda.is_available():
        prod = prod.cuda()
    return ref, prod


def assert_close(got, want, what="", rtol=RTOL, atol=ATOL, check_argmax=True):
    got = got.detach().float().cpu()
    want = want.detach().float().cpu()
    assert got.shape == want.shape, f"{what}: shape {tuple(got.shape)} vs {tuple(want.shape)}"
    if got.numel() == 0:
        return
    err = (got - want).abs()
    # "within 1e-4 relative": elementwise rtol plus an absolute floor that scales with the tensor's
    # magnitude (two fp32 evaluations of a K-term sum differ by ~eps*sqrt(K)*|terms|, not *|result|)
    atol = atol * max(1.0, float(want.abs().max()))
    tol = atol + rtol * want.abs()
    worst = (err - tol).max().item()
    assert worst <= 0, f"{what}: max abs err {err.max().item():.3e}, worst excess over tol {worst:.3e}"
    if check_argmax and got.dim() == 2 and got.size(1) > 1:
        # identical predicted labels; rows whose top-2 oracle logits tie within tolerance are exempt
        top2 = want.topk(2, dim=1).values
        decisive = (top2[:, 0] - top2[:, 1]) > 4 * (atol + rtol * top2[:, 0].abs())
        same = got.argmax(1) == want.argmax(1)
        assert bool(same[decisive].all()), f"{what}: argmax differs on {int((~same[decisive]).sum())} decisive rows"
