"""First-run check of the opt-in tcgen05 weight-gradient kernel (hydra_gnn_b200/csrc/wgrad_tc.cu, HG_WGRAD_TC=1).

Runs in its OWN process (a protocol bug in a tcgen05 kernel traps after ~2 s and poisons the CUDA context -- never
call this path from the pytest process until this script is green):

    timeout 300 python tools/check_wgrad_tc.py            # correctness on small + ragged shapes, then timing
    timeout 300 python tools/check_wgrad_tc.py --quick    # correctness only

For every shape: dW from the tcgen05 kernel vs an fp64 torch reference (bound: 2e-6 of |dY|^T |X|, the 3-term TF32
bound the default kernel is tested with) and vs the default kernel.  Prints one JSON line per shape and exits
non-zero on the first mismatch.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

SHAPES = [  # (M, N, K): one k-block, ragged tails in every dimension, both orientations, several tiles, config-5 shapes
    (32, 16, 16), (31, 16, 16), (33, 24, 8), (1000, 64, 64), (4097, 88, 64), (4096, 396, 64), (4099, 64, 396),
    (8192, 396, 306), (5000, 306, 396), (3001, 130, 209), (2048, 26, 128), (65536, 396, 306),
]
BIG = [(488696, 396, 306), (488696, 396, 64), (488696, 88, 64), (488696, 64, 396)]


def run(tc: bool, dy, x, out=None, accumulate=False):
    from hydra_gnn_b200 import ops
    if tc:
        os.environ["HG_WGRAD_TC"] = "1"
    else:
        os.environ.pop("HG_WGRAD_TC", None)
    return ops.linear_wgrad(dy, x, out=out, accumulate=accumulate)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(7)
    for M, N, K in SHAPES:
        # padded leading dimensions and an offset start: the kernel may assume nothing about alignment
        dyb = torch.randn(M, N + 3, device=dev, generator=g)
        xb = torch.randn(M, K + 1, device=dev, generator=g)
        dy, x = dyb[:, 1:N + 1], xb[:, :K]
        want = dy.double().t() @ x.double()
        bound = 2e-6 * (dy.double().abs().t() @ x.double().abs()) + 1e-30
        got = run(True, dy, x)
        torch.cuda.synchronize()
        base = run(False, dy, x)
        acc = run(True, dy, x, out=got.clone(), accumulate=True)  # accumulate = 1: second product on top
        torch.cuda.synchronize()
        err = ((got.double() - want).abs() / bound).max().item()
        err_base = ((base.double() - want).abs() / bound).max().item()
        err_acc = ((acc.double() - 2 * want).abs() / (2 * bound)).max().item()
        ok = err <= 1.0 and err_acc <= 1.0 and bool(torch.isfinite(got).all())
        print(json.dumps(dict(M=M, N=N, K=K, err_over_bound=round(err, 4), default_kernel=round(err_base, 4),
                              accumulate=round(err_acc, 4), ok=ok)), flush=True)
        if not ok:
            bad = ((got.double() - want).abs() / bound).argmax().item()
            print(f"FIRST MISMATCH at dW[{bad // K}, {bad % K}]: got {got.view(-1)[bad].item()} want {want.view(-1)[bad].item()}")
            sys.exit(1)
    if args.quick:
        return
    for M, N, K in BIG:
        dy = torch.randn(M, N, device=dev, generator=g)
        x = torch.randn(M, K, device=dev, generator=g)
        res = {}
        for tc in (False, True):
            for _ in range(3):
                run(tc, dy, x)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                out = run(tc, dy, x)
            e.record()
            torch.cuda.synchronize()
            res["tc" if tc else "default"] = (s.elapsed_time(e) / 10, out)
        diff = (res["tc"][1] - res["default"][1]).abs().max().item() / res["default"][1].abs().max().item()
        print(json.dumps(dict(M=M, N=N, K=K, default_us=round(res["default"][0] * 1e3, 1), tc_us=round(res["tc"][0] * 1e3, 1),
                              tflops_tc=round(2 * M * N * K / res["tc"][0] / 1e9, 1), rel_diff=diff)), flush=True)


if __name__ == "__main__":
    main()
