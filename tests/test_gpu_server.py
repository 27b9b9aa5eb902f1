"""f4 -- the flow of the reference's room_classification_server (bin/room_classification_server:196-204,273-299):

    model = model_class(**network_params)            # network_params as stored in model.yaml
    model.load_state_dict(torch.load(weight_path))   # a checkpoint written by PyG-based training
    model.to(device)
    pred = model(data.to(device)).argmax(dim=1).cpu()  # ONE scene graph per call, under torch.no_grad()

run through the product for the four model classes: the checkpoint is the ORACLE's state_dict re-spelled the way
PyG >= 2.4 writes it ('<a___b___c>' hetero keys, a single shared 'lin.weight' for homogeneous GATConv) and saved /
loaded with torch.save / torch.load; the batch-1 forward is replayed as a CUDA graph (the server's steady state) and
the predicted labels must equal the oracle's on every graph."""
import io
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import cases  # noqa: E402
import hydra_gnn_b200 as hg  # noqa: E402
from hydra_gnn_b200 import convert, synth  # noqa: E402
from oracle import models as om  # noqa: E402  (the checker)
from util import assert_close, clone_data  # noqa: E402

pytestmark = pytest.mark.gpu


def _pyg24_spelling(sd, shared_gat_lin: bool):
    out = {}
    for k, v in sd.items():
        parts = k.split(".")
        for i, part in enumerate(parts):
            if "__" in part and i > 0 and parts[i - 1] == "convs":
                parts[i] = "<" + "___".join(part.split("__")) + ">"
        k2 = ".".join(parts)
        if shared_gat_lin:
            if k2.endswith("lin_dst.weight"):
                continue
            k2 = k2.replace("lin_src.weight", "lin.weight")
        out[k2] = v.clone()
    return out


def _single_graphs(kind):
    if kind == "dsg":
        return [synth.synth_mp3d_batch([g]) for g in (0, 1, 2, 5)]
    if kind == "dsg_relpos":
        return [synth.synth_mp3d_batch([g], relative_pos=True) for g in (0, 1, 2, 5)]
    if kind == "dsg_hom":
        return [convert.dsg_to_homogeneous(synth.synth_mp3d_batch([g])) for g in (0, 1, 2, 5)]
    trees = cases.load_htrees("htree_synth")[:4]
    if kind == "htree":
        return trees
    return [convert.heterogeneous_htree_to_homogeneous(t) for t in trees]


SERVER_CASES = [("mgat", "dsg", False), ("mgat_edge", "dsg_relpos", False), ("hom_gat", "dsg_hom", True),
                ("nt_het_gat_init", "htree", False), ("nt_hom_gcn", "htree_hom", False)]


@pytest.mark.parametrize("case,kind,shared", SERVER_CASES)
def test_server_flow_checkpoint_to_batch1_argmax(case, kind, shared, tmp_path):
    cls, network_params, _ = cases.CASES[case]
    ref = cases.build_model(om, case, seed=11)  # the "trained" model
    buf = tmp_path / "model_weights.pth"
    torch.save(_pyg24_spelling(ref.state_dict(), shared), buf)

    model = getattr(hg, cls)(**network_params)              # GnnModel.__init__
    model.load_state_dict(torch.load(buf))
    model.to(torch.device("cuda"))
    model.eval()

    graphs = _single_graphs(kind)
    with torch.no_grad():
        for data in graphs:
            want = ref(clone_data(data))
            dev = clone_data(data, "cuda")
            pred = model(dev).argmax(dim=1).cpu()           # GnnModel.infer
            assert_close(model(dev), want, f"{case} batch-1 logits")
            assert torch.equal(pred, want.argmax(dim=1)), case
            # the server's steady state: the same batch-1 forward replayed as a CUDA graph
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                model(dev)
                s.synchronize()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    logits = model(dev)
            torch.cuda.synchronize()
            for _ in range(3):
                g.replay()
            torch.cuda.synchronize()
            assert torch.equal(logits.argmax(dim=1).cpu(), want.argmax(dim=1)), f"{case}: CUDA-graph replay"
            assert_close(logits, want, f"{case} replayed logits")
