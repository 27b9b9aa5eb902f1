"""ctypes binding of ``libhydra_gnn_b200.so`` (the C ABI in ``include/hydra_gnn_b200.h``).

The library is the ONLY compute path of this package: there is no CPU or eager-PyTorch
fallback.  ``lib()`` raises ``RuntimeError`` when the shared object has not been built
(``python __graft_entry__.py`` / ``make -C hydra_gnn_b200/csrc``) or when no CUDA device
is usable.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

HG_MAX_RELATIONS = 8
HG_MAX_SEGMENTS = 8
HG_MAX_PLAN_RELATIONS = 16
HG_TILE_ROWS = 32
HG_PLAN_STATIC_SCRATCH_BYTES = 16 << 20

REDUCE = {"sum": 0, "add": 0, "mean": 1, "max": 2}
ACT = {None: 0, "none": 0, "relu": 1, "elu": 2}
GROUP = {"sum": 0, "mean": 1}
PREC = {"fp32": 0, "tf32x3": 1}

_SO_NAME = "libhydra_gnn_b200.so"
# HYDRA_GNN_B200_SO: alternative build of the same library (A/B runs of kernel variants)
SO_PATH = os.environ.get("HYDRA_GNN_B200_SO") or os.path.join(os.path.dirname(os.path.abspath(__file__)), _SO_NAME)

vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float


class PlanRel(C.Structure):
    _fields_ = [("src", vp), ("dst", vp), ("n_edges", i64), ("n_src", i32), ("n_dst", i32),
                ("self_loops", i32), ("reserved", i32), ("rowptr", vp), ("col", vp), ("perm", vp),
                ("inv_sqrt_deg", vp), ("tile_win", vp)]


class SegArgs(C.Structure):
    _fields_ = [("x", vp), ("ldx", i64), ("rowptr", vp), ("col", vp), ("src_scale", vp), ("dst_scale", vp),
                ("row_map", vp), ("x_self", vp), ("ldself", i64), ("self_coef_ptr", vp), ("self_coef_add", f32),
                ("bias", vp), ("out", vp), ("ldout", i64), ("n_out", i32), ("F", i32), ("op", i32), ("act", i32)]


class GatRel(C.Structure):
    _fields_ = [("xs", vp), ("ldxs", i64), ("a_src", vp), ("lda_src", i64), ("a_dst", vp), ("lda_dst", i64),
                ("a_edge", vp), ("rowptr", vp), ("col", vp), ("bias", vp), ("tile_win", vp), ("n_src", i32), ("reserved", i32)]


class GatArgs(C.Structure):
    _fields_ = [("rel", GatRel * HG_MAX_RELATIONS), ("n_rel", i32), ("n_dst", i32), ("H", i32), ("C", i32),
                ("concat", i32), ("group", i32), ("act", i32), ("neg_slope", f32), ("out", vp), ("ldout", i64),
                ("alpha_saved", vp * HG_MAX_RELATIONS), ("addend", vp), ("ldadd", i64), ("src_shared", i32),
                ("dropout_p", f32), ("dropout_seed", C.c_uint64), ("dropout_seed_dev", vp)]


class LinSeg(C.Structure):
    _fields_ = [("x", vp), ("ldx", i64), ("w", vp), ("ldw", i64), ("K", i32), ("reserved", i32)]


class LinearArgs(C.Structure):
    _fields_ = [("seg", LinSeg * HG_MAX_SEGMENTS), ("n_seg", i32), ("M", i32), ("N", i32), ("bias", vp),
                ("scale", vp), ("shift", vp), ("act", i32), ("precision", i32), ("y", vp), ("ldy", i64),
                ("w_prepacked", i32), ("reserved", i32)]


class GatTrainRel(C.Structure):
    _fields_ = [("xs", vp), ("ldxs", i64), ("a_src", vp), ("lda_src", i64), ("a_dst", vp), ("lda_dst", i64),
                ("rowptr", vp), ("col", vp), ("bias", vp), ("alpha", vp), ("dlogit", vp), ("rowptr_t", vp),
                ("col_t", vp), ("pos_t", vp), ("dxs", vp), ("lddxs", i64), ("da_src", vp), ("ldda_src", i64),
                ("da_dst", vp), ("ldda_dst", i64), ("n_src", i32), ("reserved", i32), ("a_edge", vp)]


class GatTrainArgs(C.Structure):
    _fields_ = [("rel", GatTrainRel * HG_MAX_RELATIONS), ("n_rel", i32), ("n_dst", i32), ("H", i32), ("C", i32),
                ("concat", i32), ("group", i32), ("act", i32), ("neg_slope", f32), ("dropout_p", f32),
                ("reserved", i32), ("seed", C.c_uint64), ("out", vp), ("ldout", i64), ("dout", vp), ("lddout", i64),
                ("g", vp), ("ldg", i64), ("dbias", vp)]


HG_MAX_HTREE_TYPES = 8


class HtreeType(C.Structure):
    _fields_ = [("ptr", vp), ("n_nodes", i64), ("place", vp), ("is_room_virtual", i32), ("is_object_virtual", i32)]


class HtreeRel(C.Structure):
    _fields_ = [("src", vp), ("dst", vp), ("n_edges", i64), ("src_type", i32), ("dst_type", i32)]


class HtreeSplitArgs(C.Structure):
    _fields_ = [("type", HtreeType * HG_MAX_HTREE_TYPES), ("rel", HtreeRel * HG_MAX_PLAN_RELATIONS), ("n_types", i32),
                ("n_rel", i32), ("n_graphs", i32), ("reserved", i32), ("ptr", vp), ("node_type", vp), ("batch", vp),
                ("room_mask", vp), ("object_mask", vp), ("edges", vp * 3), ("status", vp)]


class SynthArgs(C.Structure):
    _fields_ = [("graph_id0", i64), ("n_graphs", i32), ("reserved", i32), ("seed", C.c_uint64), ("cdf_rooms", vp),
                ("cdf_objects", vp), ("prefix", vp), ("n_rooms", i64), ("n_objects", i64), ("n_rr_pairs", i64),
                ("n_oo_pairs", i64), ("word2vec", vp), ("x_rooms", vp), ("x_objects", vp), ("ld_objects", i64),
                ("rooms_to_rooms", vp), ("objects_to_objects", vp), ("objects_to_rooms", vp), ("rooms_to_objects", vp),
                ("room_label", vp), ("object_label", vp), ("room_batch", vp), ("object_batch", vp), ("status", vp)]


_STRUCTS = [PlanRel, SegArgs, GatRel, GatArgs, LinSeg, LinearArgs, GatTrainRel, GatTrainArgs, HtreeSplitArgs, SynthArgs]

# name -> (restype, argtypes); every symbol declared in include/hydra_gnn_b200.h
SYMBOLS = {
    "hg_strerror": (C.c_char_p, [C.c_int]),
    "hg_abi_version": (C.c_int, []),
    "hg_last_cuda_error": (C.c_int, []),
    "hg_device_sm_count": (C.c_int, []),
    "hg_sizeof": (C.c_int, [C.c_int]),
    "hg_plan_workspace_bytes": (C.c_size_t, [C.POINTER(PlanRel), C.c_int]),
    "hg_plan_build": (C.c_int, [C.POINTER(PlanRel), C.c_int, vp, vp, C.c_size_t, vp]),
    "hg_plan_build_static": (C.c_int, [C.POINTER(PlanRel), C.c_int, vp, vp]),
    "hg_seg_reduce": (C.c_int, [C.POINTER(SegArgs), vp]),
    "hg_gat_fused": (C.c_int, [C.POINTER(GatArgs), vp]),
    "hg_linear_workspace_bytes": (C.c_size_t, [C.POINTER(LinearArgs)]),
    "hg_linear_pack_weights": (C.c_int, [C.POINTER(LinearArgs), vp, C.c_size_t, vp]),
    "hg_linear": (C.c_int, [C.POINTER(LinearArgs), vp, C.c_size_t, vp]),
    "hg_edge_logits": (C.c_int, [vp, i64, i32, vp, vp, i64, vp, vp, i32, vp, vp]),
    "hg_relative_pos": (C.c_int, [vp, i64, vp, i64, vp, vp, i64, vp, vp, vp]),
    "hg_gather_rows": (C.c_int, [vp, i64, vp, i64, i32, i32, vp, i64, vp]),
    "hg_collate_offsets": (C.c_int, [vp, i64, vp, vp, vp, i32, vp]),
    "hg_batch_vector": (C.c_int, [vp, i64, vp, i32, vp]),
    "hg_htree_split_workspace_bytes": (C.c_size_t, [C.POINTER(HtreeSplitArgs)]),
    "hg_htree_split": (C.c_int, [C.POINTER(HtreeSplitArgs), vp, C.c_size_t, vp]),
    "hg_synth_mp3d_counts": (C.c_int, [C.POINTER(SynthArgs), vp]),
    "hg_synth_mp3d_fill": (C.c_int, [C.POINTER(SynthArgs), vp]),
    "hg_gat_train_fwd": (C.c_int, [C.POINTER(GatTrainArgs), vp]),
    "hg_gat_train_bwd": (C.c_int, [C.POINTER(GatTrainArgs), vp]),
    "hg_csr_transpose_map": (C.c_int, [vp, vp, vp, i64, vp, vp, vp, i64, i32, vp, vp, vp]),
    "hg_linear_wgrad": (C.c_int, [vp, i64, vp, i64, i32, i32, i32, vp, i64, i32, vp]),
    "hg_act_bwd": (C.c_int, [vp, i64, vp, i64, i64, i32, i32, vp, i64, vp, vp]),
    "hg_adam_step": (C.c_int, [vp, vp, vp, vp, i64, f32, f32, f32, f32, f32, f32, i32, vp, vp]),
}

_lock = threading.Lock()
_handle = None
launch_count = 0  # kernels-launching ABI calls made through this binding (bench's gpu_launches bookkeeping)


def load_shared_object():
    """dlopen + prototype every symbol + verify struct layouts.  Needs no GPU."""
    global _handle
    with _lock:
        if _handle is not None:
            return _handle
        if not os.path.exists(SO_PATH):
            raise RuntimeError(
                f"{_SO_NAME} is not built ({SO_PATH} missing). hydra_gnn_b200 has no CPU / eager fallback: "
                "build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`make -C hydra_gnn_b200/csrc`.")
        h = C.CDLL(SO_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(h, name)  # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        for which, st in enumerate(_STRUCTS):
            if h.hg_sizeof(which) != C.sizeof(st):
                raise RuntimeError(f"ABI mismatch: sizeof({st.__name__}) python={C.sizeof(st)} C={h.hg_sizeof(which)}")
        _handle = h
        return h


def lib():
    """The loaded library, after checking that a CUDA device is usable."""
    h = load_shared_object()
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError("hydra_gnn_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return h


def check(code: int, what: str):
    if code != 0:
        h = load_shared_object()
        msg = h.hg_strerror(code).decode()
        extra = f" (cudaError {h.hg_last_cuda_error()})" if code == -4 else ""
        raise RuntimeError(f"{what}: {msg}{extra}")


def ptr(t):
    """Device pointer of a tensor (or NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr():
    import torch

    return torch.cuda.current_stream().cuda_stream
