"""ctypes binding of the C-ABI in include/stemgnn_b200.h (libstemgnn_b200.so, built in-tree by
`__graft_entry__.build()`).  There is no CPU path: a missing library or a non-CUDA tensor is an
error, never a fallback."""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libstemgnn_b200.so")
ABI_VERSION = 6

_i64, _i32, _vp, _f32, _f64 = C.c_int64, C.c_int, C.c_void_p, C.c_float, C.c_double


class AggregateArgs(C.Structure):
    _fields_ = [("rowptr", _vp), ("col", _vp), ("perm", _vp), ("n_rows", _i64),
                ("n_gather_rows", _i64), ("num_edges", _i64), ("feat", _i64), ("gathered", _vp),
                ("row_cnt", _vp), ("gather_cnt", _vp), ("edge_feat", _vp), ("type_table", _vp),
                ("edge_type", _vp), ("edge_weight", _vp), ("row_x", _vp), ("act", _i32),
                ("backward", _i32), ("out", _vp), ("out_ld", _i64), ("row_x_ld", _i64),
                ("hub_ws", _vp), ("hub_ws_bytes", C.c_size_t), ("accumulate", _i32), ("reserved", _i32),
                ("row_add", _vp), ("row_add_ld", _i64)]


class VqAssignArgs(C.Structure):
    _fields_ = [("x", _vp), ("x_row_stride", _i64), ("x_head_stride", _i64), ("embed", _vp),
                ("heads", _i64), ("embed_heads", _i64), ("rows", _i64), ("codes", _i64),
                ("dim", _i64), ("metric", _i32), ("path", _i32), ("ind", _vp),
                ("ind_row_stride", _i64), ("ind_head_stride", _i64), ("quantize", _vp),
                ("q_row_stride", _i64), ("q_head_stride", _i64), ("sq_err", _vp), ("stats", _vp)]


class VqCodeStatsArgs(C.Structure):
    _fields_ = [("x", _vp), ("x_row_stride", _i64), ("x_head_stride", _i64), ("ind", _vp),
                ("ind_row_stride", _i64), ("ind_head_stride", _i64), ("heads", _i64),
                ("embed_heads", _i64), ("rows", _i64), ("codes", _i64), ("dim", _i64),
                ("counts", _vp), ("sums", _vp)]


class LinearArgs(C.Structure):
    _fields_ = [("m", _i64), ("n", _i64), ("n_src", _i32), ("w_transposed", _i32), ("a", _vp * 2),
                ("lda", _i64 * 2), ("k", _i64 * 2), ("w", _vp * 2), ("ldw", _i64 * 2), ("bias", _vp),
                ("out", _vp), ("ldo", _i64), ("accumulate", _i32), ("reserved", _i32), ("row_div", _vp)]


class LinearWgradArgs(C.Structure):
    _fields_ = [("m", _i64), ("n", _i64), ("n_src", _i32), ("reserved", _i32), ("g", _vp), ("ldg", _i64),
                ("a", _vp * 2), ("lda", _i64 * 2), ("k", _i64 * 2), ("gw", _vp * 2), ("ldgw", _i64 * 2),
                ("gbias", _vp)]


class NeighborSampleArgs(C.Structure):
    _fields_ = [("rowptr", _vp), ("col", _vp), ("perm", _vp), ("num_nodes", _i64), ("num_edges", _i64),
                ("seeds", _vp), ("num_seeds", _i64), ("num_hops", _i32), ("reserved", _i32),
                ("fanout_host", C.POINTER(C.c_int32)), ("seed", _vp), ("node_map", _vp),
                ("node_cap", _i64), ("edge_cap", _i64), ("n_id", _vp), ("edge_index_out", _vp),
                ("e_id", _vp), ("counts", _vp)]


# name -> (restype, argtypes); every symbol include/stemgnn_b200.h declares
SIGNATURES = {
    "stemgnn_last_error": (C.c_char_p, []),
    "stemgnn_abi_version": (_i32, []),
    "stemgnn_launch_count": (C.c_uint64, []),
    "stemgnn_set_sm_reserve": (_i32, [_i32]),
    "stemgnn_csr_build_workspace": (_i32, [_i64, _i64, C.POINTER(C.c_size_t)]),
    "stemgnn_csr_build": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_csr_build_pair_workspace": (_i32, [_i64, _i64, C.POINTER(C.c_size_t)]),
    "stemgnn_csr_build_pair": (_i32, [_vp, _i64, _i64, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                      _vp, C.c_size_t, _vp]),
    "stemgnn_gcn_edge_weight": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "stemgnn_aggregate": (_i32, [C.POINTER(AggregateArgs), _vp]),
    "stemgnn_aggregate_hub_workspace": (_i32, [_i64, C.POINTER(C.c_size_t)]),
    "stemgnn_row_scale": (_i32, [_i64, _i64, _vp, _vp, _vp, _vp]),
    "stemgnn_edge_dot": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _i32, _vp, _vp]),
    "stemgnn_linear_workspace": (_i32, [C.POINTER(LinearArgs), C.POINTER(C.c_size_t)]),
    "stemgnn_linear": (_i32, [C.POINTER(LinearArgs), _vp, C.c_size_t, _vp]),
    "stemgnn_linear_wgrad_workspace": (_i32, [C.POINTER(LinearWgradArgs), C.POINTER(C.c_size_t)]),
    "stemgnn_linear_wgrad": (_i32, [C.POINTER(LinearWgradArgs), _vp, C.c_size_t, _vp]),
    "stemgnn_bn_workspace": (_i32, [_i64, _i64, C.POINTER(C.c_size_t)]),
    "stemgnn_bn_stats": (_i32, [_i64, _i64, _vp, _i64, _vp, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_bn_train_stats": (_i32, [_i64, _i64, _vp, _i64, _f32, _f64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_bn_stats_static": (_i32, [_i64, _i64, _vp, _i64, _vp, _vp, _vp]),
    "stemgnn_bn_act_forward": (_i32, [_i64, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _f32, _vp, _vp, _i64, _vp]),
    "stemgnn_bn_act_backward": (_i32, [_i64, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _f32, _vp, _i32,
                                       _vp, _vp, _vp, _i64, _vp, C.c_size_t, _vp]),
    "stemgnn_bn_act_backward_apply": (_i32, [_i64, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _f32, _vp,
                                             _vp, _vp, _i64, _vp, _i64, _vp]),
    "stemgnn_vq_assign_workspace": (_i32, [C.POINTER(VqAssignArgs), C.POINTER(C.c_size_t)]),
    "stemgnn_vq_assign": (_i32, [C.POINTER(VqAssignArgs), _vp, C.c_size_t, _vp]),
    "stemgnn_vq_backward": (_i32, [_vp, _vp, _vp, _vp, _f32, _i64, _vp, _vp]),
    "stemgnn_vq_code_stats_workspace": (_i32, [C.POINTER(VqCodeStatsArgs), C.POINTER(C.c_size_t)]),
    "stemgnn_vq_code_stats": (_i32, [C.POINTER(VqCodeStatsArgs), _vp, C.c_size_t, _vp]),
    "stemgnn_vq_ema_apply": (_i32, [_vp, _vp, _i64, _i64, _i64, _f64, _f64, _i32, _vp, _vp, _vp,
                                    _vp, _vp]),
    "stemgnn_dropout_adj_workspace": (_i32, [_i64, C.POINTER(C.c_size_t)]),
    "stemgnn_dropout_adj_count": (_i32, [_vp, _i64, _vp, _vp, _f32, _i32, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_dropout_adj_write": (_i32, [_vp, _i64, _vp, _vp, _f32, _i32, _i64, _vp, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_mask_fill": (_i32, [_i64, _i64, _vp, _i64, _vp, _vp, _f32, _vp, _f32, _vp, _i64, _vp, _vp]),
    "stemgnn_gather_rows": (_i32, [_vp, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _vp]),
    "stemgnn_neighbor_sample_bounds": (_i32, [_i64, _i64, _i64, _i32, C.POINTER(C.c_int32), C.POINTER(_i64),
                                              C.POINTER(_i64), C.POINTER(C.c_size_t)]),
    "stemgnn_node_map_init": (_i32, [_vp, _i64, _vp]),
    "stemgnn_neighbor_sample": (_i32, [C.POINTER(NeighborSampleArgs), _vp, C.c_size_t, _vp]),
    "stemgnn_negative_sampling_workspace": (_i32, [_i64, _i64, C.POINTER(C.c_size_t)]),
    "stemgnn_negative_sampling": (_i32, [_vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp, _vp, C.c_size_t, _vp]),
    "stemgnn_l2norm_forward": (_i32, [_i64, _i64, _vp, _f32, _vp, _vp, _vp]),
    "stemgnn_l2norm_backward": (_i32, [_i64, _i64, _vp, _vp, _vp, _f32, _vp, _vp]),
}

_lib = None


class StemGnnError(RuntimeError):
    pass


def lib():
    """The loaded shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise StemGnnError(
                f"{SO_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`."
                " stemgnn_b200 has no CPU or PyTorch fallback.")
        h = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype, fn.argtypes = res, args
        if h.stemgnn_abi_version() != ABI_VERSION:
            raise StemGnnError("libstemgnn_b200.so ABI version mismatch; rebuild")
        _lib = h
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().stemgnn_last_error()
        raise StemGnnError(f"stemgnn error {rc}: {msg.decode() if msg else ''}")


def ptr(t):
    """Device pointer of a CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise StemGnnError("stemgnn_b200 operates on CUDA tensors only (no CPU fallback)")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


# STEMGNN_WS_POISON = ff | 00 | rand: fill every workspace before the call (debugging aid for
# tests/test_gpu_vq_stability.py and scripts/vq_stability.py: a kernel that reads workspace bytes it
# did not write shows up as a run-to-run difference).  Unset: plain torch.empty.
_POISON = os.environ.get("STEMGNN_WS_POISON", "")


def set_workspace_poison(mode):
    """mode: "", "ff", "00" or "rand" (see STEMGNN_WS_POISON)."""
    global _POISON
    if mode not in ("", "ff", "00", "rand"):
        raise ValueError("poison mode must be '', 'ff', '00' or 'rand'")
    _POISON = mode


def workspace(nbytes, device):
    n = max(int(nbytes), 1)
    if not _POISON:
        return torch.empty(n, dtype=torch.uint8, device=device)
    if _POISON == "rand":
        return torch.randint(0, 256, (n,), dtype=torch.uint8, device=device)
    return torch.full((n,), 255 if _POISON == "ff" else 0, dtype=torch.uint8, device=device)


# ------------------------------------------------------------------ per-call device timing
# Optional CUDA-event trace of every C-ABI call (used by bench.py for the roofline of the
# dominant kernel "live over the timed region").  Off by default: zero overhead.
TRACE = None


class traced:
    """with traced("name", algorithmic_bytes, flops): <C-ABI call(s)>"""
    __slots__ = ("rec",)

    def __init__(self, name, nbytes=0, flops=0, launches=0):
        self.rec = None
        if TRACE is not None:
            self.rec = [name, nbytes, flops, launches, torch.cuda.Event(enable_timing=True),
                        torch.cuda.Event(enable_timing=True)]

    def __enter__(self):
        if self.rec is not None:
            self.rec[3] = -launch_count()
            self.rec[4].record()
        return self

    def __exit__(self, *exc):
        if self.rec is not None:
            self.rec[5].record()
            self.rec[3] += launch_count()      # kernels this call launched, counted at the launch sites
            TRACE.append(self.rec)
        return False


def launch_count():
    """Kernels launched by this library in this process so far (counted at the launch sites)."""
    return int(lib().stemgnn_launch_count())
