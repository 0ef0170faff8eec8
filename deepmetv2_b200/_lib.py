"""ctypes binding of csrc/libgraphmet.so (the C ABI declared in include/graphmet.h).

There is NO fallback: if the shared library is missing the import of any op raises, telling the
user how to build it (``python -c 'import __graft_entry__ as g; g.build()'`` or ``make -C
deepmetv2_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.environ.get("GRAPHMET_SO") or os.path.join(CSRC, "libgraphmet.so")   # override: kernel-variant experiments
HEADER = os.path.join(os.path.dirname(_HERE), "include", "graphmet.h")

GM_OK = 0
FLAG_CAT_OUT_OF_RANGE = 1
FLAG_DEGREE_CAPPED = 2
FLAG_EDGE_OVERFLOW = 4
FLAG_EVENT_TOO_LARGE = 8
MAX_CONV_DEPTH = 4
AGGR = {"max": 0, "add": 1, "mean": 2}
PRECISION = {"fp32": 0, "tf32": 1, "bf16": 2, "tf32x3": 3, "fp32_direct": 4}
TIMERS = {"radius_graph": 0, "encode_fwd": 1, "edgeconv_fwd": 2, "head_fwd": 3, "scatter_s_bwd": 4,
          "conv_node_bwd": 5, "encode_bwd": 6, "bn": 7, "node_prep": 8, "head_bwd": 9,
          "drn_node": 10, "drn_knn": 11, "drn_graph": 12, "drn_edge": 13, "drn_pool": 14, "drn_edge_bwd": 15, "drn_bwd": 16}

_vp, _i, _ll, _f, _d, _sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double, C.c_size_t


class GmBatchNorm(C.Structure):
    _fields_ = [("weight", _vp), ("bias", _vp), ("running_mean", _vp), ("running_var", _vp),
                ("num_batches_tracked", _vp)]


class GmParams(C.Structure):
    _fields_ = [("embed_charge", _vp), ("embed_pdgid", _vp), ("embed_pv", _vp),
                ("cont_w", _vp), ("cont_b", _vp), ("cat_w", _vp), ("cat_b", _vp),
                ("enc_w", _vp), ("enc_b", _vp), ("bn_all", GmBatchNorm),
                ("conv_w", _vp * MAX_CONV_DEPTH), ("conv_b", _vp * MAX_CONV_DEPTH),
                ("conv_bn", GmBatchNorm * MAX_CONV_DEPTH),
                ("out1_w", _vp), ("out1_b", _vp), ("out2_w", _vp), ("out2_b", _vp),
                ("continuous_dim", _i), ("conv_depth", _i)]


class GmAdamW(C.Structure):
    _fields_ = [("params", _vp), ("grads", _vp), ("exp_avg", _vp), ("exp_avg_sq", _vp), ("n", _i),
                ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("weight_decay", C.c_float), ("step", _i), ("grad_scale", C.c_float),
                ("err_flags_dev", _vp), ("graph_stats_dev", _vp), ("status_host", _vp)]


class GmDrnConv(C.Structure):
    _fields_ = [("w1", _vp), ("b1", _vp), ("w2", _vp), ("b2", _vp), ("bn_weight", _vp), ("bn_bias", _vp),
                ("bn_running_mean", _vp), ("bn_running_var", _vp), ("bn_num_batches_tracked", _vp)]


class GmDrnParams(C.Structure):
    _fields_ = [("datanorm", _vp), ("in_w", _vp * 3), ("in_b", _vp * 3), ("conv", GmDrnConv * 2),
                ("out_w", _vp * 3), ("out_b", _vp * 3), ("output_dim", _i), ("k", _i)]


class GmDrnView(C.Structure):
    _fields_ = [("x0", _vp), ("rowptr1", _vp), ("col1", _vp), ("row1", _vp), ("x1", _vp), ("weight1", _vp),
                ("cluster1", _vp), ("partner1", _vp), ("x1p", _vp), ("ptr1p", _vp), ("rowptr2", _vp), ("col2", _vp),
                ("row2", _vp), ("x2", _vp), ("pooled", _vp), ("edge_capacity", _ll), ("asel1", _vp), ("asel2", _vp)]


# name -> (restype, argtypes); must list every symbol include/graphmet.h declares
SIGNATURES = {
    "gm_version": (_i, []),
    "gm_status_string": (C.c_char_p, [_i]),
    "gm_last_cuda_error": (_i, []),
    "gm_num_sms": (_i, []),
    "gm_launch_count": (_ll, [_i]),
    "gm_timing_enable": (_i, [_i]),
    "gm_timing_read": (_i, [_i, _vp, _vp]),
    "gm_bench_smem_read": (_i, [_i, _vp, _vp]),
    "gm_bench_ffma": (_i, [_i, _vp, _vp]),
    "gm_bench_ffma2": (_i, [_i, _vp, _vp]),
    "gm_prepare_etaphi": (_i, [_vp, _i, _i, _vp, _vp]),
    "gm_batch_to_ptr": (_i, [_vp, _i, _i, _vp, _vp]),
    "gm_radius_graph_ws_bytes": (_sz, [_i, _i]),
    "gm_radius_graph_count": (_i, [_vp, _i, _i, _vp, _i, _i, _d, _i, _i, _i, _i, _vp, _sz, _vp, _vp, _vp]),
    "gm_radius_graph_fill": (_i, [_vp, _i, _i, _vp, _i, _i, _d, _i, _i, _i, _i, _vp, _sz, _vp, _vp, _ll, _vp,
                                  _vp, _vp]),
    "gm_radius_graph_static": (_i, [_vp, _i, _i, _vp, _i, _i, _d, _i, _i, _i, _i, _vp, _vp, _ll, _vp, _vp]),
    "gm_knn_graph_ws_bytes": (_sz, [_i, _i, _i]),
    "gm_knn_graph_count": (_i, [_vp, _i, _i, _vp, _i, _i, _i, _i, _vp, _sz, _vp, _vp]),
    "gm_knn_graph_fill": (_i, [_i, _i, _i, _vp, _sz, _vp, _vp, _ll, _vp, _vp]),
    "gm_csr_ws_bytes": (_sz, [_i, _ll]),
    "gm_edge_index_to_csr": (_i, [_vp, _ll, _i, _vp, _sz, _vp, _vp, _vp]),
    "gm_csr_transpose": (_i, [_vp, _vp, _i, _ll, _vp, _sz, _vp, _vp, _vp, _vp]),
    "gm_edgeconv_ws_bytes": (_sz, [_i]),
    "gm_edgeconv_forward": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _sz, _vp, _vp, _vp]),
    "gm_net_ws_bytes": (_sz, [_i, _i, _i]),
    "gm_net_forward": (_i, [C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _f,
                            _vp, _sz, _vp, _vp, _vp]),
    "gm_net_bwd_ws_bytes": (_sz, [_i, _i]),
    "gm_net_backward": (_i, [C.POINTER(GmParams), C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp,
                             _i, _vp, _i, _i, _i, _f, _vp, _sz, _vp, _sz, _vp, _vp, _vp]),
    "gm_net_knn_ws_bytes": (_sz, [_i, _i, _i]),
    "gm_net_forward_knn": (_i, [C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _i, _i, _i, _i, _i, _i, _i, _f, _f, _vp, _sz,
                                _vp, _sz, _vp, _vp, _vp]),
    "gm_net_backward_knn": (_i, [C.POINTER(GmParams), C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _sz,
                                 _vp, _sz, _vp, _sz, _vp, _vp, _vp]),
    "gm_geo_ws_bytes": (_sz, [_i, _i]),
    "gm_geo_build": (_i, [_vp, _i, _i, _vp, _i, _i, _d, _i, _i, _i, _i, _vp, _sz, _vp, _vp, _vp]),
    "gm_net_forward_geo": (_i, [C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _i, _i, _vp, _i, _i, _i, _i, _i, _f, _f,
                                _vp, _sz, _vp, _vp, _vp]),
    "gm_net_backward_geo": (_i, [C.POINTER(GmParams), C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _i, _i, _vp,
                                 _i, _i, _i, _vp, _sz, _vp, _sz, _vp, _vp, _vp]),
    "gm_net_backward_geo_adamw": (_i, [C.POINTER(GmParams), C.POINTER(GmParams), _vp, _i, _vp, _vp, _i, _vp, _i, _i, _vp,
                                       _i, _i, _i, _vp, _sz, _vp, _sz, _vp, _vp, C.POINTER(GmAdamW), _vp]),
    "gm_edgeconv_forward_geo": (_i, [_vp, _vp, _i, _i, _vp, _vp, _i, _i, _vp, _sz, _vp, _vp, _vp]),
    "gm_met_loss": (_i, [_vp, _vp, _i, _vp, _i, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "gm_segment_sum": (_i, [_vp, _vp, _i, _vp, _vp]),
    "gm_scatter_add": (_i, [_vp, _vp, _ll, _vp, _ll, _vp]),
    "gm_index_stats": (_i, [_vp, _ll, _vp, _vp]),
    "gm_resolution": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "gm_resolution_summary": (_i, [_vp, _ll, _vp, _i, _i, _f, _i, _vp, _vp]),
    "gm_adamw_step": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _f, _f, _f, _i, _f, _vp]),
    "gm_met_loss_ex": (_i, [_vp, _vp, _i, _vp, _i, _vp, _i, _i, _f, _vp, _vp, _vp, _vp]),
    "gm_adamw_step_ex": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _f, _f, _f, _i, _f, _vp, _vp, _vp, _vp, _vp]),
    "gm_comm_create": (_i, [_i, _i, _i, C.POINTER(_vp), _vp]),
    "gm_comm_connect": (_i, [_vp, _vp]),
    "gm_comm_destroy": (_i, [_vp]),
    "gm_undirected_ws_bytes": (_sz, [_i, _ll]),
    "gm_csr_to_undirected": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _sz, _vp, _vp, _vp, _ll, _vp, _vp]),
    "gm_normalized_cut": (_i, [_vp, _i, _i, _vp, _vp, _vp, _i, _ll, _vp, _vp]),
    "gm_graclus": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "gm_cluster_max_pool": (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _vp, _vp, _sz, _vp, _vp, _vp, _vp]),
    "gm_drn_ws_bytes": (_sz, [_i, _i, _i]),
    "gm_drn_forward": (_i, [C.POINTER(GmDrnParams), _vp, _i, _vp, _i, _i, _i, _i, _i, _f, _f, _vp, _sz, _vp, _vp,
                            C.POINTER(GmDrnView), _vp]),
    "gm_exchange_post": (_i, [_vp, _i, _f, _vp, _vp, _vp]),
    "gm_exchange_apply_adamw": (_i, [_vp, _vp, _vp, _i, _f, _f, _f, _f, _f, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "gm_drn_met_loss": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "gm_drn_bwd_ws_bytes": (_sz, [_i, _i, _i]),
    "gm_drn_backward": (_i, [C.POINTER(GmDrnParams), C.POINTER(GmDrnParams), _vp, _i, _vp, _i, _i, _vp, _sz, _vp, _sz, _vp, _vp]),
    "gm_allreduce_adamw_step": (_i, [_vp, _vp, _vp, _vp, _i, _f, _f, _f, _f, _f, _i, _f, _vp, _vp, _vp, _vp, _vp, _vp,
                                     _vp]),
}

_lib = None


class GraphMETLibraryError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into csrc/libgraphmet.so (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j", str(os.cpu_count() or 4)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise GraphMETLibraryError("building libgraphmet.so failed")
    return SO_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise GraphMETLibraryError(
                f"{SO_PATH} is missing: build it with `make -C {CSRC}` "
                "(or __graft_entry__.build()). There is no CPU / PyTorch fallback for this path.")
        l = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(status: int, what: str = "libgraphmet") -> None:
    if status != GM_OK:
        l = lib()
        msg = l.gm_status_string(status).decode()
        if status == -4:
            msg += f" (cudaError {l.gm_last_cuda_error()})"
        raise GraphMETLibraryError(f"{what}: {msg}")


def stream(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t: torch.Tensor | None) -> int | None:
    return None if t is None else t.data_ptr()


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise GraphMETLibraryError(
            f"{name} is on {t.device}: the GraphMET kernels are CUDA (sm_100a) only; there is no CPU fallback")
