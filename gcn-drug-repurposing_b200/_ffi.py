"""ctypes binding of libgdr.so (the C ABI declared in include/gdr.h).

There is no CPU fallback and no alternative backend: if the shared library is missing
or a call fails, this module raises.  Every wrapper passes raw device pointers, sizes
and the caller's current CUDA stream, and turns a non-zero return code into
``RuntimeError(gdr_last_error_string())``.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libgdr.so")

c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int
c_f32 = ctypes.c_float
c_f64 = ctypes.c_double
c_ptr = ctypes.c_void_p

# name -> (restype, [argtypes]) ; mirrors include/gdr.h one to one
_SIGNATURES = {
    "gdr_version": (c_i32, []),
    "gdr_last_error_string": (ctypes.c_char_p, []),
    "gdr_launch_count": (c_i64, []),
    "gdr_launch_count_reset": (None, []),
    "gdr_csr_from_coo_workspace_bytes": (c_i64, [c_i64, c_i64, c_i64]),
    "gdr_csr_from_coo": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_permute_values": (c_i32, [c_ptr, c_ptr, c_i64, c_i32, c_ptr, c_ptr]),
    "gdr_csr_transpose_workspace_bytes": (c_i64, [c_i64, c_i64, c_i64]),
    "gdr_csr_transpose": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_csr_add_identity_count": (c_i32, [c_ptr, c_ptr, c_i64, c_ptr, ctypes.POINTER(c_i64), c_ptr, c_i64, c_ptr]),
    "gdr_csr_add_identity_workspace_bytes": (c_i64, [c_i64]),
    "gdr_csr_normalize_f64": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gdr_spmm_csr_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gdr_spmm_csr_planned_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32,
                                         c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gdr_spmm_csr_dual_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32,
                                      c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gdr_spmm_csr_f64": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "gdr_gcn_layer_fwd_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_f32, c_i64, c_i64,
                                      c_ptr, c_ptr, c_ptr]),
    "gdr_row_normalize_fwd_f32": (c_i32, [c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "gdr_row_normalize_bwd_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr]),
    "gdr_elu_bwd_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_f32, c_i64, c_ptr, c_ptr]),
    "gdr_linear_bwd_input_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr]),
    "gdr_linear_bwd_weight_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_linear_bwd_weight_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_gss_loss_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_gss_loss_fwd_bwd_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_f32, c_f32, c_ptr, c_ptr, c_ptr,
                                         c_i64, c_ptr]),
    "gdr_gss_loss_fwd_bwd_static_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_f32, c_f32, c_ptr, c_ptr, c_ptr]),
    "gdr_gemm_nt_tc_workspace_bytes": (c_i64, [c_i64, c_i64, c_i64]),
    "gdr_gemm_nt_tc_f32": (c_i32, [c_ptr, c_i64, c_ptr, c_i64, c_i64, c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_ptr]),
    "gdr_gcn_layer_fwd_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_gcn_layer_fwd_tc_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_f32, c_i64, c_i64, c_ptr,
                                         c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_linear_bwd_input_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_linear_bwd_input_tc_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_linear_bwd_weight_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_linear_bwd_weight_tc_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_gss_loss_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_gss_loss_fwd_bwd_tc_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_f32, c_f32, c_ptr, c_ptr, c_ptr,
                                            c_i64, c_ptr]),
    "gdr_gram_percentile_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_gram_percentile_tc_f32": (c_i32, [c_ptr, c_i64, c_i64, c_f64, ctypes.POINTER(c_f32), c_ptr, c_i64, c_ptr]),
    "gdr_linear_bwd_weight2_tc_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_linear_bwd_weight2_tc_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                              c_i64, c_ptr]),
    "gdr_gather_rows_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr]),
    "gdr_scatter_rows_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "gdr_gram_percentile_workspace_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_gram_percentile_f32": (c_i32, [c_ptr, c_i64, c_i64, c_f64, ctypes.POINTER(c_f32), c_ptr, c_i64, c_ptr]),
    "gdr_gram_topk_f64": (c_i32, [c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "gdr_gram_topk_tc_workspace_bytes": (c_i64, [c_i64, c_i64, c_i64]),
    "gdr_gram_topk_tc_f64": (c_i32, [c_ptr, c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_bias_act_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i32, c_ptr, c_ptr]),
    "gdr_act_bwd_colsum_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr]),
    "gdr_dropout_apply_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_f32, c_ptr, c_ptr, c_ptr]),
    "gdr_masked_softmax_ce_workspace_bytes": (c_i64, [c_i64]),
    "gdr_masked_softmax_ce_f32": (c_i32, [c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64,
                                          c_ptr]),
    "gdr_walk_row_cumsum_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_ptr, c_ptr]),
    "gdr_node2vec_walks": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_i64, c_f64, c_f64, c_i32, ctypes.c_uint64,
                                   c_i64, c_ptr, c_ptr]),
    "gdr_text_savetxt_bytes": (c_i64, [c_i64, c_i64]),
    "gdr_text_savetxt_e18": (c_i64, [c_ptr, c_i64, c_i64, c_i32, ctypes.c_char_p, c_i64, c_i32]),
    "gdr_text_parse_embs": (c_i64, [ctypes.c_char_p, c_i64, c_i32, c_ptr, c_i64, c_ptr, c_ptr,
                                    ctypes.POINTER(c_i64), c_i32]),
    "gdr_mc_broadcast_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i32, c_ptr]),
    "gdr_mc_reduce_add_f32": (c_i32, [c_ptr, c_ptr, c_i64, c_i32, c_ptr]),
    "gdr_push_rows_f32": (c_i32, [c_ptr, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_i32, c_ptr]),
    "gdr_adam_step_multi_f32": (c_i32, [c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_f64, c_f64, c_f64, c_f64,
                                        c_ptr]),
    "gdr_adam_step_f32": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_f32, c_f32, c_f32, c_f32, c_ptr, c_ptr]),
    "gdr_rwr_workspace_bytes": (c_i64, [c_i64, c_i64, c_i64]),
    "gdr_rwr_solve_f64": (c_i32, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i64,
                                  c_f64, c_f64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_rwr_refine_f64": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "gdr_rwr_power_iteration_workspace_bytes": (c_i64, [c_i64]),
    "gdr_rwr_power_iteration_f64": (c_i32, [c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_f64, c_f64, c_i64, c_ptr,
                                            ctypes.POINTER(c_i32), c_ptr, c_i64, c_ptr]),
    "gdr_rank_workspace_bytes": (c_i64, [c_i64]),
    "gdr_rank_embedding_f64": (c_i32, [c_ptr, c_i64, c_i64, c_i32, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr,
                                       c_i64, c_ptr]),
    "gdr_rank_values_f64": (c_i32, [c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_ptr]),
    "gdr_text_format_edgelist": (c_i64, [ctypes.c_char_p, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_i64,
                                         c_i32]),
    "gdr_text_parse_edgelist": (c_i64, [ctypes.c_char_p, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def lib():
    """Load libgdr.so once.  Raises if the extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA extension is the only implementation of this "
                "path (no CPU fallback). Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C gcn-drug-repurposing_b200/csrc`.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the header and the library disagree
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().gdr_last_error_string().decode("utf-8", "replace")
        raise RuntimeError(f"libgdr {what} failed (rc={rc}): {msg}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL).  CPU tensors are an error, not a slow path."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libgdr: expected a CUDA tensor (this path has no CPU implementation)")
    if not t.is_contiguous():
        raise RuntimeError("libgdr: expected a contiguous tensor")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


def launch_count():
    return int(lib().gdr_launch_count())


def launch_count_reset():
    lib().gdr_launch_count_reset()
