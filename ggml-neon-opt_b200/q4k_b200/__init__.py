"""q4k_b200 — Python face of the B200-native Q4_K x Q8_K mul_mat path.

Thin ctypes binding of libq4k_b200.so (include/q4k_b200.h).  torch is used only for device memory and
streams.  There is no CPU or eager fallback: if the CUDA library is missing or no sm_100 GPU is
present, the ops raise.
"""
from .lib import (op_rms_norm, op_binary, op_swiglu, op_get_rows, op_rope, debug_reset, debug_set, Q4K_BLOCK, Q6K_BLOCK, debug_partials_q6k, mul_mat_q6k, mul_mat_q6k_q8, mul_mat_q6k_dump, Q8K_BLOCK, QK_K, GraphPlan, HostPlan, Q4KError, abi_version, build_info, mul_mat_dump, DUMP_DISPATCH, DUMP_TENSOR_PIPE, DUMP_BULK, DUMP_BULK_EXACT, DUMP_BIG, DUMP_GEMM, DUMP_WIDE,
                  kernel_launches, lib_path, load, mul_mat, mul_mat_group, mul_mat_group_ws, mul_mat_host, mul_mat_q8, quantize_q8_K,
                  vec_dot_q4_K_q8_K, workspace_bytes)

__all__ = ["op_rms_norm", "op_binary", "op_swiglu", "op_get_rows", "op_rope", "debug_reset", "debug_set", "Q6K_BLOCK", "debug_partials_q6k", "mul_mat_q6k", "mul_mat_q6k_q8", "mul_mat_q6k_dump", "GraphPlan", "HostPlan", "Q4K_BLOCK", "Q8K_BLOCK", "QK_K", "Q4KError", "abi_version", "build_info", "mul_mat_dump", "DUMP_DISPATCH", "DUMP_TENSOR_PIPE", "DUMP_BULK", "DUMP_BULK_EXACT", "DUMP_BIG", "DUMP_GEMM", "DUMP_WIDE",
           "kernel_launches", "lib_path", "load", "mul_mat", "mul_mat_group", "mul_mat_group_ws", "mul_mat_host", "mul_mat_q8", "quantize_q8_K",
           "vec_dot_q4_K_q8_K", "workspace_bytes"]
