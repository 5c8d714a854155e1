"""ctypes binding of libq4k_b200.so — one Python function per entry of include/q4k_b200.h.

Tensors are torch CUDA tensors: Q4_K / Q8_K data as uint8 holding the reference's block bytes
(L/ggml/src/ggml-common.h:291-306, 338-344), activations / results as float32.  Shapes follow ggml:
weights [M, K/256*144] (M rows of K), activations [N, K], result [N, M] (M fastest).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

QK_K, Q4K_BLOCK, Q8K_BLOCK = 256, 144, 292
Q6K_BLOCK = 210
_HERE = os.path.dirname(os.path.abspath(__file__))
# Q4K_B200_LIB: another build of the same library (A/B comparisons of kernels on one GPU box); default = the in-tree build
_LIB_PATH = os.environ.get("Q4K_B200_LIB") or os.path.normpath(os.path.join(_HERE, "..", "lib", "libq4k_b200.so"))
_lib = None

_ERR = {-1: "bad shape", -2: "bad pointer", -3: "workspace too small", -4: "CUDA error", -5: "unsupported",
        -6: "beyond the static scratch limit of q4k_mul_mat_noalloc"}


class Q4KError(RuntimeError):
    def __init__(self, fn: str, rc: int, cuda: int = 0):
        self.rc, self.cuda = rc, cuda
        super().__init__(f"{fn} failed: {_ERR.get(rc, rc)} (rc={rc}, cudaError={cuda})")


def lib_path() -> str:
    return _LIB_PATH


def load():
    """Loads the CUDA library.  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(nvcc, sm_100a).  q4k_b200 has no CPU fallback.")
        L = C.CDLL(_LIB_PATH)
        vp, i64, i32, sz = C.c_void_p, C.c_int64, C.c_int, C.c_size_t
        L.q4k_abi_version.restype = i32
        L.q4k_build_info.restype = C.c_char_p
        L.q4k_device_ok.argtypes = [i32]
        L.q4k_kernel_launches.restype = i64
        L.q4k_quantize_q8_K.argtypes = [vp, vp, i64, i64, i64, vp]
        L.q4k_mul_mat_workspace_bytes.argtypes = [i64, i64]
        L.q4k_mul_mat_workspace_bytes.restype = sz
        L.q4k_mul_mat.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, sz, vp]
        L.q4k_mul_mat_noalloc.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp]
        L.q4k_mul_mat_reuse.argtypes = [vp, i64, i64, i64, i64, vp, i64, vp, sz, vp]
        L.q4k_mul_mat_q8.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp]
        L.q4k_vec_dot_q4_K_q8_K.argtypes = [i32, vp, sz, vp, sz, vp, sz, i32]
        L.q4k_mul_mat_dump.argtypes = [i32, vp, i64, i64, i64, vp, i64, i64, vp, vp, vp, vp, vp, sz, vp]
        L.q4k_mul_mat_host.argtypes = [vp, i64, i64, i64, vp, i64, vp, vp, vp, vp, sz, vp]
        L.q4k_mul_mat_host_async.argtypes = [vp, i64, i64, i64, vp, i64, vp, vp, vp, vp, sz, vp]
        L.q4k_mul_mat_peers.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i32, i64, i64, vp]
        L.q4k_peer_signal_wait.argtypes = [vp, i32, i32, vp, vp]
        L.q4k_mul_mat_peers_ll.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i32, i32, i64, i64, i64, vp, i64, vp, vp, sz, vp]
        L.q4k_ll_buffer_bytes.argtypes = [i64, i64]
        L.q4k_ll_buffer_bytes.restype = sz
        L.q4k_mul_mat_peers_ws.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i32, i64, i64, vp, sz, vp]
        L.q4k_mul_mat_peers_sync.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i32, i64, i64, vp, sz, vp, i32, i32, vp, vp, vp]
        L.q4k_mul_mat_group.argtypes = [i32, vp, vp, i64, i64, vp, i64, i64, vp, vp]
        L.q4k_mul_mat_group_ws.argtypes = [i32, vp, vp, i64, i64, vp, i64, i64, vp, vp, sz, vp]
        L.q4k_plan_create.argtypes = [i32]
        L.q4k_plan_create.restype = vp
        L.q4k_plan_add_mul_mat.argtypes = [vp, vp, i64, i64, i64, i64, i32]
        L.q4k_plan_add_mul_mat_q6k.argtypes = [vp, vp, i64, i64, i64, i64, i32]
        L.q4k_plan_finalize.argtypes = [vp]
        L.q4k_plan_input_bytes.argtypes = [vp]
        L.q4k_plan_input_bytes.restype = sz
        L.q4k_plan_output_bytes.argtypes = [vp]
        L.q4k_plan_output_bytes.restype = sz
        L.q4k_plan_input_offset.argtypes = [vp, i32]
        L.q4k_plan_input_offset.restype = i64
        L.q4k_plan_output_offset.argtypes = [vp, i32]
        L.q4k_plan_output_offset.restype = i64
        L.q4k_plan_input_dev.argtypes = [vp]
        L.q4k_plan_input_dev.restype = vp
        L.q4k_plan_output_dev.argtypes = [vp]
        L.q4k_plan_output_dev.restype = vp
        L.q4k_plan_launches.argtypes = [vp]
        L.q4k_plan_launches.restype = i64
        L.q4k_plan_run.argtypes = [vp, vp]
        L.q4k_plan_run_host.argtypes = [vp, vp, vp, vp]
        L.q4k_plan_destroy.argtypes = [vp]
        L.q4k_plan_destroy.restype = None
        L.q6k_mul_mat_workspace_bytes.argtypes = [i64, i64]
        L.q6k_mul_mat_workspace_bytes.restype = sz
        L.q6k_mul_mat.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, sz, vp]
        L.q6k_mul_mat_q8.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp]
        L.q6k_debug_partials.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, vp, vp]
        L.q6k_gemm_workspace_bytes.argtypes = [i64, i64]
        L.q6k_gemm_workspace_bytes.restype = sz
        L.q6k_mul_mat_dump.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, vp, vp, sz, vp]
        L.q4k_op_rms_norm.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_float, vp]
        L.q4k_op_binary.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        L.q4k_op_swiglu.argtypes = [vp, i64, vp, i64, vp, i64, i64, i64, vp]
        L.q4k_op_get_rows_f32.argtypes = [vp, i64, i64, vp, i64, vp, i64, i64, vp]
        f32 = C.c_float
        L.q4k_op_rope.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, f32, f32, f32, f32, f32, f32, vp]
        L.q4k_debug_set.argtypes = [C.c_char_p, i64]
        L.q4k_debug_reset.restype = None
        L.q4k_debug_set_fuse_limit.argtypes = [i64]
        L.q4k_debug_set_pdl.argtypes = [i32]
        L.q4k_debug_set_gemv_order.argtypes = [i32]
        L.q4k_debug_set_big_matrix_bytes.argtypes = [i64]
        L.q4k_debug_set_gemv_config.argtypes = [i32]
        _lib = L
    return _lib


def _check(fn: str, rc: int):
    if rc != 0:
        raise Q4KError(fn, rc, load().q4k_last_cuda_error())


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dev_u8(t: torch.Tensor, name: str):
    if not (t.is_cuda and t.dtype == torch.uint8 and t.is_contiguous()):
        raise TypeError(f"{name}: expected a contiguous CUDA uint8 tensor")


def abi_version() -> int:
    return load().q4k_abi_version()


def build_info() -> str:
    return load().q4k_build_info().decode()


def kernel_launches() -> int:
    return load().q4k_kernel_launches()


def debug_set(name: str, value: int) -> None:
    """Development / test knob of the calling thread (q4k_debug_set of include/q4k_b200.h); 0 = the launch heuristic chooses."""
    _check("q4k_debug_set", load().q4k_debug_set(name.encode(), int(value)))


def debug_reset() -> None:
    load().q4k_debug_reset()


def workspace_bytes(K: int, N: int) -> int:
    return load().q4k_mul_mat_workspace_bytes(K, N)


def quantize_q8_K(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """float32 [N, K] (rows may be strided) -> uint8 [N, K/256*292]; bit-exact quantize_row_q8_K."""
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1):
        raise TypeError("x: expected a 2-D CUDA float32 tensor with unit inner stride")
    N, K = x.shape
    if K % QK_K:
        raise ValueError("K must be a multiple of 256")
    if out is None:
        out = torch.empty((N, K // QK_K * Q8K_BLOCK), dtype=torch.uint8, device=x.device)
    _dev_u8(out, "out")
    with torch.cuda.device(x.device):
        _check("q4k_quantize_q8_K", load().q4k_quantize_q8_K(x.data_ptr(), out.data_ptr(), K, N, x.stride(0) * 4 if N > 1 else K * 4, _stream()))
    return out


def mul_mat(w: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None,
            workspace: torch.Tensor | None = None) -> torch.Tensor:
    """ggml_mul_mat(Q4_K w [M, K/256*144], F32 x [N, K]) -> float32 [N, M]."""
    _dev_u8(w, "w")
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.is_contiguous()):
        raise TypeError("x: expected a contiguous 2-D CUDA float32 tensor")
    M, N, K = w.shape[0], x.shape[0], x.shape[1]
    if K % QK_K or w.shape[1] != K // QK_K * Q4K_BLOCK:
        raise ValueError("shape mismatch between w and x")
    if out is None:
        out = torch.empty((N, M), dtype=torch.float32, device=x.device)
    need = workspace_bytes(K, N)
    if workspace is None:
        workspace = torch.empty(max(need, 16), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        _check("q4k_mul_mat", load().q4k_mul_mat(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4,
                                                 out.data_ptr(), out.stride(0) * 4 if N > 1 else M * 4,
                                                 workspace.data_ptr(), workspace.numel(), _stream()))
    return out


def mul_mat_q8(w: torch.Tensor, xq: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """Same with src1 already Q8_K: xq uint8 [N, K/256*292]."""
    _dev_u8(w, "w")
    _dev_u8(xq, "xq")
    M, N = w.shape[0], xq.shape[0]
    nb = w.shape[1] // Q4K_BLOCK
    if xq.shape[1] != nb * Q8K_BLOCK:
        raise ValueError("shape mismatch between w and xq")
    if out is None:
        out = torch.empty((N, M), dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        _check("q4k_mul_mat_q8", load().q4k_mul_mat_q8(w.data_ptr(), M, nb * QK_K, w.stride(0), xq.data_ptr(), N, xq.stride(0),
                                                       out.data_ptr(), out.stride(0) * 4 if N > 1 else M * 4, _stream()))
    return out


def vec_dot_q4_K_q8_K(n: int, s: torch.Tensor, bs: int, vx: torch.Tensor, bx: int, vy: torch.Tensor, by: int, nrc: int) -> None:
    """ggml_vec_dot_t over device tensors (same argument meaning as the reference's; nrc must be 1)."""
    with torch.cuda.device(vx.device):
        torch.cuda.current_stream().synchronize()
        _check("q4k_vec_dot_q4_K_q8_K", load().q4k_vec_dot_q4_K_q8_K(n, s.data_ptr(), bs, vx.data_ptr(), bx, vy.data_ptr(), by, nrc))


DUMP_DISPATCH, DUMP_TENSOR_PIPE, DUMP_BULK, DUMP_BULK_EXACT, DUMP_BIG, DUMP_GEMM, DUMP_WIDE = range(7)
MIDN_MAX = 32          # largest N the batched decode kernel serves by default (knob midn_max); the GEMM takes over above


def mul_mat_dump(w: torch.Tensor, x: torch.Tensor, kernel: int = DUMP_DISPATCH):
    """q4k_mul_mat through the DUMP instantiation of a PRODUCT kernel (q4k_mul_mat_dump).
    -> (dst [N, M] f32, p [N, M, nb, 8] int32 or None for the GEMM, sumi [N, M, nb], summ [N, M, nb]).
    The integer buffers are pre-filled with a sentinel, so anything the kernel did not write shows up."""
    _dev_u8(w, "w")
    M, nb = w.shape[0], w.shape[1] // Q4K_BLOCK
    N, K = x.shape
    SENT = -0x7f7f7f7f
    gemm = kernel == DUMP_GEMM or (kernel == DUMP_DISPATCH and N > MIDN_MAX)
    p = None if gemm else torch.full((N, M, nb, 8), SENT, dtype=torch.int32, device=w.device)
    si = torch.full((N, M, nb), SENT, dtype=torch.int32, device=w.device)
    sm = torch.full((N, M, nb), SENT, dtype=torch.int32, device=w.device)
    dst = torch.empty((N, M), dtype=torch.float32, device=w.device)
    ws = torch.empty(max(workspace_bytes(K, N), 16), dtype=torch.uint8, device=w.device)
    with torch.cuda.device(w.device):
        _check("q4k_mul_mat_dump", load().q4k_mul_mat_dump(kernel, w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, x.stride(0) * 4, dst.data_ptr(),
                                                           p.data_ptr() if p is not None else 0, si.data_ptr(), sm.data_ptr(),
                                                           ws.data_ptr(), ws.numel(), _stream()))
    return dst, p, si, sm


def mul_mat_q6k(w: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None,
                workspace: torch.Tensor | None = None) -> torch.Tensor:
    """ggml_mul_mat(Q6_K w [M, K/256*210], F32 x [N, K]) -> float32 [N, M]  (include/q4k_b200.h: q6k_mul_mat)."""
    _dev_u8(w, "w")
    if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.is_contiguous()):
        raise TypeError("x: expected a contiguous 2-D CUDA float32 tensor")
    M, N, K = w.shape[0], x.shape[0], x.shape[1]
    if K % QK_K or w.shape[1] != K // QK_K * Q6K_BLOCK:
        raise ValueError("shape mismatch between w and x")
    if out is None:
        out = torch.empty((N, M), dtype=torch.float32, device=x.device)
    if workspace is None:
        workspace = torch.empty(max(load().q6k_mul_mat_workspace_bytes(K, N), 16), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        _check("q6k_mul_mat", load().q6k_mul_mat(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4,
                                                 out.data_ptr(), out.stride(0) * 4 if N > 1 else M * 4,
                                                 workspace.data_ptr(), workspace.numel(), _stream()))
    return out


def mul_mat_q6k_q8(w: torch.Tensor, xq: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """Same with src1 already Q8_K: xq uint8 [N, K/256*292]."""
    _dev_u8(w, "w")
    _dev_u8(xq, "xq")
    M, N = w.shape[0], xq.shape[0]
    nb = w.shape[1] // Q6K_BLOCK
    if xq.shape[1] != nb * Q8K_BLOCK:
        raise ValueError("shape mismatch between w and xq")
    if out is None:
        out = torch.empty((N, M), dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        _check("q6k_mul_mat_q8", load().q6k_mul_mat_q8(w.data_ptr(), M, nb * QK_K, w.stride(0), xq.data_ptr(), N, xq.stride(0),
                                                       out.data_ptr(), out.stride(0) * 4 if N > 1 else M * 4, _stream()))
    return out


def debug_partials_q6k(w: torch.Tensor, xq: torch.Tensor):
    """-> (sumi int32 [N, M, nb], y float32 [N, M]) from the product kernel with its integer dump on."""
    _dev_u8(w, "w")
    _dev_u8(xq, "xq")
    M, N = w.shape[0], xq.shape[0]
    nb = w.shape[1] // Q6K_BLOCK
    si = torch.zeros((N, M, nb), dtype=torch.int32, device=w.device)
    y = torch.zeros((N, M), dtype=torch.float32, device=w.device)
    with torch.cuda.device(w.device):
        _check("q6k_debug_partials", load().q6k_debug_partials(w.data_ptr(), M, nb * QK_K, w.stride(0), xq.data_ptr(), N,
                                                               xq.stride(0), si.data_ptr(), y.data_ptr(), _stream()))
    return si, y


def mul_mat_q6k_dump(w: torch.Tensor, x: torch.Tensor):
    """The Q6_K prefill GEMM with its integer dump on (include/q4k_b200.h: q6k_mul_mat_dump): (sumi int32 [N, M, nb], y [N, M])."""
    M, N, K = w.shape[0], x.shape[0], x.shape[1]
    nb = K // QK_K
    si = torch.zeros((N, M, nb), dtype=torch.int32, device=x.device)
    y = torch.empty((N, M), dtype=torch.float32, device=x.device)
    nbytes = load().q6k_gemm_workspace_bytes(K, N)
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        _check("q6k_mul_mat_dump", load().q6k_mul_mat_dump(w.data_ptr(), M, K, w.stride(0), x.data_ptr(), N, K * 4, si.data_ptr(),
                                                           y.data_ptr(), ws.data_ptr(), nbytes, _stream()))
    return si, y


# ---- the cheap F32 ops around the mul_mats (include/q4k_b200.h: q4k_op_*).  Tensors are torch CUDA float32 with ggml's dim
# order REVERSED (torch shape [ne3, ne2, ne1, ne0], last dim dense), up to 4-D; strides may be arbitrary in the outer dims.
def _ggml_ne_nb(t: torch.Tensor):
    if not (t.is_cuda and t.dtype == torch.float32 and 1 <= t.dim() <= 4 and t.stride(-1) == 1):
        raise TypeError("expected a CUDA float32 tensor of 1..4 dims with a dense last dim")
    shape = [1] * (4 - t.dim()) + list(t.shape)
    stride = [0] * (4 - t.dim()) + list(t.stride())
    ne = (C.c_int64 * 4)(shape[3], shape[2], shape[1], shape[0])
    nb = (C.c_int64 * 3)(stride[2] * 4, stride[1] * 4, stride[0] * 4)
    return ne, nb


def op_rms_norm(x: torch.Tensor, eps: float, weight: torch.Tensor | None = None, residual: torch.Tensor | None = None,
                want_sum: bool = False):
    """GGML_OP_RMS_NORM per row of x; optional fused residual add in front (x + residual; want_sum also returns it) and
    multiply by a weight row behind.  -> y, or (y, x + residual)."""
    ne, x_nb = _ggml_ne_nb(x)
    y = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _, y_nb = _ggml_ne_nb(y)
    r_nb = s_nb = None
    ssum = None
    if residual is not None:
        if residual.shape != x.shape:
            raise ValueError("residual must have the shape of x")
        _, r_nb = _ggml_ne_nb(residual)
        if want_sum:
            ssum = torch.empty(x.shape, dtype=torch.float32, device=x.device)
            _, s_nb = _ggml_ne_nb(ssum)
    if weight is not None and not (weight.is_cuda and weight.dtype == torch.float32 and weight.is_contiguous() and weight.numel() == x.shape[-1]):
        raise TypeError("weight: expected a contiguous CUDA float32 row of x.shape[-1] elements")
    with torch.cuda.device(x.device):
        _check("q4k_op_rms_norm", load().q4k_op_rms_norm(x.data_ptr(), x_nb, residual.data_ptr() if residual is not None else None, r_nb,
                                                         weight.data_ptr() if weight is not None else None, y.data_ptr(), y_nb,
                                                         ssum.data_ptr() if ssum is not None else None, s_nb, ne, eps, _stream()))
    return (y, ssum) if want_sum else y


def op_binary(op: str, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """GGML_OP_ADD / GGML_OP_MUL with b repeated over a (ggml_can_repeat)."""
    a_ne, a_nb = _ggml_ne_nb(a)
    b_ne, b_nb = _ggml_ne_nb(b)
    d = torch.empty(a.shape, dtype=torch.float32, device=a.device)
    _, d_nb = _ggml_ne_nb(d)
    with torch.cuda.device(a.device):
        _check("q4k_op_binary", load().q4k_op_binary({"add": 0, "mul": 1}[op], a.data_ptr(), a_ne, a_nb, b.data_ptr(), b_ne, b_nb,
                                                     d.data_ptr(), d_nb, _stream()))
    return d


def op_swiglu(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    """GGML_OP_GLU / SWIGLU, split form: silu(gate) * up, 2-D [rows, nc] with a dense last dim."""
    if gate.shape != up.shape or gate.dim() != 2 or gate.stride(1) != 1 or up.stride(1) != 1:
        raise ValueError("gate / up: expected two 2-D tensors of one shape with dense rows")
    d = torch.empty(gate.shape, dtype=torch.float32, device=gate.device)
    with torch.cuda.device(gate.device):
        _check("q4k_op_swiglu", load().q4k_op_swiglu(gate.data_ptr(), gate.stride(0) * 4, up.data_ptr(), up.stride(0) * 4, d.data_ptr(),
                                                     d.stride(0) * 4, gate.shape[1], gate.shape[0], _stream()))
    return d


def op_get_rows(src: torch.Tensor, ids: torch.Tensor) -> torch.Tensor:
    """GGML_OP_GET_ROWS: F32 rows of src [rows, nc] picked by int32 ids."""
    if src.dim() != 2 or src.stride(1) != 1 or ids.dtype != torch.int32 or not ids.is_contiguous():
        raise ValueError("src: 2-D float32 with dense rows; ids: contiguous int32")
    d = torch.empty((ids.numel(), src.shape[1]), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        _check("q4k_op_get_rows_f32", load().q4k_op_get_rows_f32(src.data_ptr(), src.stride(0) * 4, src.shape[0], ids.data_ptr(), ids.numel(),
                                                                 d.data_ptr(), d.stride(0) * 4, src.shape[1], _stream()))
    return d


def op_rope(x: torch.Tensor, pos: torch.Tensor, n_dims: int, theta_scale: float, neox: bool = False, freq_scale: float = 1.0,
            ext_factor: float = 0.0, attn_factor: float = 1.0, corr_dims=(0.0, 0.0), freq_factors: torch.Tensor | None = None) -> torch.Tensor:
    """GGML_OP_ROPE (NORMAL / NEOX): x float32 [tokens, heads, head_dim] (or 4-D with a leading batch), pos int32 [tokens];
    theta_scale = powf(freq_base, -2.0f / n_dims) and the YaRN correction dims come from the caller (include/q4k_b200.h)."""
    ne, x_nb = _ggml_ne_nb(x)
    if pos.dtype != torch.int32 or not pos.is_contiguous() or pos.numel() != ne[2]:
        raise ValueError("pos: contiguous int32, one per token")
    d = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _, d_nb = _ggml_ne_nb(d)
    with torch.cuda.device(x.device):
        _check("q4k_op_rope", load().q4k_op_rope(x.data_ptr(), x_nb, pos.data_ptr(), freq_factors.data_ptr() if freq_factors is not None else None,
                                                 d.data_ptr(), d_nb, ne, n_dims, int(neox), theta_scale, freq_scale, ext_factor, attn_factor,
                                                 corr_dims[0], corr_dims[1], _stream()))
    return d


class HostPlan:
    """Device staging for the host-buffer entry (what the ggml backend plugin does per MUL_MAT node)."""

    def __init__(self, M: int, K: int, N: int, device="cuda"):
        self.M, self.K, self.N = M, K, N
        self.x_stage = torch.empty((N, K), dtype=torch.float32, device=device)
        self.dst_stage = torch.empty((N, M), dtype=torch.float32, device=device)
        self.ws = torch.empty(max(workspace_bytes(K, N), 16), dtype=torch.uint8, device=device)


def mul_mat_host(w: torch.Tensor, x_host: torch.Tensor, dst_host: torch.Tensor, plan: HostPlan) -> None:
    """x_host [N, K] / dst_host [N, M]: CPU float32 tensors (pinned for speed).  Synchronous."""
    _dev_u8(w, "w")
    with torch.cuda.device(w.device):
        _check("q4k_mul_mat_host", load().q4k_mul_mat_host(w.data_ptr(), plan.M, plan.K, w.stride(0), x_host.data_ptr(), plan.N,
                                                           dst_host.data_ptr(), plan.x_stage.data_ptr(), plan.dst_stage.data_ptr(),
                                                           plan.ws.data_ptr(), plan.ws.numel(), _stream()))


def mul_mat_group(ws, xs, outs=None):
    """Horizontally fused decode launch (q4k_mul_mat_group): up to 4 Q4_K matrices with the same K, F32 inputs xs[i]
    [N, K] (N <= 8; the same tensor may repeat) -> list of float32 [N, M_i]."""
    n = len(ws)
    if n != len(xs) or not 1 <= n <= 4:
        raise ValueError("1..4 matrices, one input each")
    N, K = xs[0].shape
    for w, x in zip(ws, xs):
        _dev_u8(w, "w")
        if not (x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and tuple(x.shape) == (N, K)):
            raise TypeError("xs: contiguous CUDA float32 [N, K], same shape for all")
        if w.shape[1] != K // QK_K * Q4K_BLOCK or w.stride(0) != ws[0].stride(0):
            raise ValueError("shape mismatch between w and x")
    if outs is None:
        outs = [torch.empty((N, w.shape[0]), dtype=torch.float32, device=w.device) for w in ws]
    vpa, i64a = C.c_void_p * n, C.c_int64 * n
    with torch.cuda.device(ws[0].device):
        _check("q4k_mul_mat_group", load().q4k_mul_mat_group(n, vpa(*[w.data_ptr() for w in ws]), i64a(*[w.shape[0] for w in ws]), K,
                                                             ws[0].stride(0), vpa(*[x.data_ptr() for x in xs]), N, K * 4,
                                                             vpa(*[o.data_ptr() for o in outs]), _stream()))
    return outs


def mul_mat_group_ws(ws, x, workspace=None):
    """q4k_mul_mat_group_ws: up to 4 Q4_K matrices with the same K on ONE activation tensor x [N, K], any N outside 9..32
    (decode group launch, or one quantizer pass + one GEMM launch over all row tiles) -> list of float32 [N, M_i]."""
    n = len(ws)
    N, K = x.shape
    for w in ws:
        _dev_u8(w, "w")
        if w.shape[1] != K // QK_K * Q4K_BLOCK or w.stride(0) != ws[0].stride(0):
            raise ValueError("shape mismatch between w and x")
    outs = [torch.empty((N, w.shape[0]), dtype=torch.float32, device=w.device) for w in ws]
    if workspace is None:
        workspace = torch.empty(max(workspace_bytes(K, N), 16), dtype=torch.uint8, device=x.device)
    vpa, i64a = C.c_void_p * n, C.c_int64 * n
    with torch.cuda.device(ws[0].device):
        _check("q4k_mul_mat_group_ws", load().q4k_mul_mat_group_ws(n, vpa(*[w.data_ptr() for w in ws]), i64a(*[w.shape[0] for w in ws]), K,
                                                                   ws[0].stride(0), x.data_ptr(), N, K * 4, vpa(*[o.data_ptr() for o in outs]),
                                                                   workspace.data_ptr(), workspace.numel(), _stream()))
    return outs


class GraphPlan:
    """q4k_plan_*: a fixed sequence of Q4_K mul_mats recorded once and replayed as one CUDA graph, with inputs and
    outputs packed into two device buffers (one H2D + one D2H per replay for a host caller).

    add(w, N, input_id) -> index; finalize(); input(input_id) / output(index) are torch views of the plan's buffers."""

    def __init__(self, device=0):
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        self._L = load()
        self._p = self._L.q4k_plan_create(self.device.index or 0)
        if not self._p:
            raise Q4KError("q4k_plan_create", -5, self._L.q4k_last_cuda_error())
        self._keep, self._shapes, self._in_shapes = [], [], {}
        self.finalized = False

    def add(self, w: torch.Tensor, N: int, input_id: int, q6k: bool = False) -> int:
        """w: block_q4_K rows (or block_q6_K rows with q6k=True)."""
        _dev_u8(w, "w")
        M, K = w.shape[0], w.shape[1] // (Q6K_BLOCK if q6k else Q4K_BLOCK) * QK_K
        fn = self._L.q4k_plan_add_mul_mat_q6k if q6k else self._L.q4k_plan_add_mul_mat
        rc = fn(self._p, w.data_ptr(), M, K, w.stride(0), N, input_id)
        if rc < 0:
            raise Q4KError("q4k_plan_add_mul_mat", rc, self._L.q4k_last_cuda_error())
        self._keep.append(w)
        self._shapes.append((N, M))
        self._in_shapes[input_id] = (N, K)
        return rc

    def finalize(self):
        with torch.cuda.device(self.device):
            _check("q4k_plan_finalize", self._L.q4k_plan_finalize(self._p))
        self.finalized = True
        self.input_bytes = self._L.q4k_plan_input_bytes(self._p)
        self.output_bytes = self._L.q4k_plan_output_bytes(self._p)
        self.launches = self._L.q4k_plan_launches(self._p)
        self._in = _wrap_device(self._L.q4k_plan_input_dev(self._p), self.input_bytes, self.device)
        self._out = _wrap_device(self._L.q4k_plan_output_dev(self._p), self.output_bytes, self.device)
        return self

    def input_offset(self, input_id: int) -> int:
        return self._L.q4k_plan_input_offset(self._p, input_id)

    def output_offset(self, index: int) -> int:
        return self._L.q4k_plan_output_offset(self._p, index)

    def input(self, input_id: int) -> torch.Tensor:
        N, K = self._in_shapes[input_id]
        off = self.input_offset(input_id)
        return self._in[off: off + N * K * 4].view(torch.float32).view(N, K)

    def output(self, index: int) -> torch.Tensor:
        N, M = self._shapes[index]
        off = self.output_offset(index)
        return self._out[off: off + N * M * 4].view(torch.float32).view(N, M)

    def run(self):
        with torch.cuda.device(self.device):
            _check("q4k_plan_run", self._L.q4k_plan_run(self._p, _stream()))

    def run_host(self, in_host: torch.Tensor, out_host: torch.Tensor):
        """in_host / out_host: CPU uint8 tensors of input_bytes / output_bytes (pinned for async copies)."""
        if in_host.numel() * in_host.element_size() != self.input_bytes or out_host.numel() * out_host.element_size() != self.output_bytes:
            raise ValueError("host buffers must be exactly input_bytes / output_bytes")
        with torch.cuda.device(self.device):
            _check("q4k_plan_run_host", self._L.q4k_plan_run_host(self._p, in_host.data_ptr(), out_host.data_ptr(), _stream()))

    def close(self):
        if self._p:
            self._in = self._out = None
            self._L.q4k_plan_destroy(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _wrap_device(ptr: int, nbytes: int, device) -> torch.Tensor:
    """A uint8 torch view of device memory owned by the library (no copy, no ownership)."""
    class _Mem:
        pass
    m = _Mem()
    m.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}
    return torch.as_tensor(m, device=device)
