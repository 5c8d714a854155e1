"""ctypes binding of oracle/q4k_oracle.c — the CPU ORACLE (test infrastructure, see q4k_oracle.h).

All arrays are numpy; Q4_K / Q8_K data are uint8 arrays holding the reference's block bytes
(144 B per 256 weights, 292 B per 256 activations — L/ggml/src/ggml-common.h:291-306, 338-344).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libq4k_oracle.so")

QK_K = 256
Q4K_BLOCK = 144
Q8K_BLOCK = 292
Q6K_BLOCK = 210
ORDER_GENERIC, ORDER_AVX2, ORDER_NEON = 0, 1, 2

_lib = None


def build() -> str:
    """Compile the restatement with gcc (seconds).  Building the checker is not using it."""
    src = [os.path.join(HERE, f) for f in ("q4k_oracle.c", "q6k_oracle.c", "q4k_oracle.h")]
    if os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in src):
        return LIB
    subprocess.check_call(["make", "-C", HERE, "oracle"], stdout=subprocess.DEVNULL)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp, i64, i32, f32p = C.c_void_p, C.c_int64, C.c_int, C.POINTER(C.c_float)
        L.ora_nearest_int.argtypes = [C.c_float]
        L.ora_nearest_int.restype = C.c_int
        L.ora_fp32_to_fp16.argtypes = [C.c_float]
        L.ora_fp32_to_fp16.restype = C.c_uint16
        L.ora_fp16_to_fp32.argtypes = [C.c_uint16]
        L.ora_fp16_to_fp32.restype = C.c_float
        L.ora_unpack_scales_mins.argtypes = [vp, vp, vp]
        L.ora_quantize_row_q8_K.argtypes = [vp, vp, i64]
        L.ora_quantize_row_q4_K.argtypes = [vp, vp, i64]
        L.ora_dequantize_row_q4_K.argtypes = [vp, vp, i64]
        L.ora_vec_dot_partials.argtypes = [i32, vp, vp, vp, vp, vp]
        for name in ("generic", "avx2order", "neonorder"):
            fn = getattr(L, f"ora_vec_dot_q4_K_q8_K_{name}")
            fn.argtypes = [i32, f32p, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, i32]
            fn.restype = None
        L.ora_vec_dot_q4_K_q8_K_f64.argtypes = [i32, vp, vp]
        L.ora_vec_dot_q4_K_q8_K_f64.restype = C.c_double
        L.ora_mul_mat_q4_K_f32.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, i32]
        L.ora_mul_mat_q4_K_f32.restype = i32
        L.ora_quantize_row_q6_K.argtypes = [vp, vp, i64]
        L.ora_dequantize_row_q6_K.argtypes = [vp, vp, i64]
        L.ora_vec_dot_q6_partials.argtypes = [i32, vp, vp, vp, vp]
        for name in ("generic", "avx2order"):
            fn = getattr(L, f"ora_vec_dot_q6_K_q8_K_{name}")
            fn.argtypes = [i32, f32p, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, i32]
            fn.restype = None
        L.ora_vec_dot_q6_K_q8_K_f64.argtypes = [i32, vp, vp]
        L.ora_vec_dot_q6_K_q8_K_f64.restype = C.c_double
        L.ora_mul_mat_q6_K_f32.argtypes = [vp, i64, i64, i64, vp, i64, i64, vp, i64, vp, i32]
        L.ora_mul_mat_q6_K_f32.restype = i32
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def nearest_int(f: float) -> int:
    return lib().ora_nearest_int(float(np.float32(f)))


def fp32_to_fp16_bits(f: float) -> int:
    return lib().ora_fp32_to_fp16(float(np.float32(f)))


def fp16_bits_to_fp32(h: int) -> float:
    return lib().ora_fp16_to_fp32(int(h))


def unpack_scales_mins(packed12: np.ndarray):
    packed12 = np.ascontiguousarray(packed12, dtype=np.uint8)
    sc = np.zeros(8, np.uint8)
    mn = np.zeros(8, np.uint8)
    lib().ora_unpack_scales_mins(_p(packed12), _p(sc), _p(mn))
    return sc, mn


def quantize_q8_K(x: np.ndarray, zero_init: bool = True) -> np.ndarray:
    """x: float32 [..., K] -> uint8 [..., K/256*292].  The reference leaves bsums of all-zero blocks
    unwritten (ggml-quants.c:2568-2573); zero_init makes those bytes 0."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    assert K % QK_K == 0
    rows = x.reshape(-1, K)
    out = (np.zeros if zero_init else np.empty)((rows.shape[0], K // QK_K * Q8K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        lib().ora_quantize_row_q8_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def quantize_q4_K(x: np.ndarray) -> np.ndarray:
    """x: float32 [M, K] -> uint8 [M, K/256*144] (quantize_row_q4_K_ref, no imatrix)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    assert K % QK_K == 0
    rows = x.reshape(-1, K)
    out = np.zeros((rows.shape[0], K // QK_K * Q4K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        lib().ora_quantize_row_q4_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def dequantize_q4_K(w: np.ndarray) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rows = w.reshape(-1, w.shape[-1])
    K = rows.shape[1] // Q4K_BLOCK * QK_K
    out = np.empty((rows.shape[0], K), np.float32)
    for r in range(rows.shape[0]):
        lib().ora_dequantize_row_q4_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*w.shape[:-1], K)


def vec_dot_partials(wrow: np.ndarray, yrow: np.ndarray):
    """-> (p[nb,8], sumi[nb], summ[nb]) int32: the bit-exact integer stage."""
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q4K_BLOCK
    assert yrow.size == nb * Q8K_BLOCK
    p = np.zeros((nb, 8), np.int32)
    si = np.zeros(nb, np.int32)
    sm = np.zeros(nb, np.int32)
    lib().ora_vec_dot_partials(nb * QK_K, _p(wrow), _p(yrow), _p(p), _p(si), _p(sm))
    return p, si, sm


def partials_matrix(w: np.ndarray, xq: np.ndarray):
    """The integer stage for every (column n, row m, superblock b) at once, in numpy (int64 arithmetic, int32 results):
    -> (p [N, M, nb, 8], sumi [N, M, nb], summ [N, M, nb]).  Same definitions as vec_dot_partials — the generic reference's
    aux32 / sumi / mins sum (L/ggml/src/ggml-cpu/quants.c:575-611) with get_scale_min_k4 (L/ggml/src/ggml-quants.c:703-710)."""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    xq = np.ascontiguousarray(xq, dtype=np.uint8)
    M, nb = w.shape[0], w.shape[1] // Q4K_BLOCK
    N = xq.shape[0]
    blk = w.reshape(M, nb, Q4K_BLOCK)
    s12 = blk[:, :, 4:16].astype(np.int64)
    sc = np.empty((M, nb, 8), np.int64); mn = np.empty((M, nb, 8), np.int64)
    sc[..., :4] = s12[..., 0:4] & 63
    mn[..., :4] = s12[..., 4:8] & 63
    sc[..., 4:] = (s12[..., 8:12] & 15) | ((s12[..., 0:4] >> 6) << 4)
    mn[..., 4:] = (s12[..., 8:12] >> 4) | ((s12[..., 4:8] >> 6) << 4)
    qs = blk[:, :, 16:].reshape(M, nb, 4, 32)
    q4 = np.stack([qs & 15, qs >> 4], axis=3).reshape(M, nb, 8, 32).astype(np.int32)      # sub-block 2g = low nibbles, 2g + 1 = high
    yb = xq.reshape(N, nb, Q8K_BLOCK)
    q8 = yb[:, :, 4:260].view(np.int8).reshape(N, nb, 8, 32).astype(np.int32)
    bs = yb[:, :, 260:292].copy().view(np.int16).reshape(N, nb, 8, 2).astype(np.int64).sum(3)   # bsums pairs = 32-element sums
    pj = np.einsum("mbjl,nbjl->nmbj", q4, q8, dtype=np.int64)
    sumi = (pj * sc[None]).sum(3)
    summ = np.einsum("mbj,nbj->nmb", mn, bs)
    return pj.astype(np.int32), sumi.astype(np.int32), summ.astype(np.int32)


def vec_dot(wrow: np.ndarray, yrow: np.ndarray, order: int = ORDER_AVX2) -> np.float32:
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q4K_BLOCK
    s = C.c_float(0)
    name = {ORDER_GENERIC: "generic", ORDER_AVX2: "avx2order", ORDER_NEON: "neonorder"}[order]
    getattr(lib(), f"ora_vec_dot_q4_K_q8_K_{name}")(nb * QK_K, C.byref(s), 0, _p(wrow), 0, _p(yrow), 0, 1)
    return np.float32(s.value)


def vec_dot_f64(wrow: np.ndarray, yrow: np.ndarray) -> float:
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q4K_BLOCK
    return lib().ora_vec_dot_q4_K_q8_K_f64(nb * QK_K, _p(wrow), _p(yrow))


def mul_mat(w: np.ndarray, x: np.ndarray, order: int = ORDER_AVX2) -> np.ndarray:
    """w: uint8 [M, K/256*144]; x: float32 [N, K] -> float32 [N, M] (ggml dst: ne0 = M fastest)."""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    M = w.shape[0]
    N, K = x.shape
    assert w.shape[1] == K // QK_K * Q4K_BLOCK
    dst = np.zeros((N, M), np.float32)
    wdata = np.zeros(N * (K // QK_K) * Q8K_BLOCK, np.uint8)
    rc = lib().ora_mul_mat_q4_K_f32(_p(w), M, K, w.shape[1], _p(x), N, K * 4, _p(dst), M * 4, _p(wdata), order)
    if rc != 0:
        raise ValueError(f"ora_mul_mat_q4_K_f32 rc={rc}")
    return dst


def mul_mat_f64(w: np.ndarray, x: np.ndarray) -> np.ndarray:
    """Double-precision combination of the exact integers: the tie-breaker for tolerance tests."""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    yq = quantize_q8_K(x)
    out = np.zeros((yq.shape[0], w.shape[0]), np.float64)
    for n in range(yq.shape[0]):
        for m in range(w.shape[0]):
            out[n, m] = vec_dot_f64(w[m], yq[n])
    return out


# ---------------------------------------------------------------------------------------- Q6_K (q6k_oracle.c)
def quantize_q6_K(x: np.ndarray) -> np.ndarray:
    """x: float32 [M, K] -> uint8 [M, K/256*210] (quantize_row_q6_K_ref, L/ggml/src/ggml-quants.c:1692-1760)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    assert K % QK_K == 0
    rows = x.reshape(-1, K)
    out = np.zeros((rows.shape[0], K // QK_K * Q6K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        lib().ora_quantize_row_q6_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def dequantize_q6_K(w: np.ndarray) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rows = w.reshape(-1, w.shape[-1])
    K = rows.shape[1] // Q6K_BLOCK * QK_K
    out = np.empty((rows.shape[0], K), np.float32)
    for r in range(rows.shape[0]):
        lib().ora_dequantize_row_q6_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*w.shape[:-1], K)


def vec_dot_q6_partials(wrow: np.ndarray, yrow: np.ndarray):
    """-> (p[nb,16], sumi[nb]) int32: the exact integer stage of ggml_vec_dot_q6_K_q8_K."""
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q6K_BLOCK
    assert yrow.size == nb * Q8K_BLOCK
    p = np.zeros((nb, 16), np.int32)
    si = np.zeros(nb, np.int32)
    lib().ora_vec_dot_q6_partials(nb * QK_K, _p(wrow), _p(yrow), _p(p), _p(si))
    return p, si


def vec_dot_q6(wrow: np.ndarray, yrow: np.ndarray, order: int = ORDER_AVX2) -> np.float32:
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q6K_BLOCK
    s = C.c_float(0)
    name = "avx2order" if order == ORDER_AVX2 else "generic"
    getattr(lib(), f"ora_vec_dot_q6_K_q8_K_{name}")(nb * QK_K, C.byref(s), 0, _p(wrow), 0, _p(yrow), 0, 1)
    return np.float32(s.value)


def mul_mat_q6(w: np.ndarray, x: np.ndarray, order: int = ORDER_AVX2) -> np.ndarray:
    """w: uint8 [M, K/256*210]; x: float32 [N, K] -> float32 [N, M]."""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    M = w.shape[0]
    N, K = x.shape
    assert w.shape[1] == K // QK_K * Q6K_BLOCK
    dst = np.zeros((N, M), np.float32)
    wdata = np.zeros(N * (K // QK_K) * Q8K_BLOCK, np.uint8)
    rc = lib().ora_mul_mat_q6_K_f32(_p(w), M, K, w.shape[1], _p(x), N, K * 4, _p(dst), M * 4, _p(wdata), order)
    if rc != 0:
        raise ValueError(f"ora_mul_mat_q6_K_f32 rc={rc}")
    return dst


def mul_mat_q6_f64(w: np.ndarray, x: np.ndarray) -> np.ndarray:
    w = np.ascontiguousarray(w, dtype=np.uint8)
    yq = quantize_q8_K(x)
    nb = w.shape[1] // Q6K_BLOCK
    out = np.zeros((yq.shape[0], w.shape[0]), np.float64)
    for n in range(yq.shape[0]):
        for m in range(w.shape[0]):
            out[n, m] = lib().ora_vec_dot_q6_K_q8_K_f64(nb * QK_K, _p(w[m]), _p(yq[n]))
    return out


def random_q6_K(rng, M: int, K: int) -> np.ndarray:
    """Random VALID block_q6_K bytes (any ql / qh / scales bytes are valid; d is a finite fp16)."""
    nb = K // QK_K
    w = rng.integers(0, 256, (M, nb, Q6K_BLOCK), dtype=np.uint8)
    d = (rng.uniform(0.0005, 0.01, (M, nb)) * rng.choice([-1.0, 1.0], (M, nb))).astype(np.float16)
    w[:, :, 208:210] = d.view(np.uint8).reshape(M, nb, 2)
    return w.reshape(M, nb * Q6K_BLOCK)


# ---- the cheap F32 ops around the mul_mats -------------------------------------------------------
def rms_norm(x: np.ndarray, eps: float) -> np.ndarray:
    """ggml_compute_forward_rms_norm_f32 (L/ggml/src/ggml-cpu/ops.cpp:3517-3570): squares rounded to float, summed in double
    in element order, mean and 1/sqrtf(mean + eps) in float, y = x * scale."""
    x = np.ascontiguousarray(x, np.float32)
    sq = (x * x).astype(np.float32).astype(np.float64)
    tot = np.zeros(x.shape[:-1], np.float64)
    for i in range(x.shape[-1]):                      # sequential, as the reference's loop
        tot += sq[..., i]
    mean = (tot / x.shape[-1]).astype(np.float32)
    scale = (np.float32(1.0) / np.sqrt((mean + np.float32(eps)).astype(np.float32))).astype(np.float32)
    return (x * scale[..., None]).astype(np.float32)


def swiglu(gate: np.ndarray, up: np.ndarray) -> np.ndarray:
    """ggml_compute_forward_swiglu_f32, split form (L/ggml/src/ggml-cpu/ops.cpp:2934-2990): silu(gate) * up with
    silu(x) = x / (1 + expf(-x)) (vec.h:957).  expf here is numpy's float32 exp; the reference's AVX2 path uses a polynomial
    (ggml_v_expf): agreement ~1e-6 relative, not bit-exact."""
    g = np.ascontiguousarray(gate, np.float32)
    with np.errstate(over="ignore"):                  # exp(90) = inf in float32: x / inf = -0, as expf gives the reference
        silu = (g / (np.float32(1.0) + np.exp(-g).astype(np.float32))).astype(np.float32)
    return (silu * np.ascontiguousarray(up, np.float32)).astype(np.float32)


def rope_theta_scale(freq_base: float, n_dims: int) -> np.float32:
    """powf(freq_base, -2.0f / n_dims) with the C library's powf, as ggml_compute_forward_rope_f32 derives it (ops.cpp:5584)."""
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.powf.restype = ctypes.c_float
    libm.powf.argtypes = [ctypes.c_float, ctypes.c_float]
    return np.float32(libm.powf(np.float32(freq_base), np.float32(-2.0) / np.float32(n_dims)))


def rope(x: np.ndarray, pos: np.ndarray, n_dims: int, neox: bool = False, freq_base: float = 10000.0, freq_scale: float = 1.0,
         attn_factor: float = 1.0) -> np.ndarray:
    """ggml_compute_forward_rope_f32 with ext_factor = 0 (L/ggml/src/ggml-cpu/ops.cpp:5436-5475, 5534-5700).
    x: float32 [tokens, heads, ne0].  theta is the reference's running float product; cos / sin in float64 rounded to float32
    (libm's cosf / sinf are within an ulp of that), the rotation in float32 without contraction."""
    x = np.ascontiguousarray(x, np.float32)
    y = x.copy()
    ts = rope_theta_scale(freq_base, n_dims)
    for t in range(x.shape[0]):
        theta = np.float32(pos[t])
        for i0 in range(0, n_dims, 2):
            th = np.float32(np.float32(freq_scale) * theta)
            c = np.float32(np.float32(np.cos(np.float64(th))) * np.float32(attn_factor))
            s = np.float32(np.float32(np.sin(np.float64(th))) * np.float32(attn_factor))
            a, b = (i0 // 2, i0 // 2 + n_dims // 2) if neox else (i0, i0 + 1)
            x0, x1 = x[t, :, a], x[t, :, b]
            y[t, :, a] = (x0 * c).astype(np.float32) - (x1 * s).astype(np.float32)
            y[t, :, b] = (x0 * s).astype(np.float32) + (x1 * c).astype(np.float32)
            theta = np.float32(theta * ts)
    return y
