"""ctypes binding of oracle/_ref/*.so — the UNMODIFIED reference ggml CPU code (test infrastructure).

oracle/_ref/ is built by `make -C oracle ref` in the build container (where /root/reference exists);
the built files travel to the GPU box with the snapshot.  Nothing here reads /root/reference at run
time.  `available()` is False when the prebuilt libraries are missing.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REFERENCE_ROOT = "/root/reference/external/llama.cpp"

QK_K, Q4K_BLOCK, Q8K_BLOCK = 256, 144, 292

_libs = None


def build() -> bool:
    """(Re)build oracle/_ref from the reference sources if they are present; returns available()."""
    if os.path.isdir(REFERENCE_ROOT):
        subprocess.check_call(["make", "-C", HERE, "-j8", "ref"], stdout=subprocess.DEVNULL)
    return available()


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, f))
               for f in ("libggml-base.so", "libggml-cpu.so", "libggml.so", "libref_shim.so"))


def libs():
    global _libs
    if _libs is None:
        if not available():
            raise RuntimeError("oracle/_ref is not built (run `make -C oracle ref` where /root/reference exists)")
        mode = C.RTLD_GLOBAL
        base = C.CDLL(os.path.join(REF_DIR, "libggml-base.so"), mode=mode)
        cpu = C.CDLL(os.path.join(REF_DIR, "libggml-cpu.so"), mode=mode)
        ggml = C.CDLL(os.path.join(REF_DIR, "libggml.so"), mode=mode)
        shim = C.CDLL(os.path.join(REF_DIR, "libref_shim.so"), mode=mode)
        vp, i64, i32 = C.c_void_p, C.c_int64, C.c_int
        base.quantize_row_q4_K_ref.argtypes = [vp, vp, i64]
        base.quantize_row_q8_K_ref.argtypes = [vp, vp, i64]
        base.dequantize_row_q4_K.argtypes = [vp, vp, i64]
        cpu.quantize_row_q8_K.argtypes = [vp, vp, i64]
        for fn in (cpu.ggml_vec_dot_q4_K_q8_K, cpu.ggml_vec_dot_q4_K_q8_K_generic):
            fn.argtypes = [i32, C.POINTER(C.c_float), C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, i32]
            fn.restype = None
        base.quantize_row_q6_K_ref.argtypes = [vp, vp, i64]
        base.dequantize_row_q6_K.argtypes = [vp, vp, i64]
        for fn in (cpu.ggml_vec_dot_q6_K_q8_K, cpu.ggml_vec_dot_q6_K_q8_K_generic):
            fn.argtypes = [i32, C.POINTER(C.c_float), C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, i32]
            fn.restype = None
        shim.ref_mul_mat_q4k.argtypes = [vp, i64, i64, vp, i64, vp, i32, i32, i32, C.POINTER(C.c_double)]
        shim.ref_mul_mat_q4k.restype = i32
        shim.ref_plugin_mul_mat_q4k.argtypes = [C.c_char_p, vp, i64, i64, vp, i64, vp, i32, C.POINTER(C.c_double)]
        shim.ref_plugin_mul_mat_q4k.restype = i32
        shim.ref_plugin_mul_mat_shared_q4k.argtypes = [C.c_char_p, i32, vp, C.POINTER(i64), i64, vp, i64, vp]
        shim.ref_plugin_mul_mat_shared_q4k.restype = i32
        shim.ref_cpu_has_avx2.restype = i32
        shim.ref_mul_mat_set_q4k.argtypes = [i32, C.POINTER(i64), C.POINTER(i64), i64, i32, i32, i32,
                                             C.POINTER(C.c_double), C.POINTER(C.c_double)]
        shim.ref_mul_mat_set_q4k.restype = i32
        shim.ref_mul_mat_set_typed.argtypes = [i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i32), i64, i32, i32, i32,
                                               C.POINTER(C.c_double), C.POINTER(C.c_double)]
        shim.ref_mul_mat_set_typed.restype = i32
        shim.ref_mul_mat_set_ex.argtypes = [i32, C.POINTER(i64), C.POINTER(i64), C.POINTER(i32), i64, i32, i32, i32, i32,
                                            C.POINTER(C.c_double), C.POINTER(C.c_double)]
        shim.ref_mul_mat_set_ex.restype = i32
        shim.ref_plugin_split_mul_mat.argtypes = [C.c_char_p, i32, vp, i64, i64, vp, i64, vp, vp, vp, C.POINTER(i32)]
        shim.ref_plugin_split_mul_mat.restype = i32
        shim.ref_cheap_ops.argtypes = [vp, vp, vp, vp, vp, i64, i64, C.c_float, i32, vp, vp, vp]
        shim.ref_cheap_ops.restype = i32
        f32 = C.c_float
        shim.ref_rope.argtypes = [vp, vp, i64, i64, i64, i32, i32, i32, f32, f32, f32, f32, f32, f32, i32, vp]
        shim.ref_rope.restype = i32
        # ggml-cpu's fp16->fp32 is a lookup table filled by ggml_cpu_init() (ggml-cpu.c:3526-3549);
        # the reference always reaches vec_dot through backend init, direct callers must do it themselves
        cpu.ggml_cpu_init()
        _libs = (base, cpu, ggml, shim)
    return _libs


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def quantize_q4_K(x: np.ndarray) -> np.ndarray:
    """quantize_row_q4_K_ref per row (L/ggml/src/ggml-quants.c:1280-1350)."""
    base = libs()[0]
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    rows = x.reshape(-1, K)
    out = np.zeros((rows.shape[0], K // QK_K * Q4K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        base.quantize_row_q4_K_ref(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def quantize_q8_K(x: np.ndarray) -> np.ndarray:
    """quantize_row_q8_K as ggml-cpu exports it (arch/x86/quants.c:493-495 -> ggml-quants.c:2555-2592).
    Output is zero-initialised: the reference leaves bsums of all-zero blocks unwritten."""
    cpu = libs()[1]
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    rows = x.reshape(-1, K)
    out = np.zeros((rows.shape[0], K // QK_K * Q8K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        cpu.quantize_row_q8_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def dequantize_q4_K(w: np.ndarray) -> np.ndarray:
    base = libs()[0]
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rows = w.reshape(-1, w.shape[-1])
    K = rows.shape[1] // Q4K_BLOCK * QK_K
    out = np.empty((rows.shape[0], K), np.float32)
    for r in range(rows.shape[0]):
        base.dequantize_row_q4_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*w.shape[:-1], K)


def vec_dot(wrow: np.ndarray, yrow: np.ndarray, generic: bool = False) -> np.float32:
    """ggml_vec_dot_q4_K_q8_K — the __AVX2__ branch in this build (arch/x86/quants.c:1760-1823)."""
    cpu = libs()[1]
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q4K_BLOCK
    s = C.c_float(0)
    fn = cpu.ggml_vec_dot_q4_K_q8_K_generic if generic else cpu.ggml_vec_dot_q4_K_q8_K
    fn(nb * QK_K, C.byref(s), 0, _p(wrow), 0, _p(yrow), 0, 1)
    return np.float32(s.value)


def mul_mat_direct(w: np.ndarray, x: np.ndarray) -> np.ndarray:
    """quantize_row_q8_K + ggml_vec_dot_q4_K_q8_K per (row, col), 1 thread -> float32 [N, M]."""
    cpu = libs()[1]
    w = np.ascontiguousarray(w, dtype=np.uint8)
    yq = quantize_q8_K(x)
    M, N = w.shape[0], yq.shape[0]
    nb = w.shape[1] // Q4K_BLOCK
    out = np.zeros((N, M), np.float32)
    s = C.c_float(0)
    for n in range(N):
        for m in range(M):
            cpu.ggml_vec_dot_q4_K_q8_K(nb * QK_K, C.byref(s), 0, _p(w[m]), 0, _p(yq[n]), 0, 1)
            out[n, m] = s.value
    return out


def mul_mat(w: np.ndarray, x: np.ndarray, n_threads: int = 1, warmup: int = 0, iters: int = 1):
    """Whole op: ggml_mul_mat graph on ggml_backend_cpu, weights in a plain CPU buffer
    (-> ggml_compute_forward_mul_mat -> vec_dot; L/ggml/src/ggml-cpu/ggml-cpu.c:1205-1401).
    Returns (dst float32 [N, M], best_seconds, mean_seconds)."""
    shim = libs()[3]
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    M = w.shape[0]
    N, K = x.shape
    assert w.shape[1] == K // QK_K * Q4K_BLOCK
    dst = np.zeros((N, M), np.float32)
    secs = (C.c_double * 2)()
    rc = shim.ref_mul_mat_q4k(_p(w), M, K, _p(x), N, _p(dst), n_threads, warmup, iters, secs)
    if rc != 0:
        raise RuntimeError(f"ref_mul_mat_q4k rc={rc}")
    return dst, secs[0], secs[1]


def plugin_mul_mat(so_path: str, w: np.ndarray, x: np.ndarray, iters: int = 1):
    """Loads our backend .so through the reference's ggml_backend_load and runs the same graph on it."""
    shim = libs()[3]
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32)
    M = w.shape[0]
    N, K = x.shape
    dst = np.zeros((N, M), np.float32)
    secs = (C.c_double * 2)()
    rc = shim.ref_plugin_mul_mat_q4k(so_path.encode(), _p(w), M, K, _p(x), N, _p(dst), iters, secs)
    if rc != 0:
        raise RuntimeError(f"ref_plugin_mul_mat_q4k rc={rc}")
    return dst, secs[0], secs[1]


def plugin_mul_mat_shared(so_path: str, ws, x: np.ndarray):
    """Several Q4_K matrices on ONE activation tensor (q/k/v of a layer) as one graph on our backend plugin."""
    shim = libs()[3]
    ws = [np.ascontiguousarray(w, dtype=np.uint8) for w in ws]
    x = np.ascontiguousarray(x, dtype=np.float32)
    N, K = x.shape
    n = len(ws)
    outs = [np.zeros((N, w.shape[0]), np.float32) for w in ws]
    rc = shim.ref_plugin_mul_mat_shared_q4k(so_path.encode(), n, (C.c_void_p * n)(*[w.ctypes.data for w in ws]),
                                            (C.c_int64 * n)(*[w.shape[0] for w in ws]), K, _p(x), N,
                                            (C.c_void_p * n)(*[o.ctypes.data for o in outs]))
    if rc != 0:
        raise RuntimeError(f"ref_plugin_mul_mat_shared_q4k rc={rc}")
    return outs


def mul_mat_set(shapes, N: int = 1, n_threads: int = 1, warmup: int = 1, iters: int = 3):
    """One ggml graph holding a Q4_K mul_mat per (M, K) in `shapes` (random valid weights made inside the
    shim), computed on ggml_backend_cpu with n_threads.  Returns (best_seconds, mean_seconds, checksum)."""
    shim = libs()[3]
    n = len(shapes)
    Ms = (C.c_int64 * n)(*[int(m) for m, _ in shapes])
    Ks = (C.c_int64 * n)(*[int(k) for _, k in shapes])
    secs = (C.c_double * 2)()
    chk = C.c_double(0)
    rc = shim.ref_mul_mat_set_q4k(n, Ms, Ks, N, n_threads, warmup, iters, secs, C.byref(chk))
    if rc != 0:
        raise RuntimeError(f"ref_mul_mat_set_q4k rc={rc}")
    return secs[0], secs[1], chk.value


Q6K_BLOCK = 210


def mul_mat_set_typed(shapes, N: int = 1, n_threads: int = 1, warmup: int = 1, iters: int = 3):
    """mul_mat_set with a weight type per matrix: shapes = [(M, K, is_q6k)].  Returns (best_s, mean_s, checksum)."""
    shim = libs()[3]
    n = len(shapes)
    Ms = (C.c_int64 * n)(*[int(s[0]) for s in shapes])
    Ks = (C.c_int64 * n)(*[int(s[1]) for s in shapes])
    q6 = (C.c_int * n)(*[int(bool(s[2])) for s in shapes])
    secs = (C.c_double * 2)()
    chk = C.c_double(0)
    rc = shim.ref_mul_mat_set_typed(n, Ms, Ks, q6, N, n_threads, warmup, iters, secs, C.byref(chk))
    if rc != 0:
        raise RuntimeError(f"ref_mul_mat_set_typed rc={rc}")
    return secs[0], secs[1], chk.value


def mul_mat_set_ex(shapes, N: int = 1, n_threads: int = 1, warmup: int = 1, iters: int = 3, repack: bool = False):
    """mul_mat_set_typed with the weights optionally in the CPU backend's CPU_REPACK buffer type — what llama-bench runs by default
    on an AVX2 host (L/ggml/src/ggml-cpu/repack.cpp:1844-1849: block_q4_Kx8 + ggml_gemv/gemm_q4_K_8x8_q8_K).
    shapes = [(M, K, is_q6k)].  Returns (best_s, mean_s, checksum)."""
    shim = libs()[3]
    n = len(shapes)
    Ms = (C.c_int64 * n)(*[int(s[0]) for s in shapes])
    Ks = (C.c_int64 * n)(*[int(s[1]) for s in shapes])
    q6 = (C.c_int * n)(*[int(bool(s[2])) if len(s) > 2 else 0 for s in shapes])
    secs = (C.c_double * 2)()
    chk = C.c_double(0)
    rc = shim.ref_mul_mat_set_ex(n, Ms, Ks, q6, N, n_threads, warmup, iters, int(repack), secs, C.byref(chk))
    if rc != 0:
        raise RuntimeError(f"ref_mul_mat_set_ex rc={rc}")
    return secs[0], secs[1], chk.value


def direct_q4k(m_sample: int, K: int, N: int = 1, warmup: int = 1, iters: int = 3):
    """BASELINE.md §5 timing (i): quantize_row_q8_K + a loop of ggml_vec_dot_q4_K_q8_K over `m_sample` rows, ONE thread, through the
    reference's type-traits table.  Returns (best_s, mean_s, checksum) per pass over the sample."""
    shim = libs()[3]
    secs = (C.c_double * 2)()
    chk = C.c_double(0)
    shim.ref_direct_q4k.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    rc = shim.ref_direct_q4k(m_sample, K, N, warmup, iters, secs, C.byref(chk))
    if rc != 0:
        raise RuntimeError(f"ref_direct_q4k rc={rc}")
    return secs[0], secs[1], chk.value


def plugin_graph_set(so_path: str, shapes, input_ids, N: int = 1, use_plan: bool = False, warmup: int = 3, iters: int = 20):
    """The whole mul_mat set as ONE ggml graph on the backend plugin, driven through the reference's public backend API per step
    (tensor_set_async -> graph_compute_async | graph_plan_compute -> tensor_get_async -> synchronize), wall-clock.
    shapes = [(M, K, is_q6k)].  Returns dict(best_s, mean_s, checksum, h2d_bytes, d2h_bytes, pinned)."""
    shim = libs()[3]
    n = len(shapes)
    Ms = (C.c_int64 * n)(*[int(s[0]) for s in shapes])
    Ks = (C.c_int64 * n)(*[int(s[1]) for s in shapes])
    q6 = (C.c_int * n)(*[int(bool(s[2])) if len(s) > 2 else 0 for s in shapes])
    ids = (C.c_int * n)(*[int(i) for i in input_ids])
    secs = (C.c_double * 2)()
    chk = C.c_double(0)
    h2d, d2h, pinned = C.c_int64(0), C.c_int64(0), C.c_int(0)
    shim.ref_plugin_graph_set.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    rc = shim.ref_plugin_graph_set(so_path.encode(), n, Ms, Ks, q6, ids, N, int(use_plan), warmup, iters, secs, C.byref(chk),
                                   C.byref(h2d), C.byref(d2h), C.byref(pinned))
    if rc != 0:
        raise RuntimeError(f"ref_plugin_graph_set rc={rc}")
    return {"best_s": secs[0], "mean_s": secs[1], "checksum": chk.value, "h2d_bytes": h2d.value, "d2h_bytes": d2h.value, "pinned": bool(pinned.value)}


def quantize_q6_K(x: np.ndarray) -> np.ndarray:
    """quantize_row_q6_K_ref per row (L/ggml/src/ggml-quants.c:1692-1760)."""
    base = libs()[0]
    x = np.ascontiguousarray(x, dtype=np.float32)
    K = x.shape[-1]
    rows = x.reshape(-1, K)
    out = np.zeros((rows.shape[0], K // QK_K * Q6K_BLOCK), np.uint8)
    for r in range(rows.shape[0]):
        base.quantize_row_q6_K_ref(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*x.shape[:-1], -1)


def dequantize_q6_K(w: np.ndarray) -> np.ndarray:
    base = libs()[0]
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rows = w.reshape(-1, w.shape[-1])
    K = rows.shape[1] // Q6K_BLOCK * QK_K
    out = np.empty((rows.shape[0], K), np.float32)
    for r in range(rows.shape[0]):
        base.dequantize_row_q6_K(_p(rows[r]), _p(out[r]), K)
    return out.reshape(*w.shape[:-1], K)


def vec_dot_q6(wrow: np.ndarray, yrow: np.ndarray, generic: bool = False) -> np.float32:
    """ggml_vec_dot_q6_K_q8_K — the __AVX2__ branch in this build (arch/x86/quants.c:2145-2218)."""
    cpu = libs()[1]
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    yrow = np.ascontiguousarray(yrow, dtype=np.uint8)
    nb = wrow.size // Q6K_BLOCK
    s = C.c_float(0)
    fn = cpu.ggml_vec_dot_q6_K_q8_K_generic if generic else cpu.ggml_vec_dot_q6_K_q8_K
    fn(nb * QK_K, C.byref(s), 0, _p(wrow), 0, _p(yrow), 0, 1)
    return np.float32(s.value)


def has_avx2() -> bool:
    return bool(libs()[3].ref_cpu_has_avx2())


def cheap_ops(x: np.ndarray, res: np.ndarray, w: np.ndarray, g: np.ndarray, u: np.ndarray, eps: float, n_threads: int = 2):
    """The reference's own ADD -> RMS_NORM -> MUL and SWIGLU(split) on ggml_backend_cpu (oracle/ref_shim.c: ref_cheap_ops).
    x, res, g, u: float32 [rows, ne0]; w: [ne0].  -> (x + res, rms_norm(x + res) * w, silu(g) * u)."""
    *_, shim = libs()
    x, res, g, u = (np.ascontiguousarray(a, np.float32) for a in (x, res, g, u))
    w = np.ascontiguousarray(w, np.float32)
    ys, yn, yg = (np.empty_like(x) for _ in range(3))
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = shim.ref_cheap_ops(p(x), p(res), p(w), p(g), p(u), x.shape[1], x.shape[0], eps, n_threads, p(ys), p(yn), p(yg))
    if rc != 0:
        raise RuntimeError(f"ref_cheap_ops failed: {rc}")
    return ys, yn, yg


def rope(x: np.ndarray, pos: np.ndarray, n_dims: int, mode: int = 0, freq_base: float = 10000.0, freq_scale: float = 1.0,
         ext_factor: float = 0.0, attn_factor: float = 1.0, n_ctx_orig: int = 2048, beta_fast: float = 32.0, beta_slow: float = 1.0,
         n_threads: int = 2) -> np.ndarray:
    """The reference's ggml_rope_ext on ggml_backend_cpu (oracle/ref_shim.c: ref_rope).  x: float32 [tokens, heads, ne0]."""
    *_, shim = libs()
    x = np.ascontiguousarray(x, np.float32)
    pos = np.ascontiguousarray(pos, np.int32)
    y = np.empty_like(x)
    rc = shim.ref_rope(x.ctypes.data_as(C.c_void_p), pos.ctypes.data_as(C.c_void_p), x.shape[2], x.shape[1], x.shape[0], n_dims, mode, n_ctx_orig,
                       freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow, n_threads, y.ctypes.data_as(C.c_void_p))
    if rc != 0:
        raise RuntimeError(f"ref_rope failed: {rc}")
    return y


def plugin_split_mul_mat(plugin_path: str, w: np.ndarray, x: np.ndarray, q6k: bool = False, tensor_split=None):
    """ggml_mul_mat with the weights in the plugin's split (row-sharded) buffer type, as llama.cpp's --split-mode row places
    them (oracle/ref_shim.c: ref_plugin_split_mul_mat).  -> (y [N, M], weights read back through get_tensor, device count)."""
    *_, shim = libs()
    blk = 210 if q6k else Q4K_BLOCK
    M, K = w.shape[0], w.shape[1] // blk * QK_K
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w)
    y = np.empty((x.shape[0], M), np.float32)
    wb = np.empty_like(w)
    nd = C.c_int32(0)
    ts = None
    if tensor_split is not None:
        ts = np.zeros(16, np.float32)
        ts[:len(tensor_split)] = tensor_split
    rc = shim.ref_plugin_split_mul_mat(plugin_path.encode(), int(q6k), w.ctypes.data_as(C.c_void_p), M, K, x.ctypes.data_as(C.c_void_p),
                                       x.shape[0], y.ctypes.data_as(C.c_void_p), ts.ctypes.data_as(C.c_void_p) if ts is not None else None,
                                       wb.ctypes.data_as(C.c_void_p), C.byref(nd))
    if rc != 0:
        raise RuntimeError(f"ref_plugin_split_mul_mat failed: rc={rc}")
    return y, wb, nd.value
