// ops.cu — the cheap F32 ops that sit between the weight mul_mats of a llama decoder layer, so that
// a residual / norm / FFN chain stays on the device and the ggml scheduler does not cut the graph around every mul_mat:
//   RMS_NORM   ggml_compute_forward_rms_norm_f32   L/ggml/src/ggml-cpu/ops.cpp:3517-3570
//   ADD / MUL  apply_binary_op<op_add / op_mul>    L/ggml/src/ggml-cpu/binary-ops.cpp:52-120 (src1 broadcast over rows / dims)
//   GLU        ggml_compute_forward_swiglu_f32     L/ggml/src/ggml-cpu/ops.cpp:2934-2990 (split form: silu(gate) * up)
//   GET_ROWS   ggml_compute_forward_get_rows_f32   L/ggml/src/ggml-cpu/ops.cpp:4626-4660
//   ROPE       ggml_compute_forward_rope_f32       L/ggml/src/ggml-cpu/ops.cpp:5436-5475, 5534-5700 (NORMAL and NEOX modes)
// All HBM / latency-bound elementwise or row kernels (a decode token moves a few KB through each); they exist for the launch
// chain, not for throughput: every kernel takes part in programmatic dependent launch like the decode kernels, so a captured
// layer is one dependency chain without host gaps.
//
// Arithmetic.  ADD / MUL are one IEEE operation per element: bit-identical to the reference.  RMS_NORM follows the reference's
// types — squares rounded to float, accumulated in double, mean and 1/sqrtf(mean + eps) in float with IEEE sqrt and divide —
// but sums in a tree, not sequentially: the double sum can differ in its last bits, the float mean almost never does.
// SWIGLU uses x / (1 + expf(-x)) (ggml_silu_f32, vec.h:957) with CUDA's expf; the reference's AVX2 path uses a polynomial
// expf (ggml_v_expf) — agreement is ~1e-6 relative, the tests say so.
#include "../../include/q4k_b200.h"

#include <cstdint>

#include "knobs.h"
#include "q4k_kernels.h"

namespace q4k {
int note_cuda_error(cudaError_t e);
}

namespace {

__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

constexpr int NORM_THREADS = 256;

// y = rms_norm(x [+ r]) [* w];  if r: s = x + r is also written (the residual stream the next ADD needs).
// One CTA per row.  w (optional) is one row of ne0 weights broadcast over all rows (the llama norm weights).
__global__ void __launch_bounds__(NORM_THREADS)
rms_norm_kernel(const float *__restrict__ x, const float *__restrict__ r, const float *__restrict__ w, float *__restrict__ y,
                float *__restrict__ s, int64_t ne0, int64_t ne1, int64_t ne2,
                int64_t xb1, int64_t xb2, int64_t xb3, int64_t rb1, int64_t rb2, int64_t rb3,
                int64_t yb1, int64_t yb2, int64_t yb3, int64_t sb1, int64_t sb2, int64_t sb3, float eps) {
    pdl_enter();
    const int64_t row = blockIdx.x;
    const int64_t i3 = row / (ne1 * ne2), i2 = (row - i3 * ne1 * ne2) / ne1, i1 = row - i3 * ne1 * ne2 - i2 * ne1;
    const float *xr = reinterpret_cast<const float *>(reinterpret_cast<const char *>(x) + i1 * xb1 + i2 * xb2 + i3 * xb3);
    const float *rr = r ? reinterpret_cast<const float *>(reinterpret_cast<const char *>(r) + i1 * rb1 + i2 * rb2 + i3 * rb3) : nullptr;
    float *yr = reinterpret_cast<float *>(reinterpret_cast<char *>(y) + i1 * yb1 + i2 * yb2 + i3 * yb3);
    float *sr = s ? reinterpret_cast<float *>(reinterpret_cast<char *>(s) + i1 * sb1 + i2 * sb2 + i3 * sb3) : nullptr;

    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < ne0; i += NORM_THREADS) {
        float v = xr[i];
        if (rr) { v = __fadd_rn(v, rr[i]); if (sr) sr[i] = v; }
        acc += (double)__fmul_rn(v, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ double part[NORM_THREADS / 32];
    __shared__ float scale_s;
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < NORM_THREADS / 32; ++k) t += part[k];
        const float mean = (float)(t / (double)ne0);
        scale_s = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    }
    __syncthreads();
    const float scale = scale_s;
    for (int64_t i = threadIdx.x; i < ne0; i += NORM_THREADS) {
        float v = rr ? (sr ? sr[i] : __fadd_rn(xr[i], rr[i])) : xr[i];
        v = __fmul_rn(v, scale);
        if (w) v = __fmul_rn(v, w[i]);
        yr[i] = v;
    }
}

struct BinArgs {
    const float *a, *b;
    float *d;
    int64_t ne0, ne1, ne2, ne3;           // shape of a and d
    int64_t bne0, bne1, bne2, bne3;       // shape of b (each divides the corresponding extent of a)
    int64_t ab1, ab2, ab3, bb1, bb2, bb3, db1, db2, db3;     // byte strides of dims 1..3 (dim 0 is dense)
};

template <int OP>
__global__ void __launch_bounds__(256) binary_kernel(const __grid_constant__ BinArgs p) {
    pdl_enter();
    const int64_t row = blockIdx.y;
    const int64_t i3 = row / (p.ne1 * p.ne2), i2 = (row - i3 * p.ne1 * p.ne2) / p.ne1, i1 = row - i3 * p.ne1 * p.ne2 - i2 * p.ne1;
    const float *ar = reinterpret_cast<const float *>(reinterpret_cast<const char *>(p.a) + i1 * p.ab1 + i2 * p.ab2 + i3 * p.ab3);
    const float *br = reinterpret_cast<const float *>(reinterpret_cast<const char *>(p.b) + (i1 % p.bne1) * p.bb1 + (i2 % p.bne2) * p.bb2 +
                                                      (i3 % p.bne3) * p.bb3);
    float *dr = reinterpret_cast<float *>(reinterpret_cast<char *>(p.d) + i1 * p.db1 + i2 * p.db2 + i3 * p.db3);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < p.ne0; i += (int64_t)gridDim.x * blockDim.x) {
        const float bv = br[p.bne0 == p.ne0 ? i : i % p.bne0];
        dr[i] = OP == 0 ? __fadd_rn(ar[i], bv) : __fmul_rn(ar[i], bv);
    }
}

__global__ void __launch_bounds__(256)
swiglu_kernel(const float *__restrict__ g, const float *__restrict__ u, float *__restrict__ d, int64_t nc,
              int64_t gb1, int64_t ub1, int64_t db1) {
    pdl_enter();
    const int64_t row = blockIdx.y;
    const float *gr = reinterpret_cast<const float *>(reinterpret_cast<const char *>(g) + row * gb1);
    const float *ur = reinterpret_cast<const float *>(reinterpret_cast<const char *>(u) + row * ub1);
    float *dr = reinterpret_cast<float *>(reinterpret_cast<char *>(d) + row * db1);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += (int64_t)gridDim.x * blockDim.x) {
        const float x = gr[i];
        const float silu = __fdiv_rn(x, __fadd_rn(1.0f, expf(-x)));
        dr[i] = __fmul_rn(silu, ur[i]);
    }
}

__global__ void __launch_bounds__(256)
get_rows_kernel(const float *__restrict__ src, const int32_t *__restrict__ ids, float *__restrict__ d, int64_t nc, int64_t nrows_src,
                int64_t sb1, int64_t db1) {
    pdl_enter();
    const int64_t row = blockIdx.y;
    int64_t id = ids[row];
    if (id < 0 || id >= nrows_src) return;                     // the reference asserts; out-of-range ids write nothing here
    const float *sr = reinterpret_cast<const float *>(reinterpret_cast<const char *>(src) + id * sb1);
    float *dr = reinterpret_cast<float *>(reinterpret_cast<char *>(d) + row * db1);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nc; i += (int64_t)gridDim.x * blockDim.x) dr[i] = sr[i];
}

// ROPE, modes NORMAL (pairs (2i, 2i+1)) and NEOX (pairs (i, i + n_dims/2)).  One thread per pair.  theta follows the reference's
// cache loop exactly: theta = (float)pos, multiplied by theta_scale once per preceding pair (sequential float products), so the
// angle has the reference's bits; cosf / sinf are CUDA's (<= 2 ulp), the rotation may contract to FMAs as gcc's does.
struct RopeArgs {
    const float *x;
    const int32_t *pos;
    const float *ff;
    float *d;
    int64_t ne0, ne1, ne2, ne3, xb1, xb2, xb3, db1, db2, db3;
    int n_dims, neox;
    float theta_scale, freq_scale, ext_factor, attn_factor, corr_lo, corr_hi;
};

__global__ void __launch_bounds__(128) rope_kernel(const __grid_constant__ RopeArgs p) {
    pdl_enter();
    const int64_t pair = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;          // pair index inside the row (ne0 / 2 pairs)
    if (2 * pair >= p.ne0) return;
    const int64_t i1 = blockIdx.y, i23 = blockIdx.z, i2 = i23 % p.ne2, i3 = i23 / p.ne2;
    const float *xr = reinterpret_cast<const float *>(reinterpret_cast<const char *>(p.x) + i1 * p.xb1 + i2 * p.xb2 + i3 * p.xb3);
    float *dr = reinterpret_cast<float *>(reinterpret_cast<char *>(p.d) + i1 * p.db1 + i2 * p.db2 + i3 * p.db3);
    const int64_t i0 = 2 * pair;
    if (i0 >= p.n_dims) {                          // channels beyond the rotated ones are copied
        dr[i0] = xr[i0];
        dr[i0 + 1] = xr[i0 + 1];
        return;
    }
    float theta = (float)(int64_t)p.pos[i2];
    for (int64_t k = 0; k < pair; ++k) theta = __fmul_rn(theta, p.theta_scale);
    const float ff = p.ff ? p.ff[pair] : 1.0f;
    const float theta_extrap = __fdiv_rn(theta, ff);
    const float theta_interp = __fmul_rn(p.freq_scale, theta_extrap);
    float th = theta_interp, mscale = p.attn_factor;
    if (p.ext_factor != 0.0f) {
        const float y = ((float)(i0 / 2) - p.corr_lo) / fmaxf(0.001f, p.corr_hi - p.corr_lo);
        const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * p.ext_factor;
        th = theta_interp * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
        mscale *= 1.0f + 0.1f * logf(1.0f / p.freq_scale);
    }
    const float c = cosf(th) * mscale, sn = sinf(th) * mscale;
    const int64_t a = p.neox ? pair : i0, b = p.neox ? pair + p.n_dims / 2 : i0 + 1;
    const float x0 = xr[a], x1 = xr[b];
    dr[a] = x0 * c - x1 * sn;
    dr[b] = x0 * sn + x1 * c;
}

template <typename... KArgs, typename... Args>
int launch(void (*kern)(KArgs...), dim3 grid, dim3 block, void *stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = q4k::knobs().pdl != 0 ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
    q4k::count_launch();
    return e == cudaSuccess ? Q4K_OK : q4k::note_cuda_error(e);
}

bool strides_ok(const int64_t *nb) { return nb && !(nb[0] & 3) && !(nb[1] & 3) && !(nb[2] & 3); }

}  // namespace

extern "C" {

int q4k_op_rms_norm(const float *x_dev, const int64_t *x_nb, const float *res_dev, const int64_t *res_nb, const float *w_dev,
                    float *y_dev, const int64_t *y_nb, float *sum_dev, const int64_t *sum_nb, const int64_t *ne, float eps, void *stream) {
    if (!ne || ne[0] < 0 || ne[1] < 0 || ne[2] < 0 || ne[3] < 0 || eps < 0.0f) return Q4K_ERR_BAD_SHAPE;
    const int64_t rows = ne[1] * ne[2] * ne[3];
    if (ne[0] == 0 || rows == 0) return Q4K_OK;
    if (!x_dev || !y_dev || (sum_dev && !res_dev)) return Q4K_ERR_BAD_POINTER;
    if (!strides_ok(x_nb) || !strides_ok(y_nb) || (res_dev && !strides_ok(res_nb)) || (sum_dev && !strides_ok(sum_nb))) return Q4K_ERR_BAD_SHAPE;
    if (rows > 0x7fffffff) return Q4K_ERR_BAD_SHAPE;
    static const int64_t zero[3] = {0, 0, 0};
    const int64_t *rb = res_dev ? res_nb : zero, *sb = sum_dev ? sum_nb : zero;
    return launch(rms_norm_kernel, dim3((unsigned)rows), dim3(NORM_THREADS), stream, x_dev, res_dev, w_dev, y_dev, sum_dev, ne[0], ne[1], ne[2],
                  x_nb[0], x_nb[1], x_nb[2], rb[0], rb[1], rb[2], y_nb[0], y_nb[1], y_nb[2], sb[0], sb[1], sb[2], eps);
}

int q4k_op_binary(int op, const float *a_dev, const int64_t *a_ne, const int64_t *a_nb, const float *b_dev, const int64_t *b_ne,
                  const int64_t *b_nb, float *dst_dev, const int64_t *dst_nb, void *stream) {
    if (op != Q4K_OP_ADD && op != Q4K_OP_MUL) return Q4K_ERR_UNSUPPORTED;
    if (!a_ne || !b_ne) return Q4K_ERR_BAD_SHAPE;
    for (int d = 0; d < 4; ++d)
        if (a_ne[d] < 0 || b_ne[d] <= 0 || (a_ne[d] && a_ne[d] % b_ne[d])) return a_ne[d] == 0 ? Q4K_OK : Q4K_ERR_BAD_SHAPE;      // ggml_can_repeat(b, a)
    const int64_t rows = a_ne[1] * a_ne[2] * a_ne[3];
    if (a_ne[0] == 0 || rows == 0) return Q4K_OK;
    if (!a_dev || !b_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    if (!strides_ok(a_nb) || !strides_ok(b_nb) || !strides_ok(dst_nb) || rows > 65535LL * 65535LL) return Q4K_ERR_BAD_SHAPE;
    if (rows > 65535) return Q4K_ERR_UNSUPPORTED;
    BinArgs p{a_dev, b_dev, dst_dev, a_ne[0], a_ne[1], a_ne[2], a_ne[3], b_ne[0], b_ne[1], b_ne[2], b_ne[3],
              a_nb[0], a_nb[1], a_nb[2], b_nb[0], b_nb[1], b_nb[2], dst_nb[0], dst_nb[1], dst_nb[2]};
    const unsigned gx = (unsigned)((a_ne[0] + 255) / 256 > 64 ? 64 : (a_ne[0] + 255) / 256);
    return op == Q4K_OP_ADD ? launch(binary_kernel<0>, dim3(gx, (unsigned)rows), dim3(256), stream, p)
                            : launch(binary_kernel<1>, dim3(gx, (unsigned)rows), dim3(256), stream, p);
}

int q4k_op_swiglu(const float *gate_dev, int64_t gate_row_stride_bytes, const float *up_dev, int64_t up_row_stride_bytes,
                  float *dst_dev, int64_t dst_row_stride_bytes, int64_t nc, int64_t nrows, void *stream) {
    if (nc < 0 || nrows < 0) return Q4K_ERR_BAD_SHAPE;
    if (nc == 0 || nrows == 0) return Q4K_OK;
    if (!gate_dev || !up_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    if ((gate_row_stride_bytes | up_row_stride_bytes | dst_row_stride_bytes) & 3) return Q4K_ERR_BAD_SHAPE;
    if (nrows > 65535) return Q4K_ERR_UNSUPPORTED;
    const unsigned gx = (unsigned)((nc + 255) / 256 > 64 ? 64 : (nc + 255) / 256);
    return launch(swiglu_kernel, dim3(gx, (unsigned)nrows), dim3(256), stream, gate_dev, up_dev, dst_dev, nc, gate_row_stride_bytes,
                  up_row_stride_bytes, dst_row_stride_bytes);
}

int q4k_op_get_rows_f32(const float *src_dev, int64_t src_row_stride_bytes, int64_t nrows_src, const int32_t *ids_dev, int64_t n_ids,
                        float *dst_dev, int64_t dst_row_stride_bytes, int64_t nc, void *stream) {
    if (nc < 0 || n_ids < 0 || nrows_src < 0) return Q4K_ERR_BAD_SHAPE;
    if (nc == 0 || n_ids == 0) return Q4K_OK;
    if (!src_dev || !ids_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    if ((src_row_stride_bytes | dst_row_stride_bytes) & 3) return Q4K_ERR_BAD_SHAPE;
    if (n_ids > 65535) return Q4K_ERR_UNSUPPORTED;
    const unsigned gx = (unsigned)((nc + 255) / 256 > 64 ? 64 : (nc + 255) / 256);
    return launch(get_rows_kernel, dim3(gx, (unsigned)n_ids), dim3(256), stream, src_dev, ids_dev, dst_dev, nc, nrows_src, src_row_stride_bytes,
                  dst_row_stride_bytes);
}

int q4k_op_rope(const float *x_dev, const int64_t *x_nb, const int32_t *pos_dev, const float *freq_factors_dev, float *dst_dev,
                const int64_t *dst_nb, const int64_t *ne, int n_dims, int neox, float theta_scale, float freq_scale, float ext_factor,
                float attn_factor, float corr_lo, float corr_hi, void *stream) {
    if (!ne || ne[0] < 0 || ne[1] < 0 || ne[2] < 0 || ne[3] < 0 || (ne[0] & 1) || n_dims < 0 || (n_dims & 1) || n_dims > ne[0]) return Q4K_ERR_BAD_SHAPE;
    if (ne[0] == 0 || ne[1] == 0 || ne[2] == 0 || ne[3] == 0) return Q4K_OK;
    if (!x_dev || !pos_dev || !dst_dev) return Q4K_ERR_BAD_POINTER;
    if (!strides_ok(x_nb) || !strides_ok(dst_nb)) return Q4K_ERR_BAD_SHAPE;
    if (ne[1] > 65535 || ne[2] * ne[3] > 65535) return Q4K_ERR_UNSUPPORTED;
    RopeArgs p{x_dev, pos_dev, freq_factors_dev, dst_dev, ne[0], ne[1], ne[2], ne[3], x_nb[0], x_nb[1], x_nb[2], dst_nb[0], dst_nb[1], dst_nb[2],
               n_dims, neox, theta_scale, freq_scale, ext_factor, attn_factor, corr_lo, corr_hi};
    const int64_t pairs = ne[0] / 2;
    const int threads = pairs >= 128 ? 128 : (int)((pairs + 31) / 32 * 32);
    return launch(rope_kernel, dim3((unsigned)((pairs + threads - 1) / threads), (unsigned)ne[1], (unsigned)(ne[2] * ne[3])), dim3(threads), stream, p);
}

}  // extern "C"
