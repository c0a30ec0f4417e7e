// fakequant2.cu - the quantize2.py flavour of the fake-quantizer.
//
// Reference: models/modules/quantize2.py.  QParams = (range, zero_point) TENSORS that stay on the device; one value per
// tensor ([1,1,1,1]) or one per row of dim 0 ([C,1,1,1], weights).  UniformQuantize.forward (:41-76):
//     qmin = -2^(b-1) if signed else 0 ; qmax = qmin + 2^b - 1 ; scale = range / (qmax - qmin)        (fp32 tensor math)
//     t = (x + (qmin*scale - zero_point)) / scale  [+ uniform(-.5,.5)] ; q = rne(clamp(t, qmin, qmax))
//     y = q * scale + (zero_point - qmin*scale)    (only when dequantize)
// No 1e-8 floor on the scale here (a zero range divides by zero exactly as the reference does).  Every operation is
// rounded separately, in the reference's order.
#include <math.h>

#include "common.cuh"

constexpr int F2_THREADS = 256;

struct Q2Params {
    float add_in, scale, add_out, qmin, qmax;
};
__device__ __forceinline__ Q2Params make_q2(float zp, float range, int bits, int is_signed) {
    Q2Params p;
    p.qmin = is_signed ? -(float)(1u << (bits - 1)) : 0.0f;
    p.qmax = p.qmin + (float)((1u << bits) - 1u);                         // exact: bits <= 24
    p.scale = __fdiv_rn(range, __fsub_rn(p.qmax, p.qmin));
    const float qs = __fmul_rn(p.qmin, p.scale);
    p.add_in = __fsub_rn(qs, zp);
    p.add_out = __fsub_rn(zp, qs);
    return p;
}
__device__ __forceinline__ float q2_one(float x, float u, bool noisy, bool deq, const Q2Params& p) {
    float t = __fdiv_rn(__fadd_rn(x, p.add_in), p.scale);
    if (noisy) t = __fadd_rn(t, u);
    t = rintf(fminf(fmaxf(t, p.qmin), p.qmax));
    return deq ? __fadd_rn(__fmul_rn(t, p.scale), p.add_out) : t;
}

// rows x cols view; params per row (per_row) or one for everything.  Thread = 4 consecutive elements of one row when
// cols % 4 == 0 and the pointers are 16-byte aligned, else scalar.
template <int NOISE, bool VEC>
__global__ void __launch_bounds__(F2_THREADS)
k_fakequant2(const float* __restrict__ x, float* __restrict__ y, int64_t n, int64_t cols, const float* __restrict__ zp,
             const float* __restrict__ range, int per_row, int bits, int is_signed, int deq, NoiseArg na) {
    const uint64_t seed = effective_seed(na);
    const int64_t stride = (int64_t)gridDim.x * F2_THREADS;
    Q2Params pu;
    if (!per_row) pu = make_q2(__ldg(zp), __ldg(range), bits, is_signed);
    if (VEC) {
        for (int64_t i = (int64_t)blockIdx.x * F2_THREADS + threadIdx.x; i < (n >> 2); i += stride) {
            const int64_t e = 4 * i;
            Q2Params p = pu;
            if (per_row) {
                const int64_t r = e / cols;
                p = make_q2(__ldg(zp + r), __ldg(range + r), bits, is_signed);
            }
            const float4 v = ld_stream4(x + e);
            float u[4] = {0.f, 0.f, 0.f, 0.f};
            if (NOISE == NOISE_INJECT) {
                const float4 w = ld_stream4(na.inject + e);
                u[0] = w.x; u[1] = w.y; u[2] = w.z; u[3] = w.w;
            } else if (NOISE == NOISE_PHILOX) {
#pragma unroll
                for (int j = 0; j < 4; ++j) u[j] = philox_noise1(seed, (uint64_t)(e + j) + na.offset);
            }
            st_stream4(y + e, make_float4(q2_one(v.x, u[0], NOISE != NOISE_NONE, deq, p), q2_one(v.y, u[1], NOISE != NOISE_NONE, deq, p),
                                          q2_one(v.z, u[2], NOISE != NOISE_NONE, deq, p), q2_one(v.w, u[3], NOISE != NOISE_NONE, deq, p)));
        }
    } else {
        for (int64_t e = (int64_t)blockIdx.x * F2_THREADS + threadIdx.x; e < n; e += stride) {
            Q2Params p = pu;
            if (per_row) {
                const int64_t r = e / cols;
                p = make_q2(__ldg(zp + r), __ldg(range + r), bits, is_signed);
            }
            float u = 0.f;
            if (NOISE == NOISE_INJECT) u = na.inject[e];
            else if (NOISE == NOISE_PHILOX) u = philox_noise1(seed, (uint64_t)e + na.offset);
            y[e] = q2_one(x[e], u, NOISE != NOISE_NONE, deq, p);
        }
    }
}

extern "C" int qb_fakequant2(const float* x, float* y, int64_t rows, int64_t cols, const float* zero_point, const float* range,
                             int per_row, int bits, int is_signed, int dequantize, const float* noise, int stochastic,
                             uint64_t seed, uint64_t offset, const uint64_t* step_dev, void* stream) {
    QB_CHECK_ARG(rows >= 0 && cols >= 0, "qb_fakequant2: rows=%lld cols=%lld", (long long)rows, (long long)cols);
    const int64_t n = rows * cols;
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(x && y && zero_point && range, "qb_fakequant2: null pointer");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_fakequant2: bits=%d outside 1..24", bits);
    const int nm = !stochastic ? NOISE_NONE : (noise ? NOISE_INJECT : NOISE_PHILOX);
    NoiseArg na{noise, seed, offset, step_dev};
    const uintptr_t al = reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(noise);
    const bool vec = (al & 15) == 0 && (cols % 4 == 0);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = ((vec ? n >> 2 : n) + F2_THREADS - 1) / F2_THREADS, cap = (int64_t)sms * 8;
    const int grid = (int)(want < cap ? (want < 1 ? 1 : want) : cap);
    cudaStream_t st = qb_stream(stream);
#define QB_F2(NM)                                                                                                        \
    do {                                                                                                                 \
        if (vec) k_fakequant2<NM, true><<<grid, F2_THREADS, 0, st>>>(x, y, n, cols, zero_point, range, per_row, bits,      \
                                                                    is_signed, dequantize, na);                          \
        else k_fakequant2<NM, false><<<grid, F2_THREADS, 0, st>>>(x, y, n, cols, zero_point, range, per_row, bits,         \
                                                                  is_signed, dequantize, na);                            \
    } while (0)
    if (nm == NOISE_NONE) QB_F2(NOISE_NONE);
    else if (nm == NOISE_INJECT) QB_F2(NOISE_INJECT);
    else QB_F2(NOISE_PHILOX);
#undef QB_F2
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// Per-tensor, unsigned, deterministic quantize2 quantizer that ALSO emits the integer codes as uint8 (operand of the int8
// tensor-core contractions) and bf16 (operand of qb_conv_wgrad): what routes quantize2.QConv2d / QLinear onto the native
// contractions.  y (optional) = the dequantised tensor, bit-identical to qb_fakequant2.
__global__ void __launch_bounds__(F2_THREADS)
k_fakequant2_codes(const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ codes, uint16_t* __restrict__ codes16,
                   int64_t n, const float* __restrict__ zp, const float* __restrict__ range, int bits) {
    const Q2Params p = make_q2(__ldg(zp), __ldg(range), bits, 0);
    for (int64_t e = (int64_t)blockIdx.x * F2_THREADS + threadIdx.x; e < n; e += (int64_t)gridDim.x * F2_THREADS) {
        const float q = q2_one(x[e], 0.f, false, false, p);
        if (y) y[e] = __fadd_rn(__fmul_rn(q, p.scale), p.add_out);
        if (codes) codes[e] = (uint8_t)q;
        if (codes16) codes16[e] = (uint16_t)(__float_as_uint(q) >> 16);
    }
}

extern "C" int qb_fakequant2_codes(const float* x, float* y, uint8_t* codes, uint16_t* codes16, int64_t n, const float* zero_point,
                                   const float* range, int bits, void* stream) {
    QB_CHECK_ARG(n >= 0, "qb_fakequant2_codes: n=%lld", (long long)n);
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(x && zero_point && range && (y || codes || codes16), "qb_fakequant2_codes: null pointer");
    QB_CHECK_ARG(bits >= 1 && bits <= 8, "qb_fakequant2_codes: 8-bit codes need bits<=8 (bits=%d)", bits);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (n + F2_THREADS - 1) / F2_THREADS, cap = (int64_t)sms * 16;
    k_fakequant2_codes<<<(int)(want < cap ? want : cap), F2_THREADS, 0, qb_stream(stream)>>>(x, y, codes, codes16, n, zero_point, range, bits);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// per-row min / max (calculate_qparams with flatten_dims=(1,-1), quantize2.py:19-28): one CTA per row
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_row_minmax(const float* __restrict__ x, int64_t cols, float* row_min, float* row_max) {
    __shared__ float sm[64];
    const float* src = x + (int64_t)blockIdx.x * cols;
    float mn = QB_INF, mx = -QB_INF;
    for (int64_t j = threadIdx.x; j < cols; j += 256) {
        const float v = __ldg(src + j);
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
    }
    mn = warp_min(mn);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = mn; sm[32 + (threadIdx.x >> 5)] = mx; }
    __syncthreads();
    if (threadIdx.x < 32) {
        mn = warp_min(threadIdx.x < 8 ? sm[threadIdx.x] : QB_INF);
        mx = warp_max(threadIdx.x < 8 ? sm[32 + threadIdx.x] : -QB_INF);
        if (threadIdx.x == 0) { row_min[blockIdx.x] = mn; row_max[blockIdx.x] = mx; }
    }
}

extern "C" int qb_row_minmax(const float* x, int64_t rows, int64_t cols, float* row_min, float* row_max, void* stream) {
    QB_CHECK_ARG(x && row_min && row_max, "qb_row_minmax: null pointer");
    QB_CHECK_ARG(rows > 0 && cols > 0 && rows <= 2147483647LL, "qb_row_minmax: rows=%lld cols=%lld", (long long)rows, (long long)cols);
    k_row_minmax<<<(unsigned)rows, 256, 0, qb_stream(stream)>>>(x, cols, row_min, row_max);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// quantize2.RangeBN (:275-349): same chunked range statistics as quantize.RangeBN, but the scale is NOT inverted or
// quantized, gamma / beta are used as they are, running_var tracks the scale itself and the normalisation DIVIDES:
//     scale = (mean_max - mean_min) * scale_fix ; out = (x - mean) / (scale + eps) * gamma + beta
// finalize: red[3][C] (qb_rangebn_stats*) -> mean, scale, den = scale + eps, rinv = 1/den (for the backward kernels,
// which are the quantize.RangeBN ones with qscale := rinv, qgamma := gamma), running-buffer update.
// ---------------------------------------------------------------------------------------------------
__global__ void k_rangebn2_finalize(const double* red, int64_t C, double K, double count, double scale_fix, float eps,
                                    double momentum, int training, float* running_mean, float* running_var, float* mean_out,
                                    float* scale_out, float* den_out, float* rinv_out) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float m_f = (float)momentum, om_f = (float)(1.0 - momentum);
    float mean, scale;
    if (training) {
        const float mean_max = (float)(red[c] / K), mean_min = (float)(red[C + c] / K);
        mean = (float)(red[2 * C + c] / count);
        scale = __fmul_rn(__fsub_rn(mean_max, mean_min), (float)scale_fix);
        running_mean[c] = __fadd_rn(__fmul_rn(running_mean[c], m_f), __fmul_rn(mean, om_f));
        running_var[c] = __fadd_rn(__fmul_rn(running_var[c], m_f), __fmul_rn(scale, om_f));
    } else {
        mean = running_mean[c];
        scale = running_var[c];
    }
    const float den = __fadd_rn(scale, eps);
    mean_out[c] = mean;
    scale_out[c] = scale;
    den_out[c] = den;
    rinv_out[c] = __fdiv_rn(1.0f, den);
}

extern "C" int qb_rangebn2_finalize(const double* red, int64_t C, int num_chunks, double count, double chunk_len, float eps,
                                    double momentum, int training, float* running_mean, float* running_var, float* mean,
                                    float* scale, float* den, float* rinv, void* stream) {
    QB_CHECK_ARG(C > 0 && running_mean && running_var && mean && scale && den && rinv, "qb_rangebn2_finalize: null pointer");
    QB_CHECK_ARG(!training || (red && num_chunks > 0 && count > 0 && chunk_len > 1),
                 "qb_rangebn2_finalize: training needs red, num_chunks>0, count>0, chunk_len>1");
    double scale_fix = 0.0;
    if (training)   // quantize2.py:315-316, Python double
        scale_fix = (0.5 * 0.35) * (1.0 + sqrt(3.14159265358979323846 * log(4.0))) / sqrt(2.0 * log(chunk_len));
    k_rangebn2_finalize<<<(unsigned)((C + 127) / 128), 128, 0, qb_stream(stream)>>>(red, C, (double)num_chunks, count, scale_fix,
                                                                                   eps, momentum, training, running_mean,
                                                                                   running_var, mean, scale, den, rinv);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// out = ((x - mean_c) / den_c) * gamma_c + beta_c ; channel of element e: NCHW (e / HW) % C, channels-last e % C
__global__ void __launch_bounds__(F2_THREADS)
k_rangebn2_apply(const float* __restrict__ x, float* __restrict__ out, int64_t n, int64_t HW, int64_t C, int nhwc,
                 const float* mean, const float* den, const float* gamma, const float* beta) {
    for (int64_t e = (int64_t)blockIdx.x * F2_THREADS + threadIdx.x; e < n; e += (int64_t)gridDim.x * F2_THREADS) {
        const int64_t c = nhwc ? e % C : (e / HW) % C;
        float t = __fdiv_rn(__fsub_rn(x[e], __ldg(mean + c)), __ldg(den + c));
        if (gamma) t = __fmul_rn(t, __ldg(gamma + c));
        if (beta) t = __fadd_rn(t, __ldg(beta + c));
        out[e] = t;
    }
}

extern "C" int qb_rangebn2_apply(const float* x, float* out, int64_t B, int64_t C, int64_t HW, int nhwc, const float* mean,
                                 const float* den, const float* gamma, const float* beta, void* stream) {
    QB_CHECK_ARG(x && out && mean && den, "qb_rangebn2_apply: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn2_apply: bad shape");
    const int64_t n = B * C * HW;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (n + F2_THREADS - 1) / F2_THREADS, cap = (int64_t)sms * 16;
    k_rangebn2_apply<<<(unsigned)(want < cap ? want : cap), F2_THREADS, 0, qb_stream(stream)>>>(x, out, n, HW, C, nhwc, mean, den,
                                                                                              gamma, beta);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
