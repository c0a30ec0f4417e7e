// rangebn.cu - Range Batch-Norm statistics / apply / backward on NCHW in place (no transpose copy).
//
// Reference: models/modules/quantize.py:461-505.  The reference materialises x.transpose(0,1).contiguous()
// (a full extra read+write), then three ATen reductions and four broadcast elementwise passes.  Here the
// (channel, chunk) runs are walked directly in the NCHW tensor.
#include <math.h>

#include "common.cuh"
#include "rangebn_common.cuh"

constexpr int RB_THREADS = 256;

// ---------------------------------------------------------------------------------------------------
// walk the elements of task (c, k): flat (b,p) indices e in [e0, e1) of channel c, run by run
// ---------------------------------------------------------------------------------------------------
template <int G, class F4, class F1>
__device__ __forceinline__ void walk_task(const float* __restrict__ x, int64_t C, int64_t HW, int64_t c, int64_t e0,
                                          int64_t e1, int lane, bool vec_ok, F4&& f4, F1&& f1) {
    const int64_t b_first = e0 / HW, b_last = (e1 - 1) / HW;
    for (int64_t b = b_first; b <= b_last; ++b) {
        const int64_t p_lo = (b == b_first) ? e0 - b * HW : 0;
        const int64_t p_hi = (b == b_last) ? e1 - b * HW : HW;
        const int64_t base = (b * C + c) * HW;
        if (vec_ok && ((p_lo & 3) == 0)) {
            const int64_t nq = (p_hi - p_lo) >> 2;
            for (int64_t q = lane; q < nq; q += G) f4(base + p_lo + 4 * q, b * HW + p_lo + 4 * q);
            for (int64_t p = p_lo + 4 * nq + lane; p < p_hi; p += G) f1(base + p, b * HW + p);
        } else {
            for (int64_t p = p_lo + lane; p < p_hi; p += G) f1(base + p, b * HW + p);
        }
    }
}

struct ArgVal {
    float v;
    int32_t i;
};
__device__ __forceinline__ void upd_max(ArgVal& a, float v, int32_t i) { if (v > a.v) { a.v = v; a.i = i; } }
__device__ __forceinline__ void upd_min(ArgVal& a, float v, int32_t i) { if (v < a.v) { a.v = v; a.i = i; } }
// merge keeping the FIRST occurrence (smallest flat index) among equal extrema - ATen's tie-break
__device__ __forceinline__ void merge_max(ArgVal& a, float v, int32_t i) { if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; } }
__device__ __forceinline__ void merge_min(ArgVal& a, float v, int32_t i) { if (v < a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; } }

// G threads per (c,k) task; RB_THREADS/G tasks per CTA
template <int G>
__global__ void __launch_bounds__(RB_THREADS)
k_rangebn_stats(const float* __restrict__ x, int64_t B, int64_t C, int64_t HW, int K, int64_t L, int vec_ok, float* cmax,
                float* cmin, double* psum, int32_t* amax, int32_t* amin) {
    constexpr int TPB = RB_THREADS / G;
    const int lane = threadIdx.x % G;
    const int64_t task = (int64_t)blockIdx.x * TPB + threadIdx.x / G;
    const int64_t ntasks = C * K;
    const bool active = task < ntasks;
    const int64_t c = active ? task / K : 0, k = active ? task % K : 0;
    ArgVal mx{-QB_INF, 0x7fffffff}, mn{QB_INF, 0x7fffffff};
    double dsum = 0.0;
    if (active) {
        float fs = 0.f;
        int cnt = 0;
        auto flush = [&]() { dsum += (double)fs; fs = 0.f; cnt = 0; };
        walk_task<G>(x, C, HW, c, k * L, (k + 1) * L, lane, vec_ok != 0,
            [&](int64_t addr, int64_t e) {
                float4 v = __ldcs(reinterpret_cast<const float4*>(x + addr));
                int32_t i = (int32_t)e;
                upd_max(mx, v.x, i); upd_max(mx, v.y, i + 1); upd_max(mx, v.z, i + 2); upd_max(mx, v.w, i + 3);
                upd_min(mn, v.x, i); upd_min(mn, v.y, i + 1); upd_min(mn, v.z, i + 2); upd_min(mn, v.w, i + 3);
                fs += (v.x + v.y) + (v.z + v.w);
                if (++cnt == 8) flush();
            },
            [&](int64_t addr, int64_t e) {
                float v = x[addr];
                upd_max(mx, v, (int32_t)e);
                upd_min(mn, v, (int32_t)e);
                fs += v;
                if (++cnt == 32) flush();
            });
        flush();
    }
    // reduce over the G threads of the task
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        float v = __shfl_xor_sync(0xffffffffu, mx.v, o); int32_t i = __shfl_xor_sync(0xffffffffu, mx.i, o);
        merge_max(mx, v, i);
        v = __shfl_xor_sync(0xffffffffu, mn.v, o); i = __shfl_xor_sync(0xffffffffu, mn.i, o);
        merge_min(mn, v, i);
        dsum += __shfl_xor_sync(0xffffffffu, dsum, o);
    }
    if (G > 32) {
        __shared__ float sv[2][RB_THREADS / 32];
        __shared__ int32_t si[2][RB_THREADS / 32];
        __shared__ double ss[RB_THREADS / 32];
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) { sv[0][w] = mx.v; si[0][w] = mx.i; sv[1][w] = mn.v; si[1][w] = mn.i; ss[w] = dsum; }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int j = 1; j < RB_THREADS / 32; ++j) {
                merge_max(mx, sv[0][j], si[0][j]);
                merge_min(mn, sv[1][j], si[1][j]);
                dsum += ss[j];
            }
        }
    }
    if (active && lane == 0) {
        cmax[task] = mx.v;
        cmin[task] = mn.v;
        psum[task] = dsum;
        if (amax) amax[task] = mx.i;
        if (amin) amin[task] = mn.i;
    }
}

// red[0*C+c] = sum_k cmax, red[1*C+c] = sum_k cmin, red[2*C+c] = sum x  (fixed order)
__global__ void k_rangebn_pack(const float* cmax, const float* cmin, const double* psum, int64_t C, int K, double* red) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double a = 0.0, b = 0.0, s = 0.0;
    for (int k = 0; k < K; ++k) {
        a += (double)cmax[c * K + k];
        b += (double)cmin[c * K + k];
        s += psum[c * K + k];
    }
    red[c] = a;
    red[C + c] = b;
    red[2 * C + c] = s;
}

extern "C" size_t qb_rangebn_ws_bytes(int64_t C, int num_chunks) {
    if (C < 0 || num_chunks < 0) return 0;
    return (size_t)C * num_chunks * 16;
}

extern "C" int qb_rangebn_stats(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                                int32_t* amin, double* red, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x && red && ws, "qb_rangebn_stats: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0 && num_chunks > 0, "qb_rangebn_stats: B=%lld C=%lld HW=%lld chunks=%d",
                 (long long)B, (long long)C, (long long)HW, num_chunks);
    QB_CHECK_ARG((B * HW) % num_chunks == 0,
                 "qb_rangebn_stats: B*H*W=%lld not divisible by num_chunks=%d (the reference's view() fails too)",
                 (long long)(B * HW), num_chunks);
    QB_CHECK_ARG(B * HW < 2147483647LL, "qb_rangebn_stats: per-channel element count exceeds int32 indices");
    if (ws_bytes < qb_rangebn_ws_bytes(C, num_chunks)) {
        qb_set_error("qb_rangebn_stats: workspace %zu < %zu bytes", ws_bytes, qb_rangebn_ws_bytes(C, num_chunks));
        return QB_ERR_WORKSPACE;
    }
    const int64_t ntasks = C * num_chunks, L = B * HW / num_chunks;
    double* psum = static_cast<double*>(ws);
    float* cmax = reinterpret_cast<float*>(psum + ntasks);
    float* cmin = cmax + ntasks;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(x) & 15) == 0) && (HW % 4 == 0);
    cudaStream_t st = qb_stream(stream);
    if (L >= 4096) {
        k_rangebn_stats<RB_THREADS><<<(unsigned)ntasks, RB_THREADS, 0, st>>>(x, B, C, HW, num_chunks, L, vec_ok, cmax, cmin,
                                                                            psum, amax, amin);
    } else {
        const int tpb = RB_THREADS / 32;
        k_rangebn_stats<32><<<(unsigned)((ntasks + tpb - 1) / tpb), RB_THREADS, 0, st>>>(x, B, C, HW, num_chunks, L, vec_ok,
                                                                                        cmax, cmin, psum, amax, amin);
    }
    QB_LAUNCH_CHECK();
    k_rangebn_pack<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(cmax, cmin, psum, C, num_chunks, red);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// per-channel finalize (quantize.py:470-498), single CTA
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(RF_THREADS) k_rangebn_finalize(const double* red, int64_t C, RbnFinalizeArg a) {
    __shared__ float sm[6 * 32];
    rbn_finalize_body(red, C, a, sm);
}

extern "C" int qb_rangebn_finalize(const double* red, int64_t C, int num_chunks, double count, double chunk_len,
                                   float eps, double momentum, int training, int bits, const float* gamma,
                                   const float* beta, float* running_mean, float* running_var, float* mean,
                                   float* scale_raw, float* qscale, float* qgamma, float* qbeta, void* stream) {
    QB_CHECK_ARG(C > 0 && running_mean && running_var && mean && scale_raw && qscale, "qb_rangebn_finalize: null pointer");
    QB_CHECK_ARG(!training || (red && num_chunks > 0 && count > 0 && chunk_len > 1),
                 "qb_rangebn_finalize: training needs red, num_chunks>0, count>0, chunk_len>1");
    QB_CHECK_ARG((!gamma || qgamma) && (!beta || qbeta), "qb_rangebn_finalize: gamma/beta without output");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_rangebn_finalize: bits=%d", bits);
    RbnFinalizeArg a{(double)num_chunks, count, training ? rbn_scale_fix(chunk_len) : 0.0, momentum, eps, training, bits, gamma, beta,
                     running_mean, running_var, mean, scale_raw, qscale, qgamma, qbeta};
    k_rangebn_finalize<<<1, RF_THREADS, 0, qb_stream(stream)>>>(red, C, a);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// apply: out = ((x - mean_c) * qscale_c) * qgamma_c + qbeta_c   (quantize.py:488-499), 8 B/elem
// ---------------------------------------------------------------------------------------------------
struct ChanParams {
    const float *mean, *qscale, *qgamma, *qbeta;
};
__device__ __forceinline__ float bn_apply1(float v, uint32_t c, const ChanParams& p) {
    float t = __fmul_rn(__fsub_rn(v, __ldg(p.mean + c)), __ldg(p.qscale + c));
    if (p.qgamma) t = __fmul_rn(t, __ldg(p.qgamma + c));
    if (p.qbeta) t = __fadd_rn(t, __ldg(p.qbeta + c));
    return t;
}

template <bool UNIFORM_QUAD>
__global__ void __launch_bounds__(RB_THREADS)
k_rangebn_apply(const float* __restrict__ x, float* out, uint32_t n, FastDiv dHW, FastDiv dC, ChanParams p) {
    const uint32_t nquads = n >> 2;
    const uint32_t stride = gridDim.x * RB_THREADS;
    for (uint32_t i = blockIdx.x * RB_THREADS + threadIdx.x; i < nquads; i += stride) {
        float4 v = ld_stream4(x + 4ull * i);
        const uint32_t e = 4u * i;
        float4 r;
        if (UNIFORM_QUAD) {   // HW % 4 == 0: the four elements share a channel
            uint32_t plane = fdiv(e, dHW);
            uint32_t c = plane - fdiv(plane, dC) * dC.d;
            const float m = __ldg(p.mean + c), s = __ldg(p.qscale + c);
            const float g = p.qgamma ? __ldg(p.qgamma + c) : 1.0f, b = p.qbeta ? __ldg(p.qbeta + c) : 0.0f;
            r.x = __fmul_rn(__fsub_rn(v.x, m), s); r.y = __fmul_rn(__fsub_rn(v.y, m), s);
            r.z = __fmul_rn(__fsub_rn(v.z, m), s); r.w = __fmul_rn(__fsub_rn(v.w, m), s);
            if (p.qgamma) { r.x = __fmul_rn(r.x, g); r.y = __fmul_rn(r.y, g); r.z = __fmul_rn(r.z, g); r.w = __fmul_rn(r.w, g); }
            if (p.qbeta) { r.x = __fadd_rn(r.x, b); r.y = __fadd_rn(r.y, b); r.z = __fadd_rn(r.z, b); r.w = __fadd_rn(r.w, b); }
        } else {
            uint32_t pl[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) { uint32_t plane = fdiv(e + j, dHW); pl[j] = plane - fdiv(plane, dC) * dC.d; }
            r.x = bn_apply1(v.x, pl[0], p); r.y = bn_apply1(v.y, pl[1], p);
            r.z = bn_apply1(v.z, pl[2], p); r.w = bn_apply1(v.w, pl[3], p);
        }
        st_stream4(out + 4ull * i, r);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        uint32_t e = (nquads << 2) + threadIdx.x;
        uint32_t plane = fdiv(e, dHW);
        out[e] = bn_apply1(x[e], plane - fdiv(plane, dC) * dC.d, p);
    }
}

static int rb_grid(int64_t items) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (items + RB_THREADS - 1) / RB_THREADS, cap = (int64_t)sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

extern "C" int qb_rangebn_apply(const float* x, float* out, int64_t B, int64_t C, int64_t HW, const float* mean,
                                const float* qscale, const float* qgamma, const float* qbeta, void* stream) {
    QB_CHECK_ARG(x && out && mean && qscale, "qb_rangebn_apply: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_apply: bad shape");
    const int64_t n = B * C * HW;
    if (n >= 4294967296LL) {
        qb_set_error("qb_rangebn_apply: %lld elements >= 2^32 not supported", (long long)n);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "qb_rangebn_apply: x/out must be 16-byte aligned");
    ChanParams p{mean, qscale, qgamma, qbeta};
    FastDiv dHW = make_fastdiv((uint32_t)HW), dC = make_fastdiv((uint32_t)C);
    const int grid = rb_grid(n >> 2);
    if (HW % 4 == 0)
        k_rangebn_apply<true><<<grid, RB_THREADS, 0, qb_stream(stream)>>>(x, out, (uint32_t)n, dHW, dC, p);
    else
        k_rangebn_apply<false><<<grid, RB_THREADS, 0, qb_stream(stream)>>>(x, out, (uint32_t)n, dHW, dC, p);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// backward.  With xq the quantized input, gq the quantized output gradient, N = count per channel:
//   out = ((xq - mu) * s^) * g^ + b^ ,  s^ = Q(r), r = 1/((M - m) * fix + eps), M/m = mean of chunk max/min
//   dbeta = sum gq ; dgamma = s^ (sum gq xq - mu sum gq) ; ds^ = g^ (sum gq xq - mu sum gq)
//   dx = (gq * g^) * s^  +  a  (+ e at the first arg-max of each chunk, - e at the first arg-min)
//   a = -(s^ g^ sum gq) / N ,  e = -(fix * r^2 * ds^) / K
// ---------------------------------------------------------------------------------------------------
constexpr int RB_BWD_SPLITS = 64;

extern "C" size_t qb_rangebn_bwd_ws_bytes(int64_t C) { return C < 0 ? 0 : (size_t)C * RB_BWD_SPLITS * 16; }

template <int G>
__global__ void __launch_bounds__(RB_THREADS)
k_rangebn_bwd_reduce(const float* __restrict__ gq, const float* __restrict__ x, int64_t C, int64_t HW, int S, int64_t N,
                     int64_t Ls, int vec_ok, double* part /* [2][C*S] */) {
    constexpr int TPB = RB_THREADS / G;
    const int lane = threadIdx.x % G;
    const int64_t task = (int64_t)blockIdx.x * TPB + threadIdx.x / G;
    const int64_t ntasks = C * S;
    const bool active = task < ntasks;
    const int64_t c = active ? task / S : 0, k = active ? task % S : 0;
    double dg = 0.0, dgx = 0.0;
    const int64_t e0 = k * Ls, e1 = (e0 + Ls < N) ? e0 + Ls : N;
    if (active && e0 < e1) {
        float fg = 0.f, fgx = 0.f;
        int cnt = 0;
        auto flush = [&]() { dg += (double)fg; dgx += (double)fgx; fg = 0.f; fgx = 0.f; cnt = 0; };
        walk_task<G>(x, C, HW, c, e0, e1, lane, vec_ok != 0,
            [&](int64_t addr, int64_t) {
                float4 a = __ldcs(reinterpret_cast<const float4*>(gq + addr));
                float4 b = __ldcs(reinterpret_cast<const float4*>(x + addr));
                fg += (a.x + a.y) + (a.z + a.w);
                fgx += (a.x * b.x + a.y * b.y) + (a.z * b.z + a.w * b.w);
                if (++cnt == 8) flush();
            },
            [&](int64_t addr, int64_t) {
                float a = gq[addr], b = x[addr];
                fg += a;
                fgx += a * b;
                if (++cnt == 32) flush();
            });
        flush();
    }
    dg = warp_sum(dg);
    dgx = warp_sum(dgx);
    if (G > 32) {
        __shared__ double ss[2][RB_THREADS / 32];
        const int w = threadIdx.x >> 5;
        if ((threadIdx.x & 31) == 0) { ss[0][w] = dg; ss[1][w] = dgx; }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int j = 1; j < RB_THREADS / 32; ++j) { dg += ss[0][j]; dgx += ss[1][j]; }
    }
    if (active && lane == 0) {
        part[task] = dg;
        part[ntasks + task] = dgx;
    }
}

__global__ void k_rangebn_bwd_pack(const double* part, int64_t C, int S, double* sg, double* sgx) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double a = 0.0, b = 0.0;
    for (int k = 0; k < S; ++k) { a += part[c * S + k]; b += part[C * S + c * S + k]; }
    sg[c] = a;
    sgx[c] = b;
}

extern "C" int qb_rangebn_bwd_reduce(const float* gq, const float* x, int64_t B, int64_t C, int64_t HW, double* sg,
                                     double* sgx, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(gq && x && sg && sgx && ws, "qb_rangebn_bwd_reduce: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_reduce: bad shape");
    if (ws_bytes < qb_rangebn_bwd_ws_bytes(C)) {
        qb_set_error("qb_rangebn_bwd_reduce: workspace %zu < %zu bytes", ws_bytes, qb_rangebn_bwd_ws_bytes(C));
        return QB_ERR_WORKSPACE;
    }
    const int64_t N = B * HW;
    // enough (c, split) tasks to fill the machine, each at least ~4096 elements
    int64_t S = (2368 + C - 1) / C;
    if (S > RB_BWD_SPLITS) S = RB_BWD_SPLITS;
    if (S * 4096 > N) S = N / 4096;
    if (S < 1) S = 1;
    const int64_t Ls = (N + S - 1) / S;
    const int64_t ntasks = C * S;
    const int vec_ok = (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gq)) & 15) == 0) && (HW % 4 == 0);
    double* part = static_cast<double*>(ws);
    cudaStream_t st = qb_stream(stream);
    if (Ls >= 4096)
        k_rangebn_bwd_reduce<RB_THREADS><<<(unsigned)ntasks, RB_THREADS, 0, st>>>(gq, x, C, HW, (int)S, N, Ls, vec_ok, part);
    else
        k_rangebn_bwd_reduce<32><<<(unsigned)((ntasks + 7) / 8), RB_THREADS, 0, st>>>(gq, x, C, HW, (int)S, N, Ls, vec_ok, part);
    QB_LAUNCH_CHECK();
    k_rangebn_bwd_pack<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(part, C, (int)S, sg, sgx);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

__global__ void k_rangebn_bwd_coeffs(const double* sg, const double* sgx, int64_t C, RbnCoefArg a) {
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    rbn_bwd_coeffs_channel(sg[c], sgx[c], c, a);
}

extern "C" int qb_rangebn_bwd_coeffs(const double* sg, const double* sgx, int64_t C, int num_chunks, double count,
                                     double chunk_len, int training, const float* mean, const float* scale_raw,
                                     const float* qscale, const float* qgamma, float* dgamma, float* dbeta, float* coef_a,
                                     float* coef_e, void* stream) {
    QB_CHECK_ARG(sg && sgx && mean && scale_raw && qscale && coef_a && coef_e && C > 0, "qb_rangebn_bwd_coeffs: null pointer");
    QB_CHECK_ARG(!training || (num_chunks > 0 && count > 0 && chunk_len > 1), "qb_rangebn_bwd_coeffs: bad chunking");
    const RbnCoefArg a{(double)num_chunks, count, training ? rbn_scale_fix(chunk_len) : 0.0, training, mean, scale_raw, qscale, qgamma,
                       dgamma, dbeta, coef_a, coef_e};
    k_rangebn_bwd_coeffs<<<(unsigned)((C + 127) / 128), 128, 0, qb_stream(stream)>>>(sg, sgx, C, a);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

template <bool UNIFORM_QUAD>
__global__ void __launch_bounds__(RB_THREADS)
k_rangebn_bwd_apply(const float* __restrict__ gq, float* dx, uint32_t n, FastDiv dHW, FastDiv dC, const float* qscale,
                    const float* qgamma, const float* coef_a) {
    const uint32_t nquads = n >> 2;
    const uint32_t stride = gridDim.x * RB_THREADS;
    auto one = [&](float g, uint32_t c) {
        float t = qgamma ? __fmul_rn(g, __ldg(qgamma + c)) : g;
        return __fadd_rn(__fmul_rn(t, __ldg(qscale + c)), __ldg(coef_a + c));
    };
    for (uint32_t i = blockIdx.x * RB_THREADS + threadIdx.x; i < nquads; i += stride) {
        float4 v = ld_stream4(gq + 4ull * i);
        const uint32_t e = 4u * i;
        uint32_t pl[4];
        if (UNIFORM_QUAD) {
            uint32_t plane = fdiv(e, dHW);
            pl[0] = pl[1] = pl[2] = pl[3] = plane - fdiv(plane, dC) * dC.d;
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) { uint32_t plane = fdiv(e + j, dHW); pl[j] = plane - fdiv(plane, dC) * dC.d; }
        }
        st_stream4(dx + 4ull * i, make_float4(one(v.x, pl[0]), one(v.y, pl[1]), one(v.z, pl[2]), one(v.w, pl[3])));
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        uint32_t e = (nquads << 2) + threadIdx.x;
        uint32_t plane = fdiv(e, dHW);
        dx[e] = one(gq[e], plane - fdiv(plane, dC) * dC.d);
    }
}

// 2*K read-modify-writes per channel: the gradient of the chunk max / min goes to ONE element each
__global__ void k_rangebn_bwd_scatter(float* dx, int64_t C, int64_t HW, int K, const float* coef_e, const int32_t* amax,
                                      const int32_t* amin) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= C * K) return;
    const int64_t c = t / K;
    const float e = coef_e[c];
    int64_t i = amax[t];
    int64_t b = i / HW, p = i - b * HW;
    float* q = dx + (b * C + c) * HW + p;
    *q = __fadd_rn(*q, e);
    i = amin[t];
    b = i / HW;
    p = i - b * HW;
    q = dx + (b * C + c) * HW + p;
    *q = __fsub_rn(*q, e);
}

extern "C" int qb_rangebn_bwd_apply(const float* gq, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                                    const float* qscale, const float* qgamma, const float* coef_a, const float* coef_e,
                                    const int32_t* amax, const int32_t* amin, void* stream) {
    QB_CHECK_ARG(gq && dx && qscale && coef_a, "qb_rangebn_bwd_apply: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_apply: bad shape");
    const int64_t n = B * C * HW;
    if (n >= 4294967296LL) {
        qb_set_error("qb_rangebn_bwd_apply: %lld elements >= 2^32 not supported", (long long)n);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(((reinterpret_cast<uintptr_t>(gq) | reinterpret_cast<uintptr_t>(dx)) & 15) == 0,
                 "qb_rangebn_bwd_apply: gq/dx must be 16-byte aligned");
    FastDiv dHW = make_fastdiv((uint32_t)HW), dC = make_fastdiv((uint32_t)C);
    cudaStream_t st = qb_stream(stream);
    const int grid = rb_grid(n >> 2);
    if (HW % 4 == 0)
        k_rangebn_bwd_apply<true><<<grid, RB_THREADS, 0, st>>>(gq, dx, (uint32_t)n, dHW, dC, qscale, qgamma, coef_a);
    else
        k_rangebn_bwd_apply<false><<<grid, RB_THREADS, 0, st>>>(gq, dx, (uint32_t)n, dHW, dC, qscale, qgamma, coef_a);
    QB_LAUNCH_CHECK();
    if (coef_e && amax && amin && num_chunks > 0) {
        const int64_t t = C * num_chunks;
        k_rangebn_bwd_scatter<<<(unsigned)((t + 127) / 128), 128, 0, st>>>(dx, C, HW, num_chunks, coef_e, amax, amin);
        QB_LAUNCH_CHECK();
    }
    return QB_OK;
}
