// RangeBN per-channel finalize (quantize.py:470-498) shared by rangebn.cu (stand-alone kernel, qb_rangebn_finalize) and
// rangebn_nhwc.cu (the same arithmetic at the tail of the channels-last statistics pass: one launch less per layer).
#pragma once
#include <math.h>

#include "common.cuh"

// 256 threads: the per-channel loops are short, and every idle warp that reaches the double-precision range arithmetic
// below costs real time on this part (fp64 issues at 1/64 rate) - only threads that own a channel compute it
constexpr int RF_THREADS = 256;

struct RbnFinalizeArg {
    double K, count, scale_fix, momentum;
    float eps;
    int training, bits;
    const float *gamma, *beta;
    float *running_mean, *running_var, *mean_out, *scale_raw, *qscale, *qgamma, *qbeta;
};

// quantize.py:473-474, in Python double
static inline double rbn_scale_fix(double chunk_len) {
    return (0.5 * 0.35) * (1.0 + sqrt(3.14159265358979323846 * log(4.0))) / sqrt(2.0 * log(chunk_len));
}

__device__ __forceinline__ void block_minmax3(float (&mn)[3], float (&mx)[3], float* sm /* 6*32 */) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = RF_THREADS / 32;
#pragma unroll
    for (int j = 0; j < 3; ++j) { mn[j] = warp_min(mn[j]); mx[j] = warp_max(mx[j]); }
    if (l == 0) {
#pragma unroll
        for (int j = 0; j < 3; ++j) { sm[j * 32 + w] = mn[j]; sm[(3 + j) * 32 + w] = mx[j]; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        mn[j] = warp_min(l < nw ? sm[j * 32 + l] : QB_INF);
        mx[j] = warp_max(l < nw ? sm[(3 + j) * 32 + l] : -QB_INF);
    }
}

// One CTA of RF_THREADS threads.  red = [3][C] doubles (sum of chunk maxima, sum of chunk minima, sum of x); thread t reads
// the channels t, t + RF_THREADS, ... - a caller that has just WRITTEN red with the same thread-to-channel map needs no fence.
__device__ __forceinline__ void rbn_finalize_body(const double* red, int64_t C, const RbnFinalizeArg& a, float* sm /* 6*32 */) {
    const float fix_f = (float)a.scale_fix, m_f = (float)a.momentum, om_f = (float)(1.0 - a.momentum);
    float mn[3] = {QB_INF, QB_INF, QB_INF}, mx[3] = {-QB_INF, -QB_INF, -QB_INF};
    for (int64_t c = threadIdx.x; c < C; c += RF_THREADS) {
        float mean, r;
        if (a.training) {
            const float mean_max = (float)(red[c] / a.K), mean_min = (float)(red[C + c] / a.K);
            mean = (float)(red[2 * C + c] / a.count);
            const float u = __fadd_rn(__fmul_rn(__fsub_rn(mean_max, mean_min), fix_f), a.eps);
            r = __fdiv_rn(1.0f, u);
            a.running_mean[c] = __fadd_rn(__fmul_rn(a.running_mean[c], m_f), __fmul_rn(mean, om_f));
            a.running_var[c] = __fadd_rn(__fmul_rn(a.running_var[c], m_f), __fmul_rn(r, om_f));
        } else {
            mean = a.running_mean[c];
            r = a.running_var[c];
        }
        a.mean_out[c] = mean;
        a.scale_raw[c] = r;
        mn[0] = fminf(mn[0], r); mx[0] = fmaxf(mx[0], r);
        if (a.gamma) { float g = a.gamma[c]; mn[1] = fminf(mn[1], g); mx[1] = fmaxf(mx[1], g); }
        if (a.beta) { float b = a.beta[c]; mn[2] = fminf(mn[2], b); mx[2] = fmaxf(mx[2], b); }
    }
    block_minmax3(mn, mx, sm);
    if ((int64_t)threadIdx.x >= C) return;                                     // no channel: skip the fp64 range arithmetic
    const QParams ps = make_qparams_d((double)mn[0], (double)mx[0], a.bits);   // float(scale.min()), float(scale.max())
    const QParams pg = a.gamma ? make_qparams_d((double)mn[1], (double)mx[1], a.bits) : ps;   // float(weight.min()), float(weight.max())
    const QParams pb = a.beta ? make_qparams_f(mn[2], mx[2], a.bits) : ps;     // quantize(bias): tensor range from data
    for (int64_t c = threadIdx.x; c < C; c += RF_THREADS) {
        a.qscale[c] = q_affine_deq(q_affine_code(a.scale_raw[c], ps), ps);
        if (a.gamma) a.qgamma[c] = q_affine_deq(q_affine_code(a.gamma[c], pg), pg);
        if (a.beta) a.qbeta[c] = q_affine_deq(q_affine_code(a.beta[c], pb), pb);
    }
}

// Per-channel coefficients of RangeBN's closed-form backward from the two reduced sums (autograd of quantize.py:467-499).
struct RbnCoefArg {
    double K, count, scale_fix;
    int training;
    const float *mean, *scale_raw, *qscale, *qgamma;
    float *dgamma, *dbeta, *coef_a, *coef_e;
};

__device__ __forceinline__ void rbn_bwd_coeffs_channel(double g, double gx, int64_t c, const RbnCoefArg& a) {
    const double mu = (double)a.mean[c], s = (double)a.qscale[c];
    const double gam = a.qgamma ? (double)a.qgamma[c] : 1.0;
    const double centered = gx - mu * g;                 // sum gq (xq - mu)
    if (a.dbeta) a.dbeta[c] = (float)g;
    if (a.dgamma) a.dgamma[c] = (float)(s * centered);
    if (a.training) {
        const double r = (double)a.scale_raw[c];
        a.coef_a[c] = (float)(-(s * gam * g) / a.count);
        a.coef_e[c] = (float)(-(a.scale_fix * r * r * gam * centered) / a.K);
    } else {
        a.coef_a[c] = 0.0f;
        a.coef_e[c] = 0.0f;
    }
}
