// measure.cu - QuantMeasure statistics in one pass (4 B/elem) and the device-side buffer update.
//
// Reference: models/modules/quantize.py:221-246.  There: min over each sample, mean over the batch;
// the same for max; global mean; global unbiased std - four full reads of the tensor, then
// float(min), float(max) host syncs.  Here: one read, per-sample extrema combined through order-
// independent integer atomics, sums in fp64 with a fixed combination order, result left on the device.
#include "common.cuh"

constexpr int MS_THREADS = 256;
constexpr int MS_VEC_PER_THREAD = 8;                               // float4 loads per thread per tile
constexpr int64_t MS_TILE = MS_THREADS * 4 * MS_VEC_PER_THREAD;    // 8192 elements
constexpr int MS_MAX_BLOCKS = QB_MS_MAX_BLOCKS;

// workspace: [counter, 64 B] [double psum[MS_MAX_BLOCKS]] [double psq[MS_MAX_BLOCKS]] [uint32 row_max_enc[B]]
// [uint32 row_min_enc[B]].  The fixed-size part comes first so that one workspace can serve calls with different
// B: only the counter and the row slots must be zero on entry (the kernel re-zeroes them), and their position
// never overlaps the partial-sum area whatever B was used before.
static const size_t MS_FIXED_BYTES = QB_MS_FIXED_BYTES;
extern "C" size_t qb_measure_ws_bytes(int64_t B) {
    if (B < 0) return 0;
    return MS_FIXED_BYTES + (size_t)B * 8;
}

__device__ __forceinline__ void acc4(float4 v, float& mn, float& mx, float& s, float& q) {
    mn = fminf(fminf(mn, fminf(v.x, v.y)), fminf(v.z, v.w));
    mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
    s += (v.x + v.y) + (v.z + v.w);
    q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
}

// RELU: y = relu(x) is written (y may alias x) and the statistics are those of y - the glue ReLU between the modules
//  then delivers the next QuantMeasure's statistics for free.  NaNs propagate like ATen's relu.
__device__ __forceinline__ float relu1(float v) { return v < 0.0f ? 0.0f : v; }

// MODE 0: statistics of x; 1: y = relu(x) + statistics of y; 2: y = relu(x + r) + statistics of y (the residual add and
// the ReLU that close a ResNet block, models/resnet_quantized.py:62-68, in one pass)
// MODE 3 / 4 (qb_rangebn_apply_relu_measure): x is NOT an fp32 tensor but the uint8 codes of RangeBN's fake-quantized input
// (channels-last, written by the statistics pass); the element is first dequantised and normalised exactly as
// k_rbn_apply_codes_nhwc does it (quantize.py:488-499, four separately rounded fp32 operations), then MODE 3 = MODE 1 and
// MODE 4 = MODE 2 on that value: the normalised tensor itself is never written.
struct ApplyArg {
    const float *mean, *qscale, *qgamma, *qbeta;   // per channel; qgamma / qbeta may be null
    const float* q_range;
    int q_bits;
    int C;
    int fixed;                                      // (MS_THREADS*4) % C == 0: a thread's four channels never change
    uint8_t* mask_out;                              // optional: bit e%8 of byte e/8 = (y[e] is not <= 0), what the ReLU backward needs
};

__device__ __forceinline__ float4 apply_quad(uint32_t word, const AffineQ& aq, const float4& mu, const float4& sc, const float4& gm,
                                             const float4& be, bool has_g, bool has_b) {
    float4 o;
    float t;
#define QB_APPLY_ONE(dst, sh, f)                                                              \
    t = __fmul_rn(__fsub_rn(q_affine_deq(code_byte_to_float<sh / 8>(word), aq.p), mu.f), sc.f);  \
    if (has_g) t = __fmul_rn(t, gm.f);                                                          \
    if (has_b) t = __fadd_rn(t, be.f);                                                          \
    dst = t;
    QB_APPLY_ONE(o.x, 0, x)
    QB_APPLY_ONE(o.y, 8, y)
    QB_APPLY_ONE(o.z, 16, z)
    QB_APPLY_ONE(o.w, 24, w)
#undef QB_APPLY_ONE
    return o;
}

template <int MODE>
__global__ void __launch_bounds__(MS_THREADS)
k_measure_stats(const float* x, const float* r, float* y, int64_t B, int64_t inner, int vec, double* stats, unsigned int* row_max,
                unsigned int* row_min, double* psum, double* psq, unsigned int* counter, ApplyArg ap) {
    constexpr bool RELU = MODE >= 1, ADD = MODE == 2 || MODE == 4, CODES = MODE >= 3;
    AffineQ aq;
    float4 mu = make_float4(0.f, 0.f, 0.f, 0.f), sc = mu, gm = mu, be = mu;
    const bool has_g = CODES && ap.qgamma != nullptr, has_b = CODES && ap.qbeta != nullptr;
    auto load_params = [&](int c) {
        mu = __ldg(reinterpret_cast<const float4*>(ap.mean + c));
        sc = __ldg(reinterpret_cast<const float4*>(ap.qscale + c));
        if (has_g) gm = __ldg(reinterpret_cast<const float4*>(ap.qgamma + c));
        if (has_b) be = __ldg(reinterpret_cast<const float4*>(ap.qbeta + c));
    };
    if (CODES) {
        aq = load_affineq(ap.q_range, ap.q_bits);
        if (ap.fixed) load_params((int)((threadIdx.x * 4u) % (unsigned)ap.C));
    }
    __shared__ float smn[8], smx[8];
    __shared__ double sd[2][8];
    __shared__ bool last;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t tiles_per_row = (inner + MS_TILE - 1) / MS_TILE;
    const int64_t T = B * tiles_per_row;
    double dsum = 0.0, dsq = 0.0;
    // a CTA owns a CONTIGUOUS range of tiles, so it crosses few sample boundaries: the block-wide min/max reduction and
    // the two atomics run once per (CTA, sample) instead of once per tile
    const int64_t per_cta = (T + gridDim.x - 1) / gridDim.x;
    const int64_t t_begin = (int64_t)blockIdx.x * per_cta, t_end = (t_begin + per_cta < T) ? t_begin + per_cta : T;
    float mn = QB_INF, mx = -QB_INF;
    int64_t cur_b = -1;
    auto flush = [&](int64_t b) {
        float a = warp_min(mn), c = warp_max(mx);
        __syncthreads();
        if (lane == 0) { smn[warp] = a; smx[warp] = c; }
        __syncthreads();
        if (threadIdx.x == 0) {
#pragma unroll
            for (int w = 1; w < MS_THREADS / 32; ++w) { a = fminf(a, smn[w]); c = fmaxf(c, smx[w]); }
            atomicMax(&row_max[b], enc_ord(c));
            atomicMax(&row_min[b], ~enc_ord(a));    // max of complemented code == min
        }
        mn = QB_INF; mx = -QB_INF;
    };
    for (int64_t t = t_begin; t < t_end; ++t) {
        const int64_t b = t / tiles_per_row, s = t - b * tiles_per_row;
        if (b != cur_b) {
            if (cur_b >= 0) flush(cur_b);
            cur_b = b;
        }
        const int64_t off = s * MS_TILE;
        const int64_t len = (inner - off < MS_TILE) ? inner - off : MS_TILE;
        const float* src = x + b * inner + off;
        float fs = 0.f, fq = 0.f;
        if (CODES && vec && ap.fixed && len == MS_TILE) {
            // whole tile, per-thread channel parameters fixed: no bounds checks, no parameter reloads, full-warp shuffles - the
            // generic body below spends ~28 instructions per element on those (ncu: issue-bound after the XU pipe was freed)
            const int64_t base = b * inner + off;
            const uint32_t* cp = reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(x) + base) + threadIdx.x;
            const float4* rp = ADD ? reinterpret_cast<const float4*>(r + base) + threadIdx.x : nullptr;
            float4* yp = reinterpret_cast<float4*>(y + base) + threadIdx.x;
            uint8_t* mp = ap.mask_out ? ap.mask_out + (base >> 3) + (threadIdx.x >> 1) : nullptr;
            uint32_t cw[MS_VEC_PER_THREAD];
            float4 rv[ADD ? MS_VEC_PER_THREAD : 1];
#pragma unroll
            for (int u = 0; u < MS_VEC_PER_THREAD; ++u) {
                cw[u] = __ldcs(cp + u * MS_THREADS);
                if (ADD) rv[u] = __ldcs(rp + u * MS_THREADS);
            }
#pragma unroll
            for (int u = 0; u < MS_VEC_PER_THREAD; ++u) {
                float4 w = apply_quad(cw[u], aq, mu, sc, gm, be, has_g, has_b);
                if (ADD) w = make_float4(__fadd_rn(w.x, rv[u].x), __fadd_rn(w.y, rv[u].y), __fadd_rn(w.z, rv[u].z), __fadd_rn(w.w, rv[u].w));
                w = make_float4(relu1(w.x), relu1(w.y), relu1(w.z), relu1(w.w));
                yp[u * MS_THREADS] = w;
                if (mp) {                                                     // warp-uniform
                    const uint32_t nib = (w.x <= 0.0f ? 0u : 1u) | (w.y <= 0.0f ? 0u : 2u) | (w.z <= 0.0f ? 0u : 4u) | (w.w <= 0.0f ? 0u : 8u);
                    const uint32_t hi = __shfl_down_sync(0xffffffffu, nib, 1);
                    if ((threadIdx.x & 1) == 0) mp[u * (MS_THREADS / 2)] = (uint8_t)(nib | (hi << 4));
                }
                acc4(w, mn, mx, fs, fq);
            }
        } else if (vec) {
            const int64_t nq = len >> 2;
            float4 v[MS_VEC_PER_THREAD];
            float4 rv[ADD ? MS_VEC_PER_THREAD : 1];
            uint32_t cw[CODES ? MS_VEC_PER_THREAD : 1];
#pragma unroll
            for (int u = 0; u < MS_VEC_PER_THREAD; ++u) {
                int64_t qi = threadIdx.x + (int64_t)u * MS_THREADS;
                if (qi < nq) {
                    if (CODES)
                        cw[u] = __ldcs(reinterpret_cast<const uint32_t*>(reinterpret_cast<const uint8_t*>(x) + b * inner + off) + qi);
                    else
                        v[u] = (RELU && x == y) ? reinterpret_cast<const float4*>(src)[qi]   // in place: the line is written next
                                                : __ldcs(reinterpret_cast<const float4*>(src) + qi);
                    if (ADD) rv[u] = __ldcs(reinterpret_cast<const float4*>(r + b * inner + off) + qi);
                }
            }
#pragma unroll
            for (int u = 0; u < MS_VEC_PER_THREAD; ++u) {
                int64_t qi = threadIdx.x + (int64_t)u * MS_THREADS;
                if (qi < nq) {
                    if (CODES) {
                        if (!ap.fixed) load_params((int)((uint64_t)(off + 4 * qi) % (uint64_t)ap.C));
                        v[u] = apply_quad(cw[u], aq, mu, sc, gm, be, has_g, has_b);
                    }
                    if (ADD) v[u] = make_float4(__fadd_rn(v[u].x, rv[u].x), __fadd_rn(v[u].y, rv[u].y),
                                                __fadd_rn(v[u].z, rv[u].z), __fadd_rn(v[u].w, rv[u].w));
                    if (RELU) {
                        v[u] = make_float4(relu1(v[u].x), relu1(v[u].y), relu1(v[u].z), relu1(v[u].w));
                        reinterpret_cast<float4*>(y + b * inner + off)[qi] = v[u];
                    }
                    if (CODES && ap.mask_out) {
                        // H*W*C % 8 == 0: quads come in (even, odd) lane pairs that are both inside the row
                        uint32_t nib = (v[u].x <= 0.0f ? 0u : 1u) | (v[u].y <= 0.0f ? 0u : 2u) | (v[u].z <= 0.0f ? 0u : 4u) |
                                       (v[u].w <= 0.0f ? 0u : 8u);
                        const uint32_t hi = __shfl_down_sync(__activemask(), nib, 1);
                        if ((qi & 1) == 0) ap.mask_out[((b * inner + off) >> 3) + (qi >> 1)] = (uint8_t)(nib | (hi << 4));
                    }
                    acc4(v[u], mn, mx, fs, fq);
                }
            }
            if ((int64_t)threadIdx.x < (len & 3)) {     // only the last tile of a row can have a tail (inner%4==0 => none)
                float e = src[(nq << 2) + threadIdx.x];
                if (ADD) e = __fadd_rn(e, r[b * inner + off + (nq << 2) + threadIdx.x]);
                if (RELU) { e = relu1(e); y[b * inner + off + (nq << 2) + threadIdx.x] = e; }
                mn = fminf(mn, e); mx = fmaxf(mx, e); fs += e; fq += e * e;
            }
        } else {
            for (int64_t j = threadIdx.x; j < len; j += MS_THREADS) {
                float e = src[j];
                if (ADD) e = __fadd_rn(e, r[b * inner + off + j]);
                if (RELU) { e = relu1(e); y[b * inner + off + j] = e; }
                mn = fminf(mn, e); mx = fmaxf(mx, e); fs += e; fq += e * e;
            }
        }
        dsum += (double)fs;
        dsq += (double)fq;
    }
    if (cur_b >= 0) flush(cur_b);
    // per-CTA partial sums, fixed slot per CTA
    dsum = warp_sum(dsum);
    dsq = warp_sum(dsq);
    if (lane == 0) { sd[0][warp] = dsum; sd[1][warp] = dsq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, c = 0.0;
        for (int w = 0; w < MS_THREADS / 32; ++w) { a += sd[0][w]; c += sd[1][w]; }
        psum[blockIdx.x] = a;
        psq[blockIdx.x] = c;
        __threadfence();
        last = (atomicAdd(counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    // ---- last CTA: combine in a fixed order, clean the workspace ----
    double smin = 0.0, smax = 0.0, tsum = 0.0, tsq = 0.0;
    float gmn = QB_INF, gmx = -QB_INF;
    for (int64_t b = threadIdx.x; b < B; b += MS_THREADS) {
        float rmx = dec_ord(__ldcg(&row_max[b]));
        float rmn = dec_ord(~__ldcg(&row_min[b]));
        row_max[b] = 0u;
        row_min[b] = 0u;
        smax += (double)rmx;
        smin += (double)rmn;
        gmn = fminf(gmn, rmn);
        gmx = fmaxf(gmx, rmx);
    }
    for (int i = threadIdx.x; i < (int)gridDim.x; i += MS_THREADS) {
        tsum += __ldcg(&psum[i]);
        tsq += __ldcg(&psq[i]);
    }
    smin = warp_sum(smin); smax = warp_sum(smax); tsum = warp_sum(tsum); tsq = warp_sum(tsq);
    gmn = warp_min(gmn); gmx = warp_max(gmx);
    __shared__ double fin[4][8];
    __shared__ float fmm[2][8];
    if (lane == 0) {
        fin[0][warp] = smin; fin[1][warp] = smax; fin[2][warp] = tsum; fin[3][warp] = tsq;
        fmm[0][warp] = gmn; fmm[1][warp] = gmx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double r[4] = {0, 0, 0, 0};
        for (int w = 0; w < MS_THREADS / 32; ++w) {
            for (int k = 0; k < 4; ++k) r[k] += fin[k][w];
            gmn = fminf(gmn, fmm[0][w]);
            gmx = fmaxf(gmx, fmm[1][w]);
        }
        stats[0] = r[0]; stats[1] = r[1]; stats[2] = r[2]; stats[3] = r[3];
        stats[4] = (double)gmn; stats[5] = (double)gmx;
        stats[6] = (double)B * (double)inner; stats[7] = (double)B;
        *counter = 0u;
    }
}

static int measure_launch(const float* x, const float* r, float* y, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes, void* stream,
                          const ApplyArg* ap = nullptr);

extern "C" int qb_measure_stats(const float* x, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes,
                                void* stream) {
    return measure_launch(x, nullptr, nullptr, B, inner, stats, ws, ws_bytes, stream);
}

/* y = relu(x) (y may be x: in place) plus the QuantMeasure statistics of y in the same pass. */
extern "C" int qb_relu_measure(const float* x, float* y, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes,
                               void* stream) {
    QB_CHECK_ARG(y, "qb_relu_measure: null output");
    return measure_launch(x, nullptr, y, B, inner, stats, ws, ws_bytes, stream);
}

/* y = relu(x + r) (y may be x: in place; r must not alias y) plus the QuantMeasure statistics of y in the same pass. */
extern "C" int qb_add_relu_measure(const float* x, const float* r, float* y, int64_t B, int64_t inner, double* stats, void* ws,
                                   size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(y && r, "qb_add_relu_measure: null pointer");
    QB_CHECK_ARG(r != y, "qb_add_relu_measure: the residual must not alias the output");
    return measure_launch(x, r, y, B, inner, stats, ws, ws_bytes, stream);
}

/* RangeBN normalisation (quantize.py:488-499) + the ReLU that follows it [+ the residual add in front of that ReLU] + the
 * QuantMeasure statistics of the result in ONE pass over the uint8 codes of RangeBN's fake-quantized input:
 *   y = relu(((deq(code) - mean_c) * qscale_c) * qgamma_c + qbeta_c [+ residual])
 * 1 (+4) B/elem read, 4 written - the normalised tensor is never materialised (qb_rangebn_apply_nhwc + qb_relu_measure /
 * qb_add_relu_measure move 5 + 8 (+4)).  Bit-identical to that composition, statistics included.  Channels-last, C % 4 == 0,
 * H*W*C % 4 == 0, 16-byte aligned tensors, q_bits <= 8. */
extern "C" int qb_rangebn_apply_relu_measure(const uint8_t* x_codes, const float* residual, float* y, int64_t B, int64_t C, int64_t HW,
                                             const float* mean, const float* qscale, const float* qgamma, const float* qbeta,
                                             const float* q_range, int q_bits, uint8_t* relu_bits_out, double* stats, void* ws,
                                             size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x_codes && y && mean && qscale && q_range, "qb_rangebn_apply_relu_measure: null pointer");
    QB_CHECK_ARG(!relu_bits_out || (HW * C) % 8 == 0, "qb_rangebn_apply_relu_measure: the ReLU bit mask needs H*W*C %% 8 == 0");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0 && C % 4 == 0 && C < (1 << 30), "qb_rangebn_apply_relu_measure: bad shape (C %% 4 == 0 required)");
    QB_CHECK_ARG(q_bits >= 1 && q_bits <= 8, "qb_rangebn_apply_relu_measure: q_bits=%d", q_bits);
    QB_CHECK_ARG(((reinterpret_cast<uintptr_t>(x_codes) & 3) | (reinterpret_cast<uintptr_t>(y) & 15) |
                  (reinterpret_cast<uintptr_t>(residual) & 15) | (reinterpret_cast<uintptr_t>(mean) & 15) |
                  (reinterpret_cast<uintptr_t>(qscale) & 15) | (reinterpret_cast<uintptr_t>(qgamma) & 15) |
                  (reinterpret_cast<uintptr_t>(qbeta) & 15)) == 0,
                 "qb_rangebn_apply_relu_measure: misaligned tensor");
    QB_CHECK_ARG(residual != y, "qb_rangebn_apply_relu_measure: the residual must not alias the output");
    const int64_t inner = HW * C;
    ApplyArg ap{mean, qscale, qgamma, qbeta, q_range, q_bits, (int)C, 0, relu_bits_out};
    ap.fixed = ((MS_THREADS * 4) % C == 0 && MS_TILE % C == 0) ? 1 : 0;
    return measure_launch(reinterpret_cast<const float*>(x_codes), residual, y, B, inner, stats, ws, ws_bytes, stream, &ap);
}

static int measure_launch(const float* x, const float* r, float* y, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes, void* stream,
                          const ApplyArg* ap) {
    QB_CHECK_ARG(x && stats && ws, "qb_measure_stats: null pointer");
    QB_CHECK_ARG(B > 0 && inner > 0, "qb_measure_stats: B=%lld inner=%lld (empty input has no range)", (long long)B,
                 (long long)inner);
    if (ws_bytes < qb_measure_ws_bytes(B)) {
        qb_set_error("qb_measure_stats: workspace %zu < %zu bytes", ws_bytes, qb_measure_ws_bytes(B));
        return QB_ERR_WORKSPACE;
    }
    char* base = static_cast<char*>(ws);
    unsigned int* counter = reinterpret_cast<unsigned int*>(base);
    double* psum = reinterpret_cast<double*>(base + 64);
    double* psq = psum + MS_MAX_BLOCKS;
    unsigned int* row_max = reinterpret_cast<unsigned int*>(base + MS_FIXED_BYTES);
    unsigned int* row_min = row_max + B;
    const int vec = ap ? 1 : (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(r)) & 15) == 0) && (inner % 4 == 0);
    const int64_t T = B * ((inner + MS_TILE - 1) / MS_TILE);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t cap = (int64_t)sms * 8;
    if (cap > MS_MAX_BLOCKS) cap = MS_MAX_BLOCKS;
    int grid = (int)(T < cap ? T : cap);
    const ApplyArg none{};
    if (ap && r) k_measure_stats<4><<<grid, MS_THREADS, 0, qb_stream(stream)>>>(x, r, y, B, inner, vec, stats, row_max, row_min, psum, psq, counter, *ap);
    else if (ap) k_measure_stats<3><<<grid, MS_THREADS, 0, qb_stream(stream)>>>(x, nullptr, y, B, inner, vec, stats, row_max, row_min, psum, psq, counter, *ap);
    else if (y && r) k_measure_stats<2><<<grid, MS_THREADS, 0, qb_stream(stream)>>>(x, r, y, B, inner, vec, stats, row_max, row_min, psum, psq, counter, none);
    else if (y) k_measure_stats<1><<<grid, MS_THREADS, 0, qb_stream(stream)>>>(x, nullptr, y, B, inner, vec, stats, row_max, row_min, psum, psq, counter, none);
    else k_measure_stats<0><<<grid, MS_THREADS, 0, qb_stream(stream)>>>(x, nullptr, nullptr, B, inner, vec, stats, row_max, row_min, psum, psq, counter, none);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// QuantMeasure buffer update + range selection (quantize.py:216-219, 228-246, 256-268), one thread.
// ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ void blend(float* running, float value, bool cumulative, float m_cum, float m_f, float om_f) {
    // running.mul_(momentum).add_(new_value * (1 - momentum))
    if (cumulative) {
        float om = __fsub_rn(1.0f, m_cum);
        *running = __fadd_rn(__fmul_rn(*running, m_cum), __fmul_rn(value, om));
    } else {
        *running = __fadd_rn(__fmul_rn(*running, m_f), __fmul_rn(value, om_f));
    }
}

__device__ __forceinline__ void aciq_range(float& std_ref, float mean, float tmin, float tmax, float alpha, float& lo,
                                           float& hi) {
    std_ref = __fadd_rn(std_ref, 1e-8f);                      // std += 1e-8 (in place in the reference)
    float clip = __fmul_rn(std_ref, alpha);
    float half_range = __fdiv_rn(__fsub_rn(tmax, tmin), 2.0f);
    clip = half_range < clip ? half_range : clip;            // Python min(max_range/2, clip_val)
    float a = __fsub_rn(mean, clip), b = __fadd_rn(mean, clip);
    lo = a > tmin ? a : tmin;                                // Python max(tmin, mean-clip): first arg wins ties
    hi = b < tmax ? b : tmax;
}

__global__ void k_measure_finalize(const double* stats, float* rmin, float* rmax, float* nmeas, float* rvar, float* rmean,
                                   double momentum, int training, int method, float alpha, float* range_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    float lo, hi;
    if (training) {
        const double B = stats[7], n = stats[6];
        lo = (float)(stats[0] / B);
        hi = (float)(stats[1] / B);
        const double mean_d = stats[2] / n;
        double var_d = (stats[3] - stats[2] * mean_d) / (n - 1.0);       // unbiased; n==1 -> nan like torch
        if (var_d < 0.0) var_d = 0.0;
        const float mean = (float)mean_d;
        float stdv = (float)sqrt(var_d);
        const bool cumulative = momentum < 0.0;
        const float cnt = *nmeas;
        const float m_cum = __fdiv_rn(cnt, __fadd_rn(cnt, 1.0f));
        const float m_f = (float)momentum, om_f = (float)(1.0 - momentum);
        blend(rmin, lo, cumulative, m_cum, m_f, om_f);
        blend(rmax, hi, cumulative, m_cum, m_f, om_f);
        blend(rmean, mean, cumulative, m_cum, m_f, om_f);
        blend(rvar, stdv, cumulative, m_cum, m_f, om_f);
        *nmeas = __fadd_rn(cnt, 1.0f);
        if (method == 1) aciq_range(stdv, mean, lo, hi, alpha, lo, hi);
    } else if (method == 1) {
        aciq_range(*rvar, *rmean, *rmin, *rmax, alpha, lo, hi);          // NB: mutates running_var like the reference
    } else {
        lo = *rmin;
        hi = *rmax;
    }
    range_out[0] = lo;
    range_out[1] = hi;
}

extern "C" int qb_measure_finalize(const double* stats, float* running_min, float* running_max, float* num_measurements,
                                   float* running_var, float* running_mean, double momentum, int training, int method,
                                   int bits, float* range_out, void* stream) {
    QB_CHECK_ARG(running_min && running_max && num_measurements && running_var && running_mean && range_out,
                 "qb_measure_finalize: null pointer");
    QB_CHECK_ARG(!training || stats, "qb_measure_finalize: training needs stats");
    QB_CHECK_ARG(method == 0 || method == 1, "qb_measure_finalize: method=%d", method);
    static const float alpha_tab[9] = {0.f, 0.f, 2.83f, 3.89f, 5.03f, 6.2f, 7.41f, 8.64f, 9.89f};   // quantize.py:214
    float alpha = 0.f;
    if (method == 1) {
        QB_CHECK_ARG(bits >= 2 && bits <= 8, "qb_measure_finalize: aciq not supported for %d bits", bits);
        alpha = alpha_tab[bits];
    }
    k_measure_finalize<<<1, 32, 0, qb_stream(stream)>>>(stats, running_min, running_max, num_measurements, running_var,
                                                        running_mean, momentum, training, method, alpha, range_out);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
