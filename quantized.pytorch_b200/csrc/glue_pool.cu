// glue_pool.cu - SURVEY.md section 8(f) row N1: the stock-torch glue between the quantized modules.  MaxPool2d over
// channels-last fp32 activations (models/resnet_quantized.py:150, after conv1 / bn1 / relu): torch's NHWC kernels take
// 0.86 ms forward and 1.75 ms backward on the 256x64x112x112 tensor of ResNet-18 - ten times the traffic bound.
//
// Same arithmetic as ATen (aten/src/ATen/native/cuda/DilatedMaxPool2d.cu): the window is scanned rows first, a value
// replaces the running maximum when it is GREATER (or NaN), so the FIRST maximal element wins ties - which matters
// here because post-ReLU windows are full of equal zeros; backward adds the output gradient at that position, windows
// visited in ascending (ho, wo) order.  Forward: thread = one output pixel x 8 adjacent channels (256-bit accesses),
// the window-relative argmax is kept as one byte per element (9 positions) instead of ATen's int64 index.  Backward is a
// GATHER: thread = one input pixel x 8 channels looks at the <= ceil(k/s)^2 windows that contain it - no atomics, every
// dx element written exactly once, coalesced.
#include <stdlib.h>

#include "common.cuh"

struct PoolGeom {
    int B, H, W, C, Ho, Wo, k, stride, pad;
    FastDiv d_c8, d_wo, d_ho, d_w, d_h;
};

// CODES: the pooled tensor is the output of a RangeBN that has not been materialised - x holds the uint8 codes of RangeBN's
// fake-quantized input (written by its statistics pass) and every window element is dequantised and normalised on the fly,
// with the arithmetic of k_rbn_apply_codes_nhwc (quantize.py:488-499, four separately rounded fp32 operations), so the result
// equals pooling the normalised tensor bit for bit while the 822 MB bn1 output of ResNet-18's stem is never written or read.
struct PoolApply {
    const float *mean, *qscale, *qgamma, *qbeta;   // per channel; qgamma / qbeta may be null
    const float* q_range;
    int q_bits;
};

template <bool CODES>
__global__ void __launch_bounds__(256) k_maxpool_fwd_nhwc(const void* __restrict__ xin, float* __restrict__ out, uint8_t* __restrict__ idx,
                                                          PoolGeom g, uint32_t total, PoolApply ap) {
    const float* x = static_cast<const float*>(xin);
    const uint8_t* xc = static_cast<const uint8_t*>(xin);
    const uint32_t t = blockIdx.x * 256u + threadIdx.x;
    if (t >= total) return;
    const uint32_t c8n = (uint32_t)g.C >> 3;
    uint32_t q = fdiv(t, g.d_c8);
    const uint32_t c0 = (t - q * c8n) << 3;
    uint32_t q2 = fdiv(q, g.d_wo);
    const int wo = (int)(q - q2 * (uint32_t)g.Wo);
    const uint32_t b = fdiv(q2, g.d_ho);
    const int ho = (int)(q2 - b * (uint32_t)g.Ho);
    const int h0 = ho * g.stride - g.pad, w0 = wo * g.stride - g.pad;
    float mx[8];
    uint32_t am[8];
    bool first = true;
#pragma unroll
    for (int j = 0; j < 8; ++j) { mx[j] = -QB_INF; am[j] = 0; }
    AffineQ aq;
    float mu[8], sc[8], gm[8], be[8];
    if (CODES) {
        aq = load_affineq(ap.q_range, ap.q_bits);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            mu[j] = __ldg(ap.mean + c0 + j);
            sc[j] = __ldg(ap.qscale + c0 + j);
            gm[j] = ap.qgamma ? __ldg(ap.qgamma + c0 + j) : 1.0f;
            be[j] = ap.qbeta ? __ldg(ap.qbeta + c0 + j) : 0.0f;
        }
    }
    // CODES, fast form.  The normalisation is a monotone map of the code (every rounded fp32 operation is monotone in its
    // operand), so the window maximum of the normalised values sits at the window maximum of the codes - or the minimum, for a
    // channel whose scale * gamma is negative - and ONE element per channel is normalised instead of nine: the window search
    // runs on packed 16-bit composites of key and tap (two channels per word).  "First maximal element wins" carries over as long
    // as distinct codes give distinct normalised values; that is checked per channel from the images of codes 0 and 255 (their
    // spacing per code step must exceed 2^-18 of the magnitudes involved, 8x the worst accumulated rounding error of the four
    // operations), a channel with gamma == 0 maps every code to one value (first valid element wins, as in ATen), and a thread
    // with any other channel takes the element-wise form below.  Without this the fused pass was instruction-bound (125
    // instructions per output) and no faster than normalise-then-pool.
    bool fast = false;
    uint2 xorm = make_uint2(0u, 0u), andm = make_uint2(0xffffffffu, 0xffffffffu);
    if (CODES) {
        fast = true;
        const float xmag = fmaxf(fabsf(aq.p.lo), fabsf(q_affine_deq(aq.p.qmax, aq.p)));
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float t0 = __fmul_rn(__fsub_rn(q_affine_deq(0.0f, aq.p), mu[j]), sc[j]);
            float t1 = __fmul_rn(__fsub_rn(q_affine_deq(aq.p.qmax, aq.p), mu[j]), sc[j]);
            if (ap.qgamma) { t0 = __fmul_rn(t0, gm[j]); t1 = __fmul_rn(t1, gm[j]); }
            if (ap.qbeta) { t0 = __fadd_rn(t0, be[j]); t1 = __fadd_rn(t1, be[j]); }
            const uint32_t byte = 0xffu << (8 * (j & 3));
            const bool flat = ap.qgamma && gm[j] == 0.0f && t0 == t1;
            const float step = fabsf(t1 - t0) / aq.p.qmax;
            const bool sep = step > 3.8146973e-6f * fmaxf(fabsf(t0), fabsf(t1)) &&
                             aq.p.scale > 3.8146973e-6f * (xmag + fabsf(mu[j]));                   // 2^-18
            if (flat) { if (j < 4) andm.x &= ~byte; else andm.y &= ~byte; }
            else if (!sep) fast = false;                                                           // also catches NaN parameters
            else if (t1 < t0) { if (j < 4) xorm.x |= byte; else xorm.y |= byte; }
        }
    }
    if (CODES && fast) {
        // per channel the 16-bit composite (key << 8) | (255 - tap): its maximum over the window is the largest key and, among
        // equal keys, the EARLIEST tap ("first maximal element wins"); two channels per word, one PRMT to build a pair from the
        // key bytes and one packed 16-bit maximum per pair and tap (the byte-wise __vmaxu4 / __vcmpgtu4 pair is emulated on this
        // architecture: ~25 instructions per tap).  Every valid tap's composite is >= 31, so 0 is the identity.
        uint32_t bp[4] = {0u, 0u, 0u, 0u};
        uint2 firstpk = make_uint2(0u, 0u);
        auto tap_in = [&](const uint2& pk, uint32_t tap) {
            const uint32_t kx = (pk.x ^ xorm.x) & andm.x, ky = (pk.y ^ xorm.y) & andm.y, tinv = 255u - tap;
            bp[0] = __vmaxu2(bp[0], __byte_perm(kx, tinv, 0x1404));
            bp[1] = __vmaxu2(bp[1], __byte_perm(kx, tinv, 0x3424));
            bp[2] = __vmaxu2(bp[2], __byte_perm(ky, tinv, 0x1404));
            bp[3] = __vmaxu2(bp[3], __byte_perm(ky, tinv, 0x3424));
            if (first) { firstpk = pk; first = false; }
        };
        if (g.k == 3) {
#pragma unroll
            for (int kh = 0; kh < 3; ++kh) {
                const int h = h0 + kh;
                if (h < 0 || h >= g.H) continue;
                const uint8_t* rowp = xc + (((size_t)b * g.H + h) * g.W) * g.C + c0;
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    const int w = w0 + kw;
                    if (w < 0 || w >= g.W) continue;
                    tap_in(__ldg(reinterpret_cast<const uint2*>(rowp + (size_t)w * g.C)), (uint32_t)(kh * 3 + kw));
                }
            }
        } else {
            for (int kh = 0; kh < g.k; ++kh) {
                const int h = h0 + kh;
                if (h < 0 || h >= g.H) continue;
                for (int kw = 0; kw < g.k; ++kw) {
                    const int w = w0 + kw;
                    if (w < 0 || w >= g.W) continue;
                    tap_in(__ldg(reinterpret_cast<const uint2*>(xc + (((size_t)b * g.H + h) * g.W + w) * g.C + c0)), (uint32_t)(kh * g.k + kw));
                }
            }
        }
        // bytes 1 and 3 of a pair word are the winning keys, bytes 0 and 2 the inverted taps
        uint2 best, bidx;
        best.x = __byte_perm(bp[0], bp[1], 0x7531);
        best.y = __byte_perm(bp[2], bp[3], 0x7531);
        bidx.x = ~__byte_perm(bp[0], bp[1], 0x6420);
        bidx.y = ~__byte_perm(bp[2], bp[3], 0x6420);
        const uint2 win = make_uint2(((best.x ^ xorm.x) & andm.x) | (firstpk.x & ~andm.x), ((best.y ^ xorm.y) & andm.y) | (firstpk.y & ~andm.y));
        float xh[8];
        affine_deq8(win, aq.p, xh);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float tt = __fmul_rn(__fsub_rn(xh[j], mu[j]), sc[j]);
            if (ap.qgamma) tt = __fmul_rn(tt, gm[j]);
            if (ap.qbeta) tt = __fadd_rn(tt, be[j]);
            mx[j] = tt;
            am[j] = ((j < 4 ? bidx.x : bidx.y) >> (8 * (j & 3))) & 0xffu;
        }
    } else
    for (int kh = 0; kh < g.k; ++kh) {
        const int h = h0 + kh;
        if (h < 0 || h >= g.H) continue;
        for (int kw = 0; kw < g.k; ++kw) {
            const int w = w0 + kw;
            if (w < 0 || w >= g.W) continue;
            const size_t e0 = (((size_t)b * g.H + h) * g.W + w) * g.C + c0;
            F8 v;
            if (CODES) {
                float xh[8];
                affine_deq8(__ldg(reinterpret_cast<const uint2*>(xc + e0)), aq.p, xh);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float tt = __fmul_rn(__fsub_rn(xh[j], mu[j]), sc[j]);
                    if (ap.qgamma) tt = __fmul_rn(tt, gm[j]);
                    if (ap.qbeta) tt = __fadd_rn(tt, be[j]);
                    v.v[j] = tt;
                }
            } else {
                v = ld256_keep(x + e0);
            }
            const uint32_t code = (uint32_t)(kh * g.k + kw);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (first) am[j] = code;                          // ATen initialises the index with the first valid element
                if (v.v[j] > mx[j] || v.v[j] != v.v[j]) { mx[j] = v.v[j]; am[j] = code; }
            }
            first = false;
        }
    }
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = mx[j];
    const size_t e = (size_t)t << 3;
    st256(out + e, o);
    uint2 pk;
    pk.x = am[0] | (am[1] << 8) | (am[2] << 16) | (am[3] << 24);
    pk.y = am[4] | (am[5] << 8) | (am[6] << 16) | (am[7] << 24);
    *reinterpret_cast<uint2*>(idx + e) = pk;
}

// FUSED: the pooled tensor is the output of a ReLU (models/resnet_quantized.py:147-150: bn1 -> relu -> maxpool), and dx is
// the gradient w.r.t. the ReLU's INPUT: a window's gradient is dropped where its maximum is not positive (the element it
// would land on has relu_out = pool_out <= 0, which is exactly where aten::threshold_backward writes 0), and {min, max} of
// dx - the range the gradient quantizer of the RangeBN in front asks for (quantize.py:126,132) - is reduced on the way:
// the 822 MB intermediate gradient of ResNet-18's stem is never written and re-read.
struct PoolRangeWs { unsigned int kmax, kmin, counter, pad; };

template <bool FUSED>
__global__ void __launch_bounds__(256) k_maxpool_bwd_nhwc(const float* __restrict__ go, const uint8_t* __restrict__ idx,
                                                          const float* __restrict__ pool_out, float* __restrict__ dx,
                                                          PoolGeom g, uint32_t total, float* range_out, PoolRangeWs* ws) {
    float mn = QB_INF, mx = -QB_INF;
    // grid-stride: the FUSED launch is a few CTAs per SM, so the range costs one reduction + atomic pair per CTA
    for (uint64_t t64 = blockIdx.x * 256u + threadIdx.x; t64 < total; t64 += (uint64_t)gridDim.x * 256u) {
    const uint32_t t = (uint32_t)t64;
    const uint32_t c8n = (uint32_t)g.C >> 3;
    uint32_t q = fdiv(t, g.d_c8);
    const uint32_t c0 = (t - q * c8n) << 3;
    uint32_t q2 = fdiv(q, g.d_w);
    const int w = (int)(q - q2 * (uint32_t)g.W);
    const uint32_t b = fdiv(q2, g.d_h);
    const int h = (int)(q2 - b * (uint32_t)g.H);
    // windows (ho, wo) with ho*stride - pad <= h <= ho*stride - pad + k - 1
    int ho_lo = (h + g.pad - g.k + g.stride) / g.stride;            // ceil((h + pad - k + 1) / stride), numerator >= 0 handled below
    if (h + g.pad - g.k + 1 <= 0) ho_lo = 0;
    int ho_hi = (h + g.pad) / g.stride;
    if (ho_hi > g.Ho - 1) ho_hi = g.Ho - 1;
    int wo_lo = (w + g.pad - g.k + g.stride) / g.stride;
    if (w + g.pad - g.k + 1 <= 0) wo_lo = 0;
    int wo_hi = (w + g.pad) / g.stride;
    if (wo_hi > g.Wo - 1) wo_hi = g.Wo - 1;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
        const uint32_t kh = (uint32_t)(h - (ho * g.stride - g.pad));
        for (int wo = wo_lo; wo <= wo_hi; ++wo) {
            const uint32_t code = kh * (uint32_t)g.k + (uint32_t)(w - (wo * g.stride - g.pad));
            const size_t e = ((((size_t)b * g.Ho + ho) * g.Wo + wo) * g.C) + c0;
            const uint2 pk = __ldg(reinterpret_cast<const uint2*>(idx + e));
            const uint32_t hit_lo = pk.x ^ (code * 0x01010101u), hit_hi = pk.y ^ (code * 0x01010101u);
            if (((hit_lo - 0x01010101u) & ~hit_lo & 0x80808080u) | ((hit_hi - 0x01010101u) & ~hit_hi & 0x80808080u)) {   // any byte equal
                const F8 gv = ld256_keep(go + e);
                F8 pv;
                if (FUSED) pv = ld256_keep(pool_out + e);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t bj = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xffu;
                    if (bj == code && (!FUSED || !(pv.v[j] <= 0.0f))) acc[j] += gv.v[j];
                }
            }
        }
    }
    F8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.v[j] = acc[j];
    st256(dx + ((size_t)t << 3), o);
    if (FUSED) {
        mn = fminf(mn, fminf(fminf(fminf(o.v[0], o.v[1]), fminf(o.v[2], o.v[3])), fminf(fminf(o.v[4], o.v[5]), fminf(o.v[6], o.v[7]))));
        mx = fmaxf(mx, fmaxf(fmaxf(fmaxf(o.v[0], o.v[1]), fmaxf(o.v[2], o.v[3])), fmaxf(fmaxf(o.v[4], o.v[5]), fmaxf(o.v[6], o.v[7]))));
    }
    }   // grid-stride loop
    if (!FUSED) return;
    // range of dx: block reduction, then one atomic pair per block and only when it can improve the running extrema
    __shared__ float s_mn[8], s_mx[8];
    __shared__ bool last;
    mn = warp_min(mn);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { s_mn[threadIdx.x >> 5] = mn; s_mx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 1; k < 8; ++k) { mn = fminf(mn, s_mn[k]); mx = fmaxf(mx, s_mx[k]); }
        const unsigned int emx = enc_ord(mx), emn = ~enc_ord(mn);
        if (emx > __ldcg(&ws->kmax)) atomicMax(&ws->kmax, emx);
        if (emn > __ldcg(&ws->kmin)) atomicMax(&ws->kmin, emn);
        __threadfence();
        last = (atomicAdd(&ws->counter, 1u) == gridDim.x - 1);
        if (last) {
            __threadfence();
            range_out[0] = dec_ord(~__ldcg(&ws->kmin));
            range_out[1] = dec_ord(__ldcg(&ws->kmax));
            ws->kmax = 0u; ws->kmin = 0u; ws->counter = 0u;
        }
    }
}

// Backward for the geometry every ResNet uses (k = 3, stride 2, pad 1), shared-memory tiled: the generic gather above re-reads
// the gradient (and pool_out) from L1/L2 once per window that MIGHT hit - about 9 GB of L2 traffic for 1.1 GB of HBM traffic on
// the 256x64x112x112 stem.  Here one CTA owns the two dx rows (2m, 2m+1) of one image: they depend on the pooled rows m
// (windows rows 2m-1..2m+1) and m+1 only.  Those two rows of the gradient - already masked with the ReLU (FUSED) - and of the
// argmax bytes are staged in shared memory with coalesced loads; every dx octet then looks at its <= 2 x 2 windows there, in
// the same ascending (ho, wo) order as the generic kernel, so the result is bit-identical.
template <bool FUSED>
__global__ void __launch_bounds__(256) k_maxpool_bwd_3s2(const float* __restrict__ go, const uint8_t* __restrict__ idx,
                                                         const float* __restrict__ pool_out, float* __restrict__ dx, PoolGeom g,
                                                         float* range_out, PoolRangeWs* ws) {
    extern __shared__ __align__(32) unsigned char pool_smem[];
    const int rowv = g.Wo * g.C;                                   // values of one pooled row
    float* s_g = reinterpret_cast<float*>(pool_smem);               // [2][Wo][C]
    uint8_t* s_i = pool_smem + (size_t)2 * rowv * 4;                // [2][Wo][C]
    const int m = blockIdx.x % g.Ho, b = blockIdx.x / g.Ho;
    const int nrows = (m + 1 < g.Ho) ? 2 : 1;
    const size_t base = ((size_t)b * g.Ho + m) * rowv;
    for (int i = threadIdx.x; i < nrows * rowv / 8; i += 256) {
        F8 gv = ld256_stream(go + base + (size_t)i * 8);
        if (FUSED) {
            const F8 pv = ld256_stream(pool_out + base + (size_t)i * 8);
#pragma unroll
            for (int j = 0; j < 8; ++j) gv.v[j] = pv.v[j] <= 0.0f ? 0.0f : gv.v[j];
        }
        *reinterpret_cast<F8*>(s_g + (size_t)i * 8) = gv;
        *reinterpret_cast<uint2*>(s_i + (size_t)i * 8) = __ldcs(reinterpret_cast<const uint2*>(idx + base + (size_t)i * 8));
    }
    __syncthreads();
    float mn = QB_INF, mx = -QB_INF;
    const int c8n = g.C >> 3;
    const int hrows = (2 * m + 1 < g.H) ? 2 : 1;
    for (int t = threadIdx.x; t < hrows * g.W * c8n; t += 256) {
        const int c0 = (t % c8n) << 3;
        const int w = (t / c8n) % g.W, dh = t / (c8n * g.W);
        const int h = 2 * m + dh;
        // row 2m lies in window row m only; row 2m+1 in window rows m and m+1 (when it exists); same for columns
        const int ho_lo = m, ho_hi = (dh && m + 1 < g.Ho) ? m + 1 : m;
        const int n = w >> 1;
        const int wo_lo = n, wo_hi = ((w & 1) && n + 1 < g.Wo) ? n + 1 : n;
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        for (int ho = ho_lo; ho <= ho_hi; ++ho) {
            const uint32_t kh = (uint32_t)(h - (2 * ho - 1));
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                const uint32_t code = kh * 3u + (uint32_t)(w - (2 * wo - 1));
                const int e = ((ho - m) * g.Wo + wo) * g.C + c0;
                const uint2 pk = *reinterpret_cast<const uint2*>(s_i + e);
                const float4 g0 = *reinterpret_cast<const float4*>(s_g + e), g1 = *reinterpret_cast<const float4*>(s_g + e + 4);
                const float gv[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t bj = ((j < 4 ? pk.x : pk.y) >> (8 * (j & 3))) & 0xffu;
                    if (bj == code) acc[j] += gv[j];
                }
            }
        }
        F8 o;
#pragma unroll
        for (int j = 0; j < 8; ++j) o.v[j] = acc[j];
        st256(dx + ((((size_t)b * g.H + h) * g.W + w) * g.C + c0), o);
        if (FUSED) {
            mn = fminf(mn, fminf(fminf(fminf(o.v[0], o.v[1]), fminf(o.v[2], o.v[3])), fminf(fminf(o.v[4], o.v[5]), fminf(o.v[6], o.v[7]))));
            mx = fmaxf(mx, fmaxf(fmaxf(fmaxf(o.v[0], o.v[1]), fmaxf(o.v[2], o.v[3])), fmaxf(fmaxf(o.v[4], o.v[5]), fmaxf(o.v[6], o.v[7]))));
        }
    }
    if (!FUSED) return;
    __shared__ float s_mn[8], s_mx[8];
    __shared__ bool last;
    mn = warp_min(mn);
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0) { s_mn[threadIdx.x >> 5] = mn; s_mx[threadIdx.x >> 5] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 1; k < 8; ++k) { mn = fminf(mn, s_mn[k]); mx = fmaxf(mx, s_mx[k]); }
        const unsigned int emx = enc_ord(mx), emn = ~enc_ord(mn);
        if (emx > __ldcg(&ws->kmax)) atomicMax(&ws->kmax, emx);
        if (emn > __ldcg(&ws->kmin)) atomicMax(&ws->kmin, emn);
        __threadfence();
        last = (atomicAdd(&ws->counter, 1u) == gridDim.x - 1);
        if (last) {
            __threadfence();
            range_out[0] = dec_ord(~__ldcg(&ws->kmin));
            range_out[1] = dec_ord(__ldcg(&ws->kmax));
            ws->kmax = 0u; ws->kmin = 0u; ws->counter = 0u;
        }
    }
}

// the tiled kernel applies: 3x3 / stride 2 / pad 1, even H and W (H = 2*Ho, W = 2*Wo), two pooled rows fit shared memory
static size_t pool_3s2_smem(const PoolGeom& g) {
    if (g.k != 3 || g.stride != 2 || g.pad != 1 || g.H != 2 * g.Ho || g.W != 2 * g.Wo) return 0;
    static const bool off = getenv("QB_POOL_TILED") && getenv("QB_POOL_TILED")[0] == '0';
    const size_t smem = (size_t)2 * g.Wo * g.C * 5;
    return (off || smem > 100 * 1024 || (int64_t)g.B * g.Ho > 2147483647LL) ? 0 : smem;
}

static int pool_geom(PoolGeom& g, const char* who, int64_t B, int64_t H, int64_t W, int64_t C, int k, int stride, int pad,
                     const void* p0, const void* p1, const void* p2) {
    QB_CHECK_ARG(p0 && p1 && p2, "%s: null pointer", who);
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0 && k >= 1 && k <= 15 && stride >= 1 && pad >= 0 && 2 * pad <= k,
                 "%s: bad geometry B=%lld H=%lld W=%lld C=%lld k=%d stride=%d pad=%d", who, (long long)B, (long long)H, (long long)W,
                 (long long)C, k, stride, pad);
    const int64_t Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    QB_CHECK_ARG(Ho > 0 && Wo > 0, "%s: empty output", who);
    if (C % 8 != 0 || B * H * W * (C / 8) >= 4294967296LL ||
        ((reinterpret_cast<uintptr_t>(p0) | reinterpret_cast<uintptr_t>(p1)) & 31) != 0 || (reinterpret_cast<uintptr_t>(p2) & 7) != 0) {
        qb_set_error("%s: needs C %% 8 == 0, < 2^32 octets and 32-byte aligned tensors", who);
        return QB_ERR_UNSUPPORTED;
    }
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.C = (int)C; g.Ho = (int)Ho; g.Wo = (int)Wo; g.k = k; g.stride = stride; g.pad = pad;
    g.d_c8 = make_fastdiv((uint32_t)(C / 8));
    g.d_wo = make_fastdiv((uint32_t)Wo); g.d_ho = make_fastdiv((uint32_t)Ho);
    g.d_w = make_fastdiv((uint32_t)W); g.d_h = make_fastdiv((uint32_t)H);
    return QB_OK;
}

extern "C" int qb_maxpool2d_nhwc_fwd(const float* x, float* out, uint8_t* idx, int64_t B, int64_t H, int64_t W, int64_t C, int k,
                                     int stride, int pad, void* stream) {
    PoolGeom g;
    int rc = pool_geom(g, "qb_maxpool2d_nhwc_fwd", B, H, W, C, k, stride, pad, x, out, idx);
    if (rc) return rc;
    const uint32_t total = (uint32_t)((int64_t)B * g.Ho * g.Wo * (C / 8));
    k_maxpool_fwd_nhwc<false><<<(total + 255u) / 256u, 256, 0, qb_stream(stream)>>>(x, out, idx, g, total, PoolApply{});
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* maxpool(RangeBN-normalise(codes)) without the normalised tensor: x_codes are the uint8 codes [B, H, W, C] the channels-last
 * RangeBN statistics pass wrote (qb_rangebn_stats_nhwc, x_codes_out), mean / qscale / qgamma / qbeta the vectors of
 * qb_rangebn_finalize, q_range / q_bits the quantizer of those codes.  out / idx as qb_maxpool2d_nhwc_fwd on the normalised
 * tensor, bit for bit (models/resnet_quantized.py:147-150: conv1 -> bn1 -> relu -> maxpool). */
extern "C" int qb_rangebn_apply_maxpool_nhwc(const uint8_t* x_codes, float* out, uint8_t* idx, int64_t B, int64_t H, int64_t W, int64_t C,
                                             int k, int stride, int pad, const float* mean, const float* qscale, const float* qgamma,
                                             const float* qbeta, const float* q_range, int q_bits, void* stream) {
    PoolGeom g;
    int rc = pool_geom(g, "qb_rangebn_apply_maxpool_nhwc", B, H, W, C, k, stride, pad, out, out, idx);
    if (rc) return rc;
    QB_CHECK_ARG(x_codes && mean && qscale && q_range, "qb_rangebn_apply_maxpool_nhwc: null pointer");
    QB_CHECK_ARG(q_bits >= 1 && q_bits <= 8, "qb_rangebn_apply_maxpool_nhwc: q_bits=%d", q_bits);
    QB_CHECK_ARG((reinterpret_cast<uintptr_t>(x_codes) & 7) == 0, "qb_rangebn_apply_maxpool_nhwc: codes must be 8-byte aligned");
    const uint32_t total = (uint32_t)((int64_t)B * g.Ho * g.Wo * (C / 8));
    const PoolApply ap{mean, qscale, qgamma, qbeta, q_range, q_bits};
    k_maxpool_fwd_nhwc<true><<<(total + 255u) / 256u, 256, 0, qb_stream(stream)>>>(x_codes, out, idx, g, total, ap);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_maxpool2d_nhwc_bwd(const float* grad_out, const uint8_t* idx, float* dx, int64_t B, int64_t H, int64_t W,
                                     int64_t C, int k, int stride, int pad, void* stream) {
    PoolGeom g;
    int rc = pool_geom(g, "qb_maxpool2d_nhwc_bwd", B, H, W, C, k, stride, pad, grad_out, dx, idx);
    if (rc) return rc;
    const uint32_t total = (uint32_t)((int64_t)B * H * W * (C / 8));
    if (const size_t smem = pool_3s2_smem(g)) {
        QB_CUDA(cudaFuncSetAttribute(k_maxpool_bwd_3s2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_maxpool_bwd_3s2<false><<<(unsigned)(B * g.Ho), 256, smem, qb_stream(stream)>>>(grad_out, idx, nullptr, dx, g, nullptr, nullptr);
        QB_LAUNCH_CHECK();
        return QB_OK;
    }
    k_maxpool_bwd_nhwc<false><<<(total + 255u) / 256u, 256, 0, qb_stream(stream)>>>(grad_out, idx, nullptr, dx, g, total, nullptr,
                                                                                   nullptr);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" size_t qb_maxpool2d_relu_bwd_ws_bytes(void) { return sizeof(PoolRangeWs); }

extern "C" int qb_maxpool2d_relu_nhwc_bwd(const float* grad_out, const uint8_t* idx, const float* pool_out, float* dx, int64_t B,
                                          int64_t H, int64_t W, int64_t C, int k, int stride, int pad, float* range_out, void* ws,
                                          size_t ws_bytes, void* stream) {
    PoolGeom g;
    int rc = pool_geom(g, "qb_maxpool2d_relu_nhwc_bwd", B, H, W, C, k, stride, pad, grad_out, dx, idx);
    if (rc) return rc;
    QB_CHECK_ARG(pool_out && range_out && ws, "qb_maxpool2d_relu_nhwc_bwd: null pointer");
    if ((reinterpret_cast<uintptr_t>(pool_out) & 31) != 0) {
        qb_set_error("qb_maxpool2d_relu_nhwc_bwd: pool_out must be 32-byte aligned");
        return QB_ERR_UNSUPPORTED;
    }
    if (ws_bytes < sizeof(PoolRangeWs)) {
        qb_set_error("qb_maxpool2d_relu_nhwc_bwd: workspace %zu < %zu bytes", ws_bytes, sizeof(PoolRangeWs));
        return QB_ERR_WORKSPACE;
    }
    if (const size_t smem = pool_3s2_smem(g)) {
        QB_CUDA(cudaFuncSetAttribute(k_maxpool_bwd_3s2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_maxpool_bwd_3s2<true><<<(unsigned)(B * g.Ho), 256, smem, qb_stream(stream)>>>(grad_out, idx, pool_out, dx, g, range_out,
                                                                                       static_cast<PoolRangeWs*>(ws));
        QB_LAUNCH_CHECK();
        return QB_OK;
    }
    const uint32_t total = (uint32_t)((int64_t)B * H * W * (C / 8));
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const uint32_t want = (total + 255u) / 256u, cap = (uint32_t)sms * 16u;
    k_maxpool_bwd_nhwc<true><<<(want < cap ? want : cap), 256, 0, qb_stream(stream)>>>(grad_out, idx, pool_out, dx, g, total, range_out,
                                                                                  static_cast<PoolRangeWs*>(ws));
    QB_LAUNCH_CHECK();
    return QB_OK;
}
