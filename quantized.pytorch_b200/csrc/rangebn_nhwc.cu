// rangebn_nhwc.cu - RangeBN kernels for channels-last activations: x viewed as [R = B*H*W rows, C columns].
//
// Channels-last is the native layout of this library (the int8 implicit-GEMM reads K = C innermost and writes
// pixel-major tiles), so the per-channel statistics of models/modules/quantize.py:467-476 become COLUMN
// reductions: thread t owns 4 adjacent channels (one 128-bit load per row), row-lanes stride over the rows of a
// chunk segment, rows are contiguous in memory, every load is fully coalesced.  The reference's chunk k of
// channel c (elements in (b,h,w) order) is simply rows [k*L, (k+1)*L) of column c; the flat (b,h,w) index the
// arg-max bookkeeping needs is the row index itself.
#include <math.h>

#include "common.cuh"

constexpr int NH_THREADS = 256;
constexpr int NH_UNROLL = 4;
constexpr int NH_MAX_SEG = 32;      // row segments per chunk (partials per (chunk, channel))

struct NhwcGeom {
    int64_t R, C;         // rows, channels
    int K;                // chunks (stats) or 1 (backward reduce)
    int64_t L;            // rows per chunk
    int S;                // segments per chunk
    int64_t seg_rows;     // rows per segment (last one shorter)
    int CT;               // column threads = ceil(min(C, tileC) / VEC)
    int RL;               // row lanes per CTA
    int tileC;            // channels per CTA (gridDim.y tiles)
};

template <int VEC>
static NhwcGeom make_geom(int64_t R, int64_t C, int K, int max_seg = NH_MAX_SEG) {
    NhwcGeom g;
    g.R = R; g.C = C; g.K = K; g.L = R / K;
    g.tileC = (int)(C < (int64_t)NH_THREADS * VEC ? C : (int64_t)NH_THREADS * VEC);
    g.CT = (g.tileC + VEC - 1) / VEC;
    g.RL = NH_THREADS / g.CT;
    int ctiles = (int)((C + g.tileC - 1) / g.tileC);
    // enough CTAs to fill 148 SMs a few times over, each segment at least RL*NH_UNROLL rows
    int64_t want = (148 * 4 + (int64_t)K * ctiles - 1) / ((int64_t)K * ctiles);
    int64_t max_by_rows = g.L / ((int64_t)g.RL * NH_UNROLL);
    if (want > max_by_rows) want = max_by_rows;
    if (want > max_seg) want = max_seg;
    if (want < 1) want = 1;
    g.S = (int)want;
    g.seg_rows = (g.L + g.S - 1) / g.S;
    return g;
}

struct ArgV { float v; int32_t i; };
__device__ __forceinline__ void amax_upd(ArgV& a, float v, int32_t i) { if (v > a.v) { a.v = v; a.i = i; } }
__device__ __forceinline__ void amin_upd(ArgV& a, float v, int32_t i) { if (v < a.v) { a.v = v; a.i = i; } }
__device__ __forceinline__ void amax_merge(ArgV& a, float v, int32_t i) { if (v > a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; } }
__device__ __forceinline__ void amin_merge(ArgV& a, float v, int32_t i) { if (v < a.v || (v == a.v && i < a.i)) { a.v = v; a.i = i; } }

template <int VEC> struct VecT;
template <> struct VecT<4> { using T = float4; };
template <> struct VecT<1> { using T = float; };
__device__ __forceinline__ void unpack(const float4& v, float (&a)[4]) { a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w; }
__device__ __forceinline__ void unpack(const float& v, float (&a)[1]) { a[0] = v; }

// ---------------------------------------------------------------------------------------------------
// forward statistics: partials per (chunk k, segment s, channel c)
// ---------------------------------------------------------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_stats_nhwc(const float* __restrict__ x, NhwcGeom g, float* pmax, float* pmin, int32_t* imax, int32_t* imin,
                 double* psum) {
    using V = typename VecT<VEC>::T;
    extern __shared__ unsigned char smem_raw[];
    const int ct = threadIdx.x % g.CT, rl = threadIdx.x / g.CT;
    const int64_t c0 = (int64_t)blockIdx.y * g.tileC + (int64_t)ct * VEC;
    const int k = blockIdx.x / g.S, s = blockIdx.x % g.S;
    const int64_t r_begin = (int64_t)k * g.L + (int64_t)s * g.seg_rows;
    int64_t r_end = r_begin + g.seg_rows;
    if (r_end > (int64_t)(k + 1) * g.L) r_end = (int64_t)(k + 1) * g.L;
    const bool active = rl < g.RL && c0 < g.C;
    ArgV mx[VEC], mn[VEC];
    double dsum[VEC];
    float fs[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) { mx[j] = {-QB_INF, 0x7fffffff}; mn[j] = {QB_INF, 0x7fffffff}; dsum[j] = 0.0; fs[j] = 0.f; }
    if (active) {
        int64_t r = r_begin + rl;
        const int64_t step = g.RL;
        int cnt = 0;
        for (; r + (NH_UNROLL - 1) * step < r_end; r += NH_UNROLL * step) {
            V v[NH_UNROLL];
#pragma unroll
            for (int u = 0; u < NH_UNROLL; ++u) v[u] = __ldcs(reinterpret_cast<const V*>(x + (r + u * step) * g.C + c0));
#pragma unroll
            for (int u = 0; u < NH_UNROLL; ++u) {
                float a[VEC];
                unpack(v[u], a);
                const int32_t idx = (int32_t)(r + u * step);
#pragma unroll
                for (int j = 0; j < VEC; ++j) { amax_upd(mx[j], a[j], idx); amin_upd(mn[j], a[j], idx); fs[j] += a[j]; }
            }
            if (++cnt == 8) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) { dsum[j] += (double)fs[j]; fs[j] = 0.f; }
                cnt = 0;
            }
        }
        for (; r < r_end; r += step) {
            float a[VEC];
            unpack(__ldcs(reinterpret_cast<const V*>(x + r * g.C + c0)), a);
#pragma unroll
            for (int j = 0; j < VEC; ++j) { amax_upd(mx[j], a[j], (int32_t)r); amin_upd(mn[j], a[j], (int32_t)r); fs[j] += a[j]; }
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) dsum[j] += (double)fs[j];
    }
    // combine the row lanes through shared memory: layout [RL][CT*VEC]
    const int W = g.CT * VEC;
    double* s_sum = reinterpret_cast<double*>(smem_raw);
    float* s_mx = reinterpret_cast<float*>(s_sum + (size_t)g.RL * W);
    float* s_mn = s_mx + (size_t)g.RL * W;
    int32_t* s_ix = reinterpret_cast<int32_t*>(s_mn + (size_t)g.RL * W);
    int32_t* s_in = s_ix + (size_t)g.RL * W;
    if (rl < g.RL) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const int o = rl * W + ct * VEC + j;
            s_sum[o] = dsum[j]; s_mx[o] = mx[j].v; s_mn[o] = mn[j].v; s_ix[o] = mx[j].i; s_in[o] = mn[j].i;
        }
    }
    __syncthreads();
    if (rl == 0 && c0 < g.C) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (c0 + j >= g.C) break;
            for (int q = 1; q < g.RL; ++q) {
                const int o = q * W + ct * VEC + j;
                amax_merge(mx[j], s_mx[o], s_ix[o]);
                amin_merge(mn[j], s_mn[o], s_in[o]);
                dsum[j] += s_sum[o];
            }
            const int64_t o = ((int64_t)blockIdx.x) * g.C + c0 + j;     // [(k*S+s)][c]
            pmax[o] = mx[j].v; pmin[o] = mn[j].v; imax[o] = mx[j].i; imin[o] = mn[j].i; psum[o] = dsum[j];
        }
    }
}

// merge the S segment partials of every (chunk k, channel c): one thread each, c fastest (coalesced)
__global__ void k_rbn_stats_nhwc_merge(const float* pmax, const float* pmin, const int32_t* imax, const int32_t* imin,
                                       const double* psum, int64_t C, int K, int S, int32_t* amax, int32_t* amin,
                                       float* cmax, float* cmin, double* csum) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= C * K) return;
    const int64_t k = t / C, c = t - k * C;
    ArgV mx{-QB_INF, 0x7fffffff}, mn{QB_INF, 0x7fffffff};
    double ssum = 0.0;
    for (int s = 0; s < S; ++s) {
        const int64_t o = (k * S + s) * C + c;
        amax_merge(mx, pmax[o], imax[o]);
        amin_merge(mn, pmin[o], imin[o]);
        ssum += psum[o];
    }
    cmax[t] = mx.v;
    cmin[t] = mn.v;
    csum[t] = ssum;
    if (amax) amax[c * K + k] = mx.i;
    if (amin) amin[c * K + k] = mn.i;
}

// red[3][C]: sum over the K chunks in a fixed order
__global__ void k_rbn_stats_nhwc_pack(const float* cmax, const float* cmin, const double* csum, int64_t C, int K,
                                      double* red) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    double smax = 0.0, smin = 0.0, ssum = 0.0;
    for (int k = 0; k < K; ++k) {
        smax += (double)cmax[(int64_t)k * C + c];
        smin += (double)cmin[(int64_t)k * C + c];
        ssum += csum[(int64_t)k * C + c];
    }
    red[c] = smax;
    red[C + c] = smin;
    red[2 * C + c] = ssum;
}

extern "C" size_t qb_rangebn_nhwc_ws_bytes(int64_t C, int num_chunks) {
    if (C < 0 || num_chunks < 0) return 0;
    return (size_t)C * num_chunks * (NH_MAX_SEG * 24 + 16);
}

static size_t nhwc_smem(const NhwcGeom& g, int vec, int per_elem) { return (size_t)g.RL * g.CT * vec * per_elem; }

extern "C" int qb_rangebn_stats_nhwc(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                                     int32_t* amin, double* red, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x && red && ws, "qb_rangebn_stats_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0 && num_chunks > 0, "qb_rangebn_stats_nhwc: bad shape");
    const int64_t R = B * HW;
    QB_CHECK_ARG(R % num_chunks == 0,
                 "qb_rangebn_stats_nhwc: B*H*W=%lld not divisible by num_chunks=%d (the reference's view() fails too)",
                 (long long)R, num_chunks);
    QB_CHECK_ARG(R < 2147483647LL, "qb_rangebn_stats_nhwc: per-channel element count exceeds int32 indices");
    if (ws_bytes < qb_rangebn_nhwc_ws_bytes(C, num_chunks)) {
        qb_set_error("qb_rangebn_stats_nhwc: workspace %zu < %zu bytes", ws_bytes, qb_rangebn_nhwc_ws_bytes(C, num_chunks));
        return QB_ERR_WORKSPACE;
    }
    const bool vec4 = (C % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    NhwcGeom g = vec4 ? make_geom<4>(R, C, num_chunks) : make_geom<1>(R, C, num_chunks);
    const size_t n = (size_t)num_chunks * g.S * C;
    double* psum = static_cast<double*>(ws);
    float* pmax = reinterpret_cast<float*>(psum + n);
    float* pmin = pmax + n;
    int32_t* imax = reinterpret_cast<int32_t*>(pmin + n);
    int32_t* imin = imax + n;
    const size_t nk = (size_t)num_chunks * C;
    double* csum = reinterpret_cast<double*>(static_cast<char*>(ws) + (size_t)C * num_chunks * NH_MAX_SEG * 24);
    float* cmax = reinterpret_cast<float*>(csum + nk);
    float* cmin = cmax + nk;
    dim3 grid((unsigned)(num_chunks * g.S), (unsigned)((C + g.tileC - 1) / g.tileC));
    cudaStream_t st = qb_stream(stream);
    if (vec4)
        k_rbn_stats_nhwc<4><<<grid, NH_THREADS, nhwc_smem(g, 4, 24), st>>>(x, g, pmax, pmin, imax, imin, psum);
    else
        k_rbn_stats_nhwc<1><<<grid, NH_THREADS, nhwc_smem(g, 1, 24), st>>>(x, g, pmax, pmin, imax, imin, psum);
    QB_LAUNCH_CHECK();
    k_rbn_stats_nhwc_merge<<<(unsigned)((nk + 127) / 128), 128, 0, st>>>(pmax, pmin, imax, imin, psum, C, num_chunks, g.S,
                                                                        amax, amin, cmax, cmin, csum);
    QB_LAUNCH_CHECK();
    k_rbn_stats_nhwc_pack<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(cmax, cmin, csum, C, num_chunks, red);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// apply: out[r][c] = ((x - mean_c) * qscale_c) * qgamma_c + qbeta_c      (quantize.py:488-499)
// backward apply: dx[r][c] = (gq * qgamma_c) * qscale_c + a_c            (+ scatter of e_c, below)
// One kernel template: MODE 0 = forward normalise, MODE 1 = backward.
// ---------------------------------------------------------------------------------------------------
struct NhwcVecs {
    const float *p0, *p1, *p2, *p3;   // fwd: mean, qscale, qgamma|null, qbeta|null ; bwd: qgamma|null, qscale, coef_a, unused
};

template <int MODE>
__device__ __forceinline__ float nhwc_one(float v, uint32_t c, const NhwcVecs& p) {
    if (MODE == 0) {
        float t = __fmul_rn(__fsub_rn(v, __ldg(p.p0 + c)), __ldg(p.p1 + c));
        if (p.p2) t = __fmul_rn(t, __ldg(p.p2 + c));
        if (p.p3) t = __fadd_rn(t, __ldg(p.p3 + c));
        return t;
    } else {
        float t = p.p0 ? __fmul_rn(v, __ldg(p.p0 + c)) : v;
        return __fadd_rn(__fmul_rn(t, __ldg(p.p1 + c)), __ldg(p.p2 + c));
    }
}

template <int MODE, bool C4>
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_apply_nhwc(const float* __restrict__ x, float* out, uint32_t n, FastDiv dC, NhwcVecs p) {
    const uint32_t nquads = n >> 2;
    const uint32_t stride = gridDim.x * NH_THREADS;
    for (uint32_t i = blockIdx.x * NH_THREADS + threadIdx.x; i < nquads; i += stride) {
        const float4 v = ld_stream4(x + 4ull * i);
        const uint32_t e = 4u * i;
        uint32_t c = e - fdiv(e, dC) * dC.d;
        float4 r;
        if (C4) {            // C % 4 == 0: the quad holds channels c..c+3 -> 128-bit parameter loads (L1 resident)
            const float4 a0 = p.p0 ? __ldg(reinterpret_cast<const float4*>(p.p0 + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
            const float4 a1 = __ldg(reinterpret_cast<const float4*>(p.p1 + c));
            const float4 a2 = p.p2 ? __ldg(reinterpret_cast<const float4*>(p.p2 + c)) : make_float4(1.f, 1.f, 1.f, 1.f);
            if (MODE == 0) {
                r.x = __fmul_rn(__fsub_rn(v.x, a0.x), a1.x); r.y = __fmul_rn(__fsub_rn(v.y, a0.y), a1.y);
                r.z = __fmul_rn(__fsub_rn(v.z, a0.z), a1.z); r.w = __fmul_rn(__fsub_rn(v.w, a0.w), a1.w);
                if (p.p2) { r.x = __fmul_rn(r.x, a2.x); r.y = __fmul_rn(r.y, a2.y); r.z = __fmul_rn(r.z, a2.z); r.w = __fmul_rn(r.w, a2.w); }
                if (p.p3) {
                    const float4 a3 = __ldg(reinterpret_cast<const float4*>(p.p3 + c));
                    r.x = __fadd_rn(r.x, a3.x); r.y = __fadd_rn(r.y, a3.y); r.z = __fadd_rn(r.z, a3.z); r.w = __fadd_rn(r.w, a3.w);
                }
            } else {
                r = v;
                if (p.p0) { r.x = __fmul_rn(r.x, a0.x); r.y = __fmul_rn(r.y, a0.y); r.z = __fmul_rn(r.z, a0.z); r.w = __fmul_rn(r.w, a0.w); }
                r.x = __fadd_rn(__fmul_rn(r.x, a1.x), a2.x); r.y = __fadd_rn(__fmul_rn(r.y, a1.y), a2.y);
                r.z = __fadd_rn(__fmul_rn(r.z, a1.z), a2.z); r.w = __fadd_rn(__fmul_rn(r.w, a1.w), a2.w);
            }
        } else {
            r.x = nhwc_one<MODE>(v.x, c, p); c = (c + 1 == dC.d) ? 0 : c + 1;
            r.y = nhwc_one<MODE>(v.y, c, p); c = (c + 1 == dC.d) ? 0 : c + 1;
            r.z = nhwc_one<MODE>(v.z, c, p); c = (c + 1 == dC.d) ? 0 : c + 1;
            r.w = nhwc_one<MODE>(v.w, c, p);
        }
        st_stream4(out + 4ull * i, r);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
        const uint32_t e = (nquads << 2) + threadIdx.x;
        out[e] = nhwc_one<MODE>(x[e], e - fdiv(e, dC) * dC.d, p);
    }
}

static int nh_grid(int64_t items) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (items + NH_THREADS - 1) / NH_THREADS, cap = (int64_t)sms * 8;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

template <int MODE>
static int launch_apply_nhwc(const char* who, const float* x, float* out, int64_t R, int64_t C, NhwcVecs p, cudaStream_t st) {
    const int64_t n = R * C;
    if (n >= 4294967296LL) {
        qb_set_error("%s: %lld elements >= 2^32 not supported", who, (long long)n);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0,
                 "%s: tensors must be 16-byte aligned", who);
    FastDiv dC = make_fastdiv((uint32_t)C);
    const int grid = nh_grid(n >> 2);
    if (C % 4 == 0)
        k_rbn_apply_nhwc<MODE, true><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    else
        k_rbn_apply_nhwc<MODE, false><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_rangebn_apply_nhwc(const float* x, float* out, int64_t B, int64_t C, int64_t HW, const float* mean,
                                     const float* qscale, const float* qgamma, const float* qbeta, void* stream) {
    QB_CHECK_ARG(x && out && mean && qscale, "qb_rangebn_apply_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_apply_nhwc: bad shape");
    return launch_apply_nhwc<0>("qb_rangebn_apply_nhwc", x, out, B * HW, C, NhwcVecs{mean, qscale, qgamma, qbeta},
                                qb_stream(stream));
}

// ---------------------------------------------------------------------------------------------------
// backward reduce: sg[c] = sum_r gq[r][c], sgx[c] = sum_r gq[r][c] * x[r][c]   (column sums, fp64)
// ---------------------------------------------------------------------------------------------------
template <int VEC>
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_bwd_reduce_nhwc(const float* __restrict__ gq, const float* __restrict__ x, NhwcGeom g, double* part /*[2][S][C]*/) {
    using V = typename VecT<VEC>::T;
    extern __shared__ unsigned char smem_raw[];
    const int ct = threadIdx.x % g.CT, rl = threadIdx.x / g.CT;
    const int64_t c0 = (int64_t)blockIdx.y * g.tileC + (int64_t)ct * VEC;
    const int s = blockIdx.x;
    const int64_t r_begin = (int64_t)s * g.seg_rows;
    const int64_t r_end = (r_begin + g.seg_rows < g.R) ? r_begin + g.seg_rows : g.R;
    const bool active = rl < g.RL && c0 < g.C;
    double dg[VEC], dgx[VEC];
    float fg[VEC], fgx[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) { dg[j] = 0.0; dgx[j] = 0.0; fg[j] = 0.f; fgx[j] = 0.f; }
    if (active) {
        int64_t r = r_begin + rl;
        const int64_t step = g.RL;
        int cnt = 0;
        for (; r + (NH_UNROLL - 1) * step < r_end; r += NH_UNROLL * step) {
            V a[NH_UNROLL], b[NH_UNROLL];
#pragma unroll
            for (int u = 0; u < NH_UNROLL; ++u) {
                a[u] = __ldcs(reinterpret_cast<const V*>(gq + (r + u * step) * g.C + c0));
                b[u] = __ldcs(reinterpret_cast<const V*>(x + (r + u * step) * g.C + c0));
            }
#pragma unroll
            for (int u = 0; u < NH_UNROLL; ++u) {
                float av[VEC], bv[VEC];
                unpack(a[u], av);
                unpack(b[u], bv);
#pragma unroll
                for (int j = 0; j < VEC; ++j) { fg[j] += av[j]; fgx[j] += av[j] * bv[j]; }
            }
            if (++cnt == 8) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) { dg[j] += (double)fg[j]; dgx[j] += (double)fgx[j]; fg[j] = 0.f; fgx[j] = 0.f; }
                cnt = 0;
            }
        }
        for (; r < r_end; r += step) {
            float av[VEC], bv[VEC];
            unpack(__ldcs(reinterpret_cast<const V*>(gq + r * g.C + c0)), av);
            unpack(__ldcs(reinterpret_cast<const V*>(x + r * g.C + c0)), bv);
#pragma unroll
            for (int j = 0; j < VEC; ++j) { fg[j] += av[j]; fgx[j] += av[j] * bv[j]; }
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) { dg[j] += (double)fg[j]; dgx[j] += (double)fgx[j]; }
    }
    const int W = g.CT * VEC;
    double* s_g = reinterpret_cast<double*>(smem_raw);
    double* s_gx = s_g + (size_t)g.RL * W;
    if (rl < g.RL) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) { s_g[rl * W + ct * VEC + j] = dg[j]; s_gx[rl * W + ct * VEC + j] = dgx[j]; }
    }
    __syncthreads();
    if (rl == 0 && c0 < g.C) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            if (c0 + j >= g.C) break;
            for (int q = 1; q < g.RL; ++q) { dg[j] += s_g[q * W + ct * VEC + j]; dgx[j] += s_gx[q * W + ct * VEC + j]; }
            part[(int64_t)s * g.C + c0 + j] = dg[j];
            part[((int64_t)g.S + s) * g.C + c0 + j] = dgx[j];
        }
    }
}

// one warp per channel: lanes stride over the segments, fixed-order shuffle tree => deterministic
__global__ void k_rbn_bwd_pack_nhwc(const double* part, int64_t C, int S, double* sg, double* sgx) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (c >= C) return;
    double a = 0.0, b = 0.0;
    for (int s = lane; s < S; s += 32) { a += part[(int64_t)s * C + c]; b += part[((int64_t)S + s) * C + c]; }
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0) { sg[c] = a; sgx[c] = b; }
}

constexpr int NH_MAX_SEG_BWD = 1024;
extern "C" size_t qb_rangebn_bwd_nhwc_ws_bytes(int64_t C) { return C < 0 ? 0 : (size_t)C * NH_MAX_SEG_BWD * 16; }

extern "C" int qb_rangebn_bwd_reduce_nhwc(const float* gq, const float* x, int64_t B, int64_t C, int64_t HW, double* sg,
                                          double* sgx, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(gq && x && sg && sgx && ws, "qb_rangebn_bwd_reduce_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_reduce_nhwc: bad shape");
    if (ws_bytes < qb_rangebn_bwd_nhwc_ws_bytes(C)) {
        qb_set_error("qb_rangebn_bwd_reduce_nhwc: workspace %zu < %zu bytes", ws_bytes, qb_rangebn_bwd_nhwc_ws_bytes(C));
        return QB_ERR_WORKSPACE;
    }
    const int64_t R = B * HW;
    const bool vec4 = (C % 4 == 0) && (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gq)) & 15) == 0);
    NhwcGeom g = vec4 ? make_geom<4>(R, C, 1, NH_MAX_SEG_BWD) : make_geom<1>(R, C, 1, NH_MAX_SEG_BWD);
    double* part = static_cast<double*>(ws);
    dim3 grid((unsigned)g.S, (unsigned)((C + g.tileC - 1) / g.tileC));
    cudaStream_t st = qb_stream(stream);
    if (vec4)
        k_rbn_bwd_reduce_nhwc<4><<<grid, NH_THREADS, nhwc_smem(g, 4, 16), st>>>(gq, x, g, part);
    else
        k_rbn_bwd_reduce_nhwc<1><<<grid, NH_THREADS, nhwc_smem(g, 1, 16), st>>>(gq, x, g, part);
    QB_LAUNCH_CHECK();
    k_rbn_bwd_pack_nhwc<<<(unsigned)((C * 32 + 127) / 128), 128, 0, st>>>(part, C, g.S, sg, sgx);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// scatter: flat (b,h,w) index == row index in channels-last
__global__ void k_rbn_bwd_scatter_nhwc(float* dx, int64_t C, int K, const float* coef_e, const int32_t* amax,
                                       const int32_t* amin) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= C * K) return;
    const int64_t c = t / K;
    const float e = coef_e[c];
    float* q = dx + (int64_t)amax[t] * C + c;
    *q = __fadd_rn(*q, e);
    q = dx + (int64_t)amin[t] * C + c;
    *q = __fsub_rn(*q, e);
}

extern "C" int qb_rangebn_bwd_apply_nhwc(const float* gq, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                                         const float* qscale, const float* qgamma, const float* coef_a,
                                         const float* coef_e, const int32_t* amax, const int32_t* amin, void* stream) {
    QB_CHECK_ARG(gq && dx && qscale && coef_a, "qb_rangebn_bwd_apply_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_apply_nhwc: bad shape");
    cudaStream_t st = qb_stream(stream);
    int rc = launch_apply_nhwc<1>("qb_rangebn_bwd_apply_nhwc", gq, dx, B * HW, C, NhwcVecs{qgamma, qscale, coef_a, nullptr}, st);
    if (rc) return rc;
    if (coef_e && amax && amin && num_chunks > 0) {
        const int64_t t = C * num_chunks;
        k_rbn_bwd_scatter_nhwc<<<(unsigned)((t + 127) / 128), 128, 0, st>>>(dx, C, num_chunks, coef_e, amax, amin);
        QB_LAUNCH_CHECK();
    }
    return QB_OK;
}
