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
#include "rangebn_common.cuh"

constexpr int NH_THREADS = 256;
constexpr int NH_UNROLL = 4;
constexpr int NH_MAX_SEG = 32;      // row segments per chunk (partials per (chunk, channel))
constexpr int NH_MAX_SEG_BWD = 1024;

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

// CTAs that are resident at once (SMs x occupancy of `kernel`): grids are sized to ONE full wave of them - 512 CTAs on 444
// slots ran as 1.15 waves, i.e. a second wave of 68 CTAs on an otherwise idle GPU (profiles/ncu_rbn_r02d.txt)
template <typename Kern>
static int nh_slots(Kern kernel, size_t smem) {
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, NH_THREADS, smem) != cudaSuccess || occ < 1) occ = 2;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    return occ * sms;
}

template <int VEC>
static NhwcGeom make_geom(int64_t R, int64_t C, int K, int max_seg = NH_MAX_SEG, int slots = 0) {
    NhwcGeom g;
    g.R = R; g.C = C; g.K = K; g.L = R / K;
    g.tileC = (int)(C < (int64_t)NH_THREADS * VEC ? C : (int64_t)NH_THREADS * VEC);
    g.CT = (g.tileC + VEC - 1) / VEC;
    g.RL = NH_THREADS / g.CT;
    int ctiles = (int)((C + g.tileC - 1) / g.tileC);
    // enough CTAs to fill 148 SMs a few times over, each segment at least RL*NH_UNROLL rows
    int64_t want = slots > 0 ? slots / ((int64_t)K * ctiles) : (148 * 4 + (int64_t)K * ctiles - 1) / ((int64_t)K * ctiles);
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
template <> struct VecT<8> { using T = F8; };
__device__ __forceinline__ void unpack(const F8& v, float (&a)[8]) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = v.v[j];
}
template <typename V> __device__ __forceinline__ V ld_vec_stream(const float* p) { return __ldcs(reinterpret_cast<const V*>(p)); }
template <> __device__ __forceinline__ F8 ld_vec_stream<F8>(const float* p) { return ld256_stream(p); }
__device__ __forceinline__ void unpack(const float4& v, float (&a)[4]) { a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w; }
__device__ __forceinline__ void unpack(const float& v, float (&a)[1]) { a[0] = v; }

template <int VEC>
__device__ __forceinline__ void store_codes(uint8_t* dst, const uint32_t (&cd)[VEC]) {
    if (VEC == 1) {
        *dst = (uint8_t)cd[0];
    } else {
#pragma unroll
        for (int j = 0; j < VEC; j += 4)
            *reinterpret_cast<uint32_t*>(dst + j) = cd[j] | (cd[j + 1 < VEC ? j + 1 : 0] << 8) | (cd[j + 2 < VEC ? j + 2 : 0] << 16) |
                                                    (cd[j + 3 < VEC ? j + 3 : 0] << 24);
    }
}

// ---------------------------------------------------------------------------------------------------
// forward statistics: partials per (chunk k, segment s, channel c)
// ---------------------------------------------------------------------------------------------------
template <int VEC, bool QUANT>
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_stats_nhwc(const float* __restrict__ x, NhwcGeom g, float* pmax, float* pmin, int32_t* imax, int32_t* imin,
                 double* psum, const float* q_range, int q_bits, uint8_t* __restrict__ codes_out) {
    using V = typename VecT<VEC>::T;
    AffineQ aq;
    if (QUANT) aq = load_affineq(q_range, q_bits);
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
                if (QUANT) {
                    if (codes_out) {          // x_hat leaves as uint8 codes: the later passes read 1 B/elem instead of 4
                        uint32_t cd[VEC];
                        affine_fq_vec<VEC>(a, aq, cd);
                        store_codes<VEC>(codes_out + (r + u * step) * g.C + c0, cd);
                    } else {
                        affine_fq_vec<VEC>(a, aq);
                    }
                }
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
            if (QUANT) {
                if (codes_out) {
                    uint32_t cd[VEC];
                    affine_fq_vec<VEC>(a, aq, cd);
                    store_codes<VEC>(codes_out + r * g.C + c0, cd);
                } else {
                    affine_fq_vec<VEC>(a, aq);
                }
            }
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
                                       int32_t* amax_kc, int32_t* amin_kc, float* cmax, float* cmin, double* csum) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= C * K) return;
    const int64_t k = t / C, c = t - k * C;
    ArgV mx{-QB_INF, 0x7fffffff}, mn{QB_INF, 0x7fffffff};
    double ssum = 0.0;
#pragma unroll 8
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
    if (amax_kc) amax_kc[t] = mx.i;          // [k][c]: what the fused backward reads with one vector load per octet
    if (amin_kc) amin_kc[t] = mn.i;
}

// ---------------------------------------------------------------------------------------------------
// forward statistics in CODE SPACE (input quantizer fused, q_bits <= 8, C % 8 == 0, rows per chunk <= 2^24).
// x_hat = deq(code) is a monotone function of the integer code, so max / min / first-arg of x_hat over a chunk are
// max / min / first-arg of the code, and sum(x_hat) follows from the integer sum of the codes.  Per element the kernel
// only divides, clamps and adds the rounding constant (packed fp32x2); the trackers are integer keys
//     key_max = code << 24 | (2^24 - 1 - row_in_chunk)      largest code, smallest row on ties
//     key_min = code << 24 | row_in_chunk                   smallest code, smallest row on ties
// built with one PRMT from the raw bits of (t + 1.5*2^23), reduced with VIMNMX (two rows per instruction), and the
// sums are plain integer adds of those bits - exact and order independent, so no fp64 and no index bookkeeping.
// Optionally writes the codes (uint8) for the later passes.
// ---------------------------------------------------------------------------------------------------

struct Q8Row {
    uint32_t bits[8];
};

// One accumulator per (chunk k, channel c), combined across the segment CTAs with integer atomics (max / max of the
// complement / add: exact and order independent, hence deterministic).  All-zero is the identity; k_rbn_stats_q8_finish
// resets what it has read, so the workspace region stays clean between calls.
struct Q8Acc {
    uint32_t kmax;               // largest key_max
    uint32_t nkmin;              // largest ~key_min
    unsigned long long sum;      // SUM of the codes
};
constexpr size_t NH_ACC_BYTES = (size_t)1 << 20;                     // fixed region at the head of the workspace: 65 536 (k, c) pairs

__device__ __forceinline__ Q8Row q8_row(const F8& v, const AffineQ& aq, uint64_t nlo2, uint64_t r2, uint64_t nb2, uint64_t magic2) {
    Q8Row o;
    float q[8];
    bool ok = aq.d.fast;
    uint64_t t2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        t2[j] = add2(pack2(v.v[2 * j], v.v[2 * j + 1]), nlo2);           // x - lo
        float a, b;
        unpack2(t2[j], a, b);
        ok = ok && qb_div_ok(a) && qb_div_ok(b);
    }
    if (ok) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint64_t qq = mul2(t2[j], r2);                         // qb_div_fast, two lanes at a time
            const uint64_t e = fma2(qq, nb2, t2[j]);
            unpack2(fma2(r2, e, qq), q[2 * j], q[2 * j + 1]);
        }
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float a, b;
            unpack2(t2[j], a, b);
            q[2 * j] = __fdiv_rn(a, aq.d.b);
            q[2 * j + 1] = __fdiv_rn(b, aq.d.b);
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float a = fminf(fmaxf(q[2 * j], 0.0f), aq.p.qmax), b = fminf(fmaxf(q[2 * j + 1], 0.0f), aq.p.qmax);
        float ma, mb;
        unpack2(add2(pack2(a, b), magic2), ma, mb);                      // RNE to the integer code, in the low mantissa bits
        o.bits[2 * j] = __float_as_uint(ma);
        o.bits[2 * j + 1] = __float_as_uint(mb);
    }
    return o;
}

__global__ void __launch_bounds__(NH_THREADS, 3)
k_rbn_stats_q8_nhwc(const float* __restrict__ x, NhwcGeom g, uint32_t* __restrict__ pkmax, uint32_t* __restrict__ pkmin,
                    uint32_t* __restrict__ pksum, const float* q_range, int q_bits, uint8_t* __restrict__ codes_out,
                    Q8Acc* __restrict__ acc_kc) {
    extern __shared__ unsigned char smem_raw[];
    const AffineQ aq = load_affineq(q_range, q_bits);
    const uint64_t nlo2 = pack2(-aq.p.lo, -aq.p.lo), r2 = pack2(aq.d.r, aq.d.r), nb2 = pack2(aq.d.nb, aq.d.nb),
                   magic2 = pack2(QB_MAGIC, QB_MAGIC);
    const int ct = threadIdx.x % g.CT, rl = threadIdx.x / g.CT;
    const int64_t c0 = (int64_t)blockIdx.y * g.tileC + (int64_t)ct * 8;
    const int k = blockIdx.x / g.S, sgm = blockIdx.x % g.S;
    const int64_t chunk0 = (int64_t)k * g.L;
    const int64_t r_begin = chunk0 + (int64_t)sgm * g.seg_rows;
    int64_t r_end = r_begin + g.seg_rows;
    if (r_end > chunk0 + g.L) r_end = chunk0 + g.L;
    uint32_t kmx[8], kmn[8], acc[8], nrows = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) { kmx[j] = 0u; kmn[j] = 0xffffffffu; acc[j] = 0u; }
    if (rl < g.RL && c0 < g.C) {
        auto store = [&](const Q8Row& q, int64_t r) {
            uint2 pk;
            pk.x = __byte_perm(__byte_perm(q.bits[0], q.bits[1], 0x0040), __byte_perm(q.bits[2], q.bits[3], 0x0040), 0x5410);
            pk.y = __byte_perm(__byte_perm(q.bits[4], q.bits[5], 0x0040), __byte_perm(q.bits[6], q.bits[7], 0x0040), 0x5410);
            *reinterpret_cast<uint2*>(codes_out + r * g.C + c0) = pk;
        };
        const int64_t step = g.RL;
        int64_t r = r_begin + rl;
        for (; r + step < r_end; r += 2 * step) {                           // two rows per iteration: 3-input min / max
            const F8 v0 = ld256_stream(x + r * g.C + c0), v1 = ld256_stream(x + (r + step) * g.C + c0);
            const Q8Row q0 = q8_row(v0, aq, nlo2, r2, nb2, magic2), q1 = q8_row(v1, aq, nlo2, r2, nb2, magic2);
            const uint32_t lo0 = (uint32_t)(r - chunk0), lo1 = lo0 + (uint32_t)step;
            const uint32_t hi0 = 0xffffffu - lo0, hi1 = 0xffffffu - lo1;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                kmx[j] = max(max(kmx[j], __byte_perm(hi0, q0.bits[j], 0x4210)), __byte_perm(hi1, q1.bits[j], 0x4210));
                kmn[j] = min(min(kmn[j], __byte_perm(lo0, q0.bits[j], 0x4210)), __byte_perm(lo1, q1.bits[j], 0x4210));
                acc[j] += q0.bits[j] + q1.bits[j];
            }
            nrows += 2;
            if (codes_out) { store(q0, r); store(q1, r + step); }
        }
        for (; r < r_end; r += step) {
            const Q8Row q0 = q8_row(ld256_stream(x + r * g.C + c0), aq, nlo2, r2, nb2, magic2);
            const uint32_t lo0 = (uint32_t)(r - chunk0), hi0 = 0xffffffu - lo0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                kmx[j] = max(kmx[j], __byte_perm(hi0, q0.bits[j], 0x4210));
                kmn[j] = min(kmn[j], __byte_perm(lo0, q0.bits[j], 0x4210));
                acc[j] += q0.bits[j];
            }
            nrows += 1;
            if (codes_out) store(q0, r);
        }
    }
    // combine the row lanes: [RL][CT*8] of (key_max, key_min, code sum); column `col` is finished by thread `col`
    const int W = g.CT * 8;
    uint32_t* s_mx = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_mn = s_mx + (size_t)g.RL * W;
    uint32_t* s_sm = s_mn + (size_t)g.RL * W;
    if (rl < g.RL) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int o = rl * W + ct * 8 + j;
            s_mx[o] = kmx[j]; s_mn[o] = kmn[j];
            s_sm[o] = acc[j] - nrows * QB_MAGIC_BITS;               // = sum of the codes (mod 2^32; < 2^32 by the row bound)
        }
    }
    __syncthreads();
    for (int col = threadIdx.x; col < W; col += NH_THREADS) {
        const int64_t c = (int64_t)blockIdx.y * g.tileC + col;
        if (c >= g.C) break;
        uint32_t a = 0u, b = 0xffffffffu, sm = 0u;
        for (int q = 0; q < g.RL; ++q) { a = max(a, s_mx[q * W + col]); b = min(b, s_mn[q * W + col]); sm += s_sm[q * W + col]; }
        if (acc_kc) {
            Q8Acc* dst = acc_kc + (int64_t)k * g.C + c;
            atomicMax(&dst->kmax, a);
            atomicMax(&dst->nkmin, ~b);
            atomicAdd(&dst->sum, (unsigned long long)sm);
            continue;
        }
        const int64_t o = (int64_t)blockIdx.x * g.C + c;            // [(k*S+s)][c]
        pkmax[o] = a; pkmin[o] = b; pksum[o] = sm;
    }
}

// Atomic-combine path: per channel, walk the K chunk accumulators in order - arg rows, chunk extrema and the chunk sums the
// merge kernel produced, plus the red[3][C] block of k_rbn_stats_nhwc_pack - and reset them.  One thread per channel
// (coalesced over c); K is small (16).
struct Q8FinishArg {
    Q8Acc* acc_kc;
    int64_t C, L;
    int K;
    const float* q_range;
    int q_bits;
    int32_t *amax, *amin, *amax_kc, *amin_kc;
    double* red;
};

// channel c: the K accumulators are fetched KB at a time before anything is stored (the resets alias the loads, so a plain
// loop serialises K round trips to L2), then folded in chunk order - the summation order of k_rbn_stats_nhwc_pack
template <int KB>
__device__ __forceinline__ void q8_finish_channel(const Q8FinishArg& f, const QParams& p, int64_t c) {
    double smax = 0.0, smin = 0.0, ssum = 0.0;
    for (int k0 = 0; k0 < f.K; k0 += KB) {
        uint32_t a[KB], b[KB];
        unsigned long long sm[KB];
#pragma unroll
        for (int j = 0; j < KB; ++j) {
            if (k0 + j < f.K) {
                const Q8Acc* q = f.acc_kc + (int64_t)(k0 + j) * f.C + c;
                a[j] = q->kmax; b[j] = ~q->nkmin; sm[j] = q->sum;
            }
        }
#pragma unroll
        for (int j = 0; j < KB; ++j) {
            if (k0 + j < f.K) {
                const int k = k0 + j;
                const int64_t t = (int64_t)k * f.C + c;
                f.acc_kc[t].kmax = 0u; f.acc_kc[t].nkmin = 0u; f.acc_kc[t].sum = 0ull;
                const int32_t imx = (int32_t)(k * f.L + (int64_t)(0xffffffu - (a[j] & 0xffffffu)));
                const int32_t imn = (int32_t)(k * f.L + (int64_t)(b[j] & 0xffffffu));
                smax += (double)q_affine_deq((float)(a[j] >> 24), p);
                smin += (double)q_affine_deq((float)(b[j] >> 24), p);
                ssum += (double)f.L * (double)p.lo + (double)p.scale * (double)sm[j];          // as k_rbn_stats_q8_merge
                if (f.amax) f.amax[c * f.K + k] = imx;
                if (f.amin) f.amin[c * f.K + k] = imn;
                if (f.amax_kc) f.amax_kc[t] = imx;
                if (f.amin_kc) f.amin_kc[t] = imn;
            }
        }
    }
    f.red[c] = smax;
    f.red[f.C + c] = smin;
    f.red[2 * f.C + c] = ssum;
}

__global__ void k_rbn_stats_q8_finish(Q8FinishArg f) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= f.C) return;
    const QParams p = make_qparams_d((double)__ldg(f.q_range), (double)__ldg(f.q_range + 1), f.q_bits);
    q8_finish_channel<16>(f, p, c);
}

// ... and with the per-channel finalize of quantize.py:470-498 behind it in the same (single) CTA: one launch instead of two
// per RangeBN layer when no cross-rank exchange sits between the two (qb_rangebn_stats_finalize_nhwc)
__global__ void __launch_bounds__(RF_THREADS) k_rbn_stats_q8_finish_finalize(Q8FinishArg f, RbnFinalizeArg a) {
    __shared__ float sm[6 * 32];
    if ((int64_t)threadIdx.x < f.C) {
        const QParams p = make_qparams_d((double)__ldg(f.q_range), (double)__ldg(f.q_range + 1), f.q_bits);
        for (int64_t c = threadIdx.x; c < f.C; c += RF_THREADS) q8_finish_channel<16>(f, p, c);
    }
    rbn_finalize_body(f.red, f.C, a, sm);            // thread t reads back the red[] entries thread t wrote
}

__global__ void __launch_bounds__(RF_THREADS) k_rbn_finalize_after(const double* red, int64_t C, RbnFinalizeArg a) {
    __shared__ float sm[6 * 32];
    rbn_finalize_body(red, C, a, sm);
}

// merge the S segment partials of every (chunk k, channel c) and leave them in the generic kernel's output format
__global__ void k_rbn_stats_q8_merge(const uint32_t* pkmax, const uint32_t* pkmin, const uint32_t* pksum, int64_t C, int K, int S,
                                     int64_t L, const float* q_range, int q_bits, int32_t* amax, int32_t* amin, int32_t* amax_kc,
                                     int32_t* amin_kc, float* cmax, float* cmin, double* csum) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= C * K) return;
    const QParams p = make_qparams_d((double)__ldg(q_range), (double)__ldg(q_range + 1), q_bits);
    const int64_t k = t / C, c = t - k * C;
    uint32_t a = 0u, b = 0xffffffffu;
    uint64_t sm = 0;
#pragma unroll 8
    for (int s = 0; s < S; ++s) {
        const int64_t o = (k * S + s) * C + c;
        a = max(a, pkmax[o]); b = min(b, pkmin[o]); sm += pksum[o];
    }
    const int32_t imx = (int32_t)(k * L + (int64_t)(0xffffffu - (a & 0xffffffu))), imn = (int32_t)(k * L + (int64_t)(b & 0xffffffu));
    cmax[t] = q_affine_deq((float)(a >> 24), p);
    cmin[t] = q_affine_deq((float)(b >> 24), p);
    // sum over the chunk of x_hat = code*scale + lo, from the exact integer code sum (each x_hat carries an fp32 rounding
    // of at most half an ulp that averages out over the L rows; the chunk mean stays within an ulp of the fp64 sum of
    // the rounded values, the same agreement the generic kernel has with ATen's fp32 sums)
    csum[t] = (double)L * (double)p.lo + (double)p.scale * (double)sm;
    if (amax) amax[c * K + k] = imx;
    if (amin) amin[c * K + k] = imn;
    if (amax_kc) amax_kc[t] = imx;
    if (amin_kc) amin_kc[t] = imn;
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
    return NH_ACC_BYTES + (size_t)C * num_chunks * (NH_MAX_SEG * 24 + 16);
}

static size_t nhwc_smem(const NhwcGeom& g, int vec, int per_elem) { return (size_t)g.RL * g.CT * vec * per_elem; }

static int stats_nhwc_impl(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax, int32_t* amin,
                           int32_t* amax_kc, int32_t* amin_kc, double* red, const float* q_range, int q_bits, uint8_t* x_codes_out,
                           void* ws, size_t ws_bytes, void* stream, const RbnFinalizeArg* fin) {
    QB_CHECK_ARG(!x_codes_out || (q_range && q_bits <= 8), "qb_rangebn_stats_nhwc: x_codes_out needs q_range and q_bits <= 8");
    QB_CHECK_ARG(!q_range || (q_bits >= 1 && q_bits <= 24), "qb_rangebn_stats_nhwc: q_bits=%d", q_bits);
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
    Q8Acc* acc_kc = static_cast<Q8Acc*>(ws);                          // fixed-size head region, zero between calls
    ws = static_cast<char*>(ws) + NH_ACC_BYTES;
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
    cudaStream_t st = qb_stream(stream);
    const bool q8 = q_range && q_bits <= 8 && C % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 31) == 0 && R / num_chunks <= (1 << 24) &&
                    (!x_codes_out || (reinterpret_cast<uintptr_t>(x_codes_out) & 7) == 0);
    if (q8) {
        static const int slots8 = nh_slots(k_rbn_stats_q8_nhwc, (size_t)NH_THREADS * 8 * 12);
        NhwcGeom g8 = make_geom<8>(R, C, num_chunks, NH_MAX_SEG, slots8);
        if ((size_t)num_chunks * g8.S * C * 12 <= (size_t)C * num_chunks * NH_MAX_SEG * 24) {
            uint32_t* pkmax = static_cast<uint32_t*>(ws);
            uint32_t* pkmin = pkmax + (size_t)num_chunks * g8.S * C;
            uint32_t* pksum = pkmin + (size_t)num_chunks * g8.S * C;
            dim3 grid8((unsigned)(num_chunks * g8.S), (unsigned)((C + g8.tileC - 1) / g8.tileC));
            if (nk * sizeof(Q8Acc) <= NH_ACC_BYTES) {                // two launches: the segment CTAs combine with integer atomics
                k_rbn_stats_q8_nhwc<<<grid8, NH_THREADS, (size_t)g8.RL * g8.CT * 8 * 12, st>>>(x, g8, nullptr, nullptr, nullptr, q_range,
                                                                                          q_bits, x_codes_out, acc_kc);
                QB_LAUNCH_CHECK();
                const Q8FinishArg f{acc_kc, C, R / num_chunks, num_chunks, q_range, q_bits, amax, amin, amax_kc, amin_kc, red};
                if (fin) k_rbn_stats_q8_finish_finalize<<<1, RF_THREADS, 0, st>>>(f, *fin);
                else k_rbn_stats_q8_finish<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(f);
                QB_LAUNCH_CHECK();
                return QB_OK;
            }
            k_rbn_stats_q8_nhwc<<<grid8, NH_THREADS, (size_t)g8.RL * g8.CT * 8 * 12, st>>>(x, g8, pkmax, pkmin, pksum, q_range, q_bits,
                                                                                      x_codes_out, nullptr);
            QB_LAUNCH_CHECK();
            k_rbn_stats_q8_merge<<<(unsigned)((nk + 127) / 128), 128, 0, st>>>(pkmax, pkmin, pksum, C, num_chunks, g8.S, R / num_chunks,
                                                                              q_range, q_bits, amax, amin, amax_kc, amin_kc, cmax,
                                                                              cmin, csum);
            QB_LAUNCH_CHECK();
            k_rbn_stats_nhwc_pack<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(cmax, cmin, csum, C, num_chunks, red);
            QB_LAUNCH_CHECK();
            if (fin) {
                k_rbn_finalize_after<<<1, RF_THREADS, 0, st>>>(red, C, *fin);
                QB_LAUNCH_CHECK();
            }
            return QB_OK;
        }
    }
    dim3 grid((unsigned)(num_chunks * g.S), (unsigned)((C + g.tileC - 1) / g.tileC));
    if (vec4 && q_range)
        k_rbn_stats_nhwc<4, true><<<grid, NH_THREADS, nhwc_smem(g, 4, 24), st>>>(x, g, pmax, pmin, imax, imin, psum, q_range, q_bits, x_codes_out);
    else if (vec4)
        k_rbn_stats_nhwc<4, false><<<grid, NH_THREADS, nhwc_smem(g, 4, 24), st>>>(x, g, pmax, pmin, imax, imin, psum, nullptr, 0, nullptr);
    else if (q_range)
        k_rbn_stats_nhwc<1, true><<<grid, NH_THREADS, nhwc_smem(g, 1, 24), st>>>(x, g, pmax, pmin, imax, imin, psum, q_range, q_bits, x_codes_out);
    else
        k_rbn_stats_nhwc<1, false><<<grid, NH_THREADS, nhwc_smem(g, 1, 24), st>>>(x, g, pmax, pmin, imax, imin, psum, nullptr, 0, nullptr);
    QB_LAUNCH_CHECK();
    k_rbn_stats_nhwc_merge<<<(unsigned)((nk + 127) / 128), 128, 0, st>>>(pmax, pmin, imax, imin, psum, C, num_chunks, g.S,
                                                                        amax, amin, amax_kc, amin_kc, cmax, cmin, csum);
    QB_LAUNCH_CHECK();
    k_rbn_stats_nhwc_pack<<<(unsigned)((C + 127) / 128), 128, 0, st>>>(cmax, cmin, csum, C, num_chunks, red);
    QB_LAUNCH_CHECK();
    if (fin) {
        k_rbn_finalize_after<<<1, RF_THREADS, 0, st>>>(red, C, *fin);
        QB_LAUNCH_CHECK();
    }
    return QB_OK;
}

extern "C" int qb_rangebn_stats_nhwc(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                                     int32_t* amin, int32_t* amax_kc, int32_t* amin_kc, double* red, const float* q_range,
                                     int q_bits, uint8_t* x_codes_out, void* ws, size_t ws_bytes, void* stream) {
    return stats_nhwc_impl(x, B, C, HW, num_chunks, amax, amin, amax_kc, amin_kc, red, q_range, q_bits, x_codes_out, ws, ws_bytes,
                           stream, nullptr);
}

/* qb_rangebn_stats_nhwc followed by qb_rangebn_finalize(training = 1) on the same stream, with the finalize folded into the
 * tail kernel of the statistics pass where that is a single CTA.  Same results as the two calls (same arithmetic, same
 * order).  For the single-process case: across ranks `red` has to be exchanged between the two, so the calls stay apart. */
extern "C" int qb_rangebn_stats_finalize_nhwc(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                                              int32_t* amin, int32_t* amax_kc, int32_t* amin_kc, double* red,
                                              const float* q_range, int q_bits, uint8_t* x_codes_out, float eps, double momentum,
                                              int bits, const float* gamma, const float* beta, float* running_mean,
                                              float* running_var, float* mean, float* scale_raw, float* qscale, float* qgamma,
                                              float* qbeta, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(running_mean && running_var && mean && scale_raw && qscale, "qb_rangebn_stats_finalize_nhwc: null pointer");
    QB_CHECK_ARG((!gamma || qgamma) && (!beta || qbeta), "qb_rangebn_stats_finalize_nhwc: gamma/beta without output");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_rangebn_stats_finalize_nhwc: bits=%d", bits);
    QB_CHECK_ARG(B > 0 && HW > 0 && num_chunks > 0 && (B * HW) / num_chunks > 1,
                 "qb_rangebn_stats_finalize_nhwc: chunk length must exceed 1");
    const double count = (double)B * (double)HW, chunk_len = count / num_chunks;
    const RbnFinalizeArg a{(double)num_chunks, count, rbn_scale_fix(chunk_len), momentum, eps, 1, bits, gamma, beta,
                           running_mean, running_var, mean, scale_raw, qscale, qgamma, qbeta};
    return stats_nhwc_impl(x, B, C, HW, num_chunks, amax, amin, amax_kc, amin_kc, red, q_range, q_bits, x_codes_out, ws, ws_bytes,
                           stream, &a);
}

// ---------------------------------------------------------------------------------------------------
// apply: out[r][c] = ((x - mean_c) * qscale_c) * qgamma_c + qbeta_c      (quantize.py:488-499)
// backward apply: dx[r][c] = (gq * qgamma_c) * qscale_c + a_c            (+ scatter of e_c, below)
// One kernel template: MODE 0 = forward normalise, MODE 1 = backward.
// ---------------------------------------------------------------------------------------------------
struct NhwcVecs {
    const float *p0, *p1, *p2, *p3;   // fwd: mean, qscale, qgamma|null, qbeta|null ; bwd: qgamma|null, qscale, coef_a, unused
    const float* q_range;             // fwd only: quantize x on the fly with this device range (or null)
    int q_bits;
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

template <int MODE, bool C4, bool QUANT>
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_apply_nhwc(const float* __restrict__ x, float* out, uint32_t n, FastDiv dC, NhwcVecs p) {
    const uint32_t nquads = n >> 2;
    const uint32_t stride = gridDim.x * NH_THREADS;
    AffineQ aq;
    if (QUANT) aq = load_affineq(p.q_range, p.q_bits);
    for (uint32_t i = blockIdx.x * NH_THREADS + threadIdx.x; i < nquads; i += stride) {
        float4 v = ld_stream4(x + 4ull * i);
        if (QUANT) {
            float a[4] = {v.x, v.y, v.z, v.w};
            affine_fq_vec<4>(a, aq);
            v = make_float4(a[0], a[1], a[2], a[3]);
        }
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
        out[e] = nhwc_one<MODE>(QUANT ? affine_fq(x[e], aq) : x[e], e - fdiv(e, dC) * dC.d, p);
    }
}

// Forward normalisation from the uint8 codes the statistics pass wrote (C % 8 == 0): 1 + 4 B/elem.  Column-persistent:
// a thread keeps mean / qscale / qgamma / qbeta of its 8 channels in registers and strides over rows.
__global__ void __launch_bounds__(NH_THREADS)
k_rbn_apply_codes_nhwc(const uint8_t* __restrict__ codes, float* __restrict__ out, NhwcGeom geo, NhwcVecs p) {
    const AffineQ aq = load_affineq(p.q_range, p.q_bits);
    const int ct = threadIdx.x % geo.CT, rl = threadIdx.x / geo.CT;
    const int64_t c0 = (int64_t)blockIdx.y * geo.tileC + (int64_t)ct * 8;
    const int64_t r_begin = (int64_t)blockIdx.x * geo.seg_rows;
    const int64_t r_end = (r_begin + geo.seg_rows < geo.R) ? r_begin + geo.seg_rows : geo.R;
    if (rl >= geo.RL || c0 >= geo.C) return;
    float mu[8], sc[8], gm[8], be[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        mu[j] = __ldg(p.p0 + c0 + j);
        sc[j] = __ldg(p.p1 + c0 + j);
        gm[j] = p.p2 ? __ldg(p.p2 + c0 + j) : 1.0f;
        be[j] = p.p3 ? __ldg(p.p3 + c0 + j) : 0.0f;
    }
    auto one_row = [&](const uint2& pk, int64_t r) {
        F8 o;
        float xh[8];
        affine_deq8(pk, aq.p, xh);                       // PRMT + FADD per code: no I2F (the XU pipe was this pass' bound too)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float t = __fmul_rn(__fsub_rn(xh[j], mu[j]), sc[j]);
            if (p.p2) t = __fmul_rn(t, gm[j]);
            if (p.p3) t = __fadd_rn(t, be[j]);
            o.v[j] = t;
        }
        st256(out + r * geo.C + c0, o);
    };
    const int64_t step = geo.RL;
    int64_t r = r_begin + rl;
    for (; r + 3 * step < r_end; r += 4 * step) {
        uint2 pk[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) pk[u] = __ldcs(reinterpret_cast<const uint2*>(codes + (r + u * step) * geo.C + c0));
#pragma unroll
        for (int u = 0; u < 4; ++u) one_row(pk[u], r + u * step);
    }
    for (; r < r_end; r += step) one_row(__ldcs(reinterpret_cast<const uint2*>(codes + r * geo.C + c0)), r);
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
    const bool quant = MODE == 0 && p.q_range != nullptr;
    if (C % 4 == 0 && quant)
        k_rbn_apply_nhwc<MODE, true, MODE == 0><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    else if (C % 4 == 0)
        k_rbn_apply_nhwc<MODE, true, false><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    else if (quant)
        k_rbn_apply_nhwc<MODE, false, MODE == 0><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    else
        k_rbn_apply_nhwc<MODE, false, false><<<grid, NH_THREADS, 0, st>>>(x, out, (uint32_t)n, dC, p);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_rangebn_apply_nhwc(const float* x, const uint8_t* x_codes, float* out, int64_t B, int64_t C, int64_t HW,
                                     const float* mean, const float* qscale, const float* qgamma, const float* qbeta,
                                     const float* q_range, int q_bits, void* stream) {
    QB_CHECK_ARG((x || x_codes) && out && mean && qscale, "qb_rangebn_apply_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_apply_nhwc: bad shape");
    QB_CHECK_ARG(!q_range || (q_bits >= 1 && q_bits <= 24), "qb_rangebn_apply_nhwc: q_bits=%d", q_bits);
    if (x_codes) {
        QB_CHECK_ARG(q_range && q_bits <= 8 && C % 8 == 0 && (reinterpret_cast<uintptr_t>(x_codes) & 7) == 0 &&
                         (reinterpret_cast<uintptr_t>(out) & 31) == 0,
                     "qb_rangebn_apply_nhwc: x_codes needs q_range, q_bits <= 8, C %% 8 == 0, aligned tensors");
        const int64_t R = B * HW;
        NhwcGeom geo = make_geom<8>(R, C, 1, 1 << 20);
        const int ctiles = (int)((C + geo.tileC - 1) / geo.tileC);
        static const int slots = nh_slots(k_rbn_apply_codes_nhwc, 0);
        int64_t want = (2 * slots) / ctiles;                                 // two full waves of resident CTAs
        if (want < 1) want = 1;
        const int64_t max_by_rows = R / ((int64_t)geo.RL * 4) > 0 ? R / ((int64_t)geo.RL * 4) : 1;
        if (want > max_by_rows) want = max_by_rows;
        geo.S = (int)want;
        geo.seg_rows = (R + geo.S - 1) / geo.S;
        dim3 grid((unsigned)geo.S, (unsigned)ctiles);
        k_rbn_apply_codes_nhwc<<<grid, NH_THREADS, 0, qb_stream(stream)>>>(x_codes, out, geo,
                                                                           NhwcVecs{mean, qscale, qgamma, qbeta, q_range, q_bits});
        QB_LAUNCH_CHECK();
        return QB_OK;
    }
    return launch_apply_nhwc<0>("qb_rangebn_apply_nhwc", x, out, B * HW, C,
                                NhwcVecs{mean, qscale, qgamma, qbeta, q_range, q_bits}, qb_stream(stream));
}

// ---------------------------------------------------------------------------------------------------
// backward reduce: sg[c] = sum_r gq[r][c], sgx[c] = sum_r gq[r][c] * x_hat[r][c]   (column sums, fp64)
// Fused form: g is the RAW output gradient and x the RAW RangeBN input; gq = Q_grad(g) (zero-point stochastic
// quantizer, noise regenerated from the element index) and x_hat = Q_act(x) are recomputed in registers, so neither
// quantized tensor is ever written to or read back from HBM.
// ---------------------------------------------------------------------------------------------------
struct GradQuantArg {
    const float* range;   // device {min,max} of g, or nullptr: g is already quantized
    int bits;
    int noise_mode;
    NoiseArg noise;
    uint8_t* codes;       // reduce pass: optional uint8 copy of the integer codes of Q(g) (bits <= 8), same indexing as g;
                          // apply pass (NOISE == -2): the codes to dequantize INSTEAD of reading and re-quantizing g
    const uint8_t* relu_bits = nullptr;   // reduce pass, VEC == 8: g is the gradient w.r.t. the OUTPUT of the ReLU that follows
                          // this RangeBN; bit e%8 of byte e/8 says whether it passes (written by qb_rangebn_apply_relu_measure)
};

// QUANT: 0 = x is x_hat already, 1 = x is the raw input, quantized on the fly with q_range, 2 = x_hat comes as the uint8
// codes the forward statistics pass wrote (x_codes; VEC == 8 only): 1 B/elem instead of 4 and no re-quantization
template <int VEC, int QUANT, int NOISE>   // NOISE < 0: g already quantized
__global__ void __launch_bounds__(NH_THREADS, 3)
k_rbn_bwd_reduce_nhwc(const float* __restrict__ gq, const float* __restrict__ x, const uint8_t* __restrict__ x_codes, NhwcGeom g,
                      double* part /*[2][S][C]*/, const float* q_range, int q_bits, GradQuantArg ga) {
    using V = typename VecT<VEC>::T;
    constexpr int UN = VEC == 8 ? 2 : NH_UNROLL;
    extern __shared__ unsigned char smem_raw[];
    AffineQ aq;
    if (QUANT) aq = load_affineq(q_range, q_bits);
    GradQ GQ;
    uint64_t seed = 0;
    if (NOISE >= 0) {
        GQ = make_gradq((double)__ldg(ga.range), (double)__ldg(ga.range + 1), ga.bits);
        seed = effective_seed(ga.noise);
    }
    const int ct = threadIdx.x % g.CT, rl = threadIdx.x / g.CT;
    const int64_t c0 = (int64_t)blockIdx.y * g.tileC + (int64_t)ct * VEC;
    const int s = blockIdx.x;
    const int64_t r_begin = (int64_t)s * g.seg_rows;
    const int64_t r_end = (r_begin + g.seg_rows < g.R) ? r_begin + g.seg_rows : g.R;
    const bool active = rl < g.RL && c0 < g.C;
    // thread-local partials stay in fp32: the geometry gives a thread at most ~100 rows (S up to 1024 segments), the
    // cross-thread / cross-segment combination below is fp64 in a fixed order
    float fg[VEC], fgx[VEC];
    uint64_t fg2[4] = {0ull, 0ull, 0ull, 0ull}, fgx2[4] = {0ull, 0ull, 0ull, 0ull};     // VEC == 8: packed (0.f, 0.f) pairs
#pragma unroll
    for (int j = 0; j < VEC; ++j) { fg[j] = 0.f; fgx[j] = 0.f; }
    auto one_row = [&](const V& av_, const V& bv_, const uint2& pk, int64_t r, uint32_t rmask = 0xffu) {
        float av[VEC], bv[VEC];
        unpack(av_, av);
        if constexpr (VEC == 8) {
            if (rmask != 0xffu) {
#pragma unroll
                for (int j = 0; j < 8; ++j) av[j] = ((rmask >> j) & 1u) ? av[j] : 0.0f;
            }
            if (QUANT == 2) affine_deq8(pk, aq.p, bv);                  // x_hat from the forward pass's codes
            else unpack(bv_, bv);
            if (NOISE >= 0) {
                float u[8];
                const int64_t e = r * g.C + c0;                         // a multiple of 8
                if (NOISE == NOISE_INJECT) {
                    const F8 w = ld256_stream(ga.noise.inject + e);
#pragma unroll
                    for (int j = 0; j < 8; ++j) u[j] = w.v[j];
                } else if (NOISE == NOISE_PHILOX) {
                    philox_noise8(seed, (uint64_t)(e >> 3) + (ga.noise.offset >> 3), u);
                }
                uint2 cd;
                grad_fq8<(NOISE < 0 ? 0 : NOISE)>(av, u, GQ, ga.codes ? &cd : nullptr);
                if (ga.codes) *reinterpret_cast<uint2*>(ga.codes + e) = cd;
            }
            if (QUANT == 1) affine_fq_vec<VEC>(bv, aq);
#pragma unroll
            for (int j = 0; j < 4; ++j) {                               // packed accumulation: FADD2 / FFMA2
                const uint64_t a2 = pack2(av[2 * j], av[2 * j + 1]);
                fg2[j] = add2(fg2[j], a2);
                fgx2[j] = fma2(a2, pack2(bv[2 * j], bv[2 * j + 1]), fgx2[j]);
            }
        } else {
            unpack(bv_, bv);
            if (QUANT == 1) affine_fq_vec<VEC>(bv, aq);
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                fg[j] += av[j];
                fgx[j] += av[j] * bv[j];
            }
        }
    };
    if (active) {
        int64_t r = r_begin + rl;
        const int64_t step = g.RL;
        for (; r + (UN - 1) * step < r_end; r += UN * step) {
            V a[UN], b[UN];
            uint2 pk[UN];
            uint32_t rm[UN];
#pragma unroll
            for (int u = 0; u < UN; ++u) {
                a[u] = ld_vec_stream<V>(gq + (r + u * step) * g.C + c0);
                if (QUANT == 2) pk[u] = __ldcs(reinterpret_cast<const uint2*>(x_codes + (r + u * step) * g.C + c0));
                else b[u] = ld_vec_stream<V>(x + (r + u * step) * g.C + c0);
                rm[u] = (VEC == 8 && ga.relu_bits) ? __ldcs(ga.relu_bits + (((r + u * step) * g.C + c0) >> 3)) : 0xffu;
            }
#pragma unroll
            for (int u = 0; u < UN; ++u) one_row(a[u], b[u], pk[u], r + u * step, rm[u]);
        }
        for (; r < r_end; r += step) {
            V b1;
            uint2 pk1 = make_uint2(0u, 0u);
            if (QUANT == 2) pk1 = __ldcs(reinterpret_cast<const uint2*>(x_codes + r * g.C + c0));
            else b1 = ld_vec_stream<V>(x + r * g.C + c0);
            const uint32_t rm1 = (VEC == 8 && ga.relu_bits) ? __ldcs(ga.relu_bits + ((r * g.C + c0) >> 3)) : 0xffu;
            one_row(ld_vec_stream<V>(gq + r * g.C + c0), b1, pk1, r, rm1);
        }
    }
    if constexpr (VEC == 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { unpack2(fg2[j], fg[2 * j], fg[2 * j + 1]); unpack2(fgx2[j], fgx[2 * j], fgx[2 * j + 1]); }
    }
    const int W = g.CT * VEC;
    double* s_g = reinterpret_cast<double*>(smem_raw);
    double* s_gx = s_g + (size_t)g.RL * W;
    if (rl < g.RL) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) { s_g[rl * W + ct * VEC + j] = (double)fg[j]; s_gx[rl * W + ct * VEC + j] = (double)fgx[j]; }
    }
    __syncthreads();
    // column c of the tile is finished by thread c: coalesced shared reads, RL terms in a fixed order
    for (int col = threadIdx.x; col < W; col += NH_THREADS) {
        const int64_t c = (int64_t)blockIdx.y * g.tileC + col;
        if (c >= g.C) break;
        double a = 0.0, b2 = 0.0;
        for (int q = 0; q < g.RL; ++q) { a += s_g[q * W + col]; b2 += s_gx[q * W + col]; }
        part[(int64_t)s * g.C + c] = a;
        part[((int64_t)g.S + s) * g.C + c] = b2;
    }
}

// one warp per channel: lanes stride over the segments, fixed-order shuffle tree => deterministic
template <bool COEFFS>
__global__ void k_rbn_bwd_pack_nhwc(const double* part, int64_t C, int S, double* sg, double* sgx, RbnCoefArg ca) {
    const int64_t c = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (c >= C) return;
    double a = 0.0, b = 0.0;
#pragma unroll 8                                                          // independent loads in flight; fixed order
    for (int s = lane; s < S; s += 32) { a += part[(int64_t)s * C + c]; b += part[((int64_t)S + s) * C + c]; }
    a = warp_sum(a);
    b = warp_sum(b);
    if (lane == 0) {
        sg[c] = a; sgx[c] = b;
        if (COEFFS) rbn_bwd_coeffs_channel(a, b, c, ca);                   // qb_rangebn_bwd_reduce_coeffs_nhwc: one launch less
    }
}

extern "C" size_t qb_rangebn_bwd_nhwc_ws_bytes(int64_t C) { return C < 0 ? 0 : (size_t)C * NH_MAX_SEG_BWD * 16; }

template <int VEC, int QUANT, int NOISE>
static void launch_bwd_reduce(dim3 grid, size_t smem, cudaStream_t st, const float* gq, const float* x, const uint8_t* x_codes,
                              const NhwcGeom& g, double* part, const float* q_range, int q_bits, const GradQuantArg& ga) {
    k_rbn_bwd_reduce_nhwc<VEC, QUANT, NOISE><<<grid, NH_THREADS, smem, st>>>(gq, x, x_codes, g, part, q_range, q_bits, ga);
}

static int grad_noise_mode(const float* g_range, const float* noise, int stochastic) {
    if (!g_range) return -1;
    return !stochastic ? NOISE_NONE : (noise ? NOISE_INJECT : NOISE_PHILOX);
}

/* g_range != NULL: g is the raw gradient, quantized on the fly to g_bits with this device range (needs C % 8 == 0 and
 * 32-byte aligned tensors); noise/stochastic/seed/offset/step_dev as in qb_fakequant_grad. */
static int bwd_reduce_nhwc_impl(const float* gq, const float* x, const uint8_t* x_codes, int64_t B, int64_t C, int64_t HW, double* sg,
                                double* sgx, const float* q_range, int q_bits, const float* g_range, int g_bits, const float* noise,
                                int stochastic, uint64_t seed, uint64_t offset, const uint64_t* step_dev, uint8_t* gq_codes_out,
                                const uint8_t* relu_bits, void* ws, size_t ws_bytes, void* stream, const RbnCoefArg* ca) {
    QB_CHECK_ARG(gq && (x || x_codes) && sg && sgx && ws, "qb_rangebn_bwd_reduce_nhwc: null pointer");
    QB_CHECK_ARG(!x_codes || (q_range && q_bits <= 8 && C % 8 == 0 && (reinterpret_cast<uintptr_t>(x_codes) & 7) == 0),
                 "qb_rangebn_bwd_reduce_nhwc: x_codes needs q_range, q_bits <= 8, C %% 8 == 0 and 8-byte alignment");
    QB_CHECK_ARG(!gq_codes_out || (g_range && g_bits <= 8 && (reinterpret_cast<uintptr_t>(gq_codes_out) & 7) == 0),
                 "qb_rangebn_bwd_reduce_nhwc: codes output needs the fused gradient quantizer, bits <= 8, 8-byte alignment");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_reduce_nhwc: bad shape");
    QB_CHECK_ARG(!q_range || (q_bits >= 1 && q_bits <= 24), "qb_rangebn_bwd_reduce_nhwc: q_bits=%d", q_bits);
    QB_CHECK_ARG(!g_range || (g_bits >= 1 && g_bits <= 24), "qb_rangebn_bwd_reduce_nhwc: g_bits=%d", g_bits);
    if (ws_bytes < qb_rangebn_bwd_nhwc_ws_bytes(C)) {
        qb_set_error("qb_rangebn_bwd_reduce_nhwc: workspace %zu < %zu bytes", ws_bytes, qb_rangebn_bwd_nhwc_ws_bytes(C));
        return QB_ERR_WORKSPACE;
    }
    const int64_t R = B * HW;
    const uintptr_t al = (x_codes ? 0 : reinterpret_cast<uintptr_t>(x)) | reinterpret_cast<uintptr_t>(gq) |
                         reinterpret_cast<uintptr_t>(noise);
    const bool vec8 = (C % 8 == 0) && ((al & 31) == 0) && (offset % 8 == 0);
    const bool vec4 = (C % 4 == 0) && ((al & 15) == 0);
    const int nm = grad_noise_mode(g_range, noise, stochastic);
    if ((nm >= 0 || x_codes) && !vec8) {
        qb_set_error("qb_rangebn_bwd_reduce_nhwc: on-the-fly gradient quantization / x_codes need C %% 8 == 0 and 32-byte alignment");
        return QB_ERR_UNSUPPORTED;
    }
    static const int slots8 = nh_slots(k_rbn_bwd_reduce_nhwc<8, 2, NOISE_PHILOX>, (size_t)NH_THREADS * 8 * 16);
    NhwcGeom g = vec8 ? make_geom<8>(R, C, 1, NH_MAX_SEG_BWD, slots8) : vec4 ? make_geom<4>(R, C, 1, NH_MAX_SEG_BWD)
                                                                              : make_geom<1>(R, C, 1, NH_MAX_SEG_BWD);
    double* part = static_cast<double*>(ws);
    dim3 grid((unsigned)g.S, (unsigned)((C + g.tileC - 1) / g.tileC));
    cudaStream_t st = qb_stream(stream);
    GradQuantArg ga{g_range, g_bits, nm, NoiseArg{noise, seed, offset, step_dev}, gq_codes_out, relu_bits};
    if (relu_bits && !(vec8 && nm >= 0)) {
        qb_set_error("qb_rangebn_bwd_reduce_nhwc: relu_bits needs the fused gradient quantizer on the 8-channel path");
        return QB_ERR_UNSUPPORTED;
    }
    const bool qx = q_range != nullptr;
    if (vec8) {
        const size_t sm = nhwc_smem(g, 8, 16);
#define QB_RED8(NM) (x_codes ? launch_bwd_reduce<8, 2, NM>(grid, sm, st, gq, x, x_codes, g, part, q_range, q_bits, ga) \
                     : qx ? launch_bwd_reduce<8, 1, NM>(grid, sm, st, gq, x, x_codes, g, part, q_range, q_bits, ga)    \
                          : launch_bwd_reduce<8, 0, NM>(grid, sm, st, gq, x, x_codes, g, part, q_range, q_bits, ga))
        if (nm < 0) QB_RED8(-1);
        else if (nm == NOISE_NONE) QB_RED8(NOISE_NONE);
        else if (nm == NOISE_INJECT) QB_RED8(NOISE_INJECT);
        else QB_RED8(NOISE_PHILOX);
#undef QB_RED8
    } else if (vec4) {
        if (qx) launch_bwd_reduce<4, 1, -1>(grid, nhwc_smem(g, 4, 16), st, gq, x, nullptr, g, part, q_range, q_bits, ga);
        else launch_bwd_reduce<4, 0, -1>(grid, nhwc_smem(g, 4, 16), st, gq, x, nullptr, g, part, q_range, q_bits, ga);
    } else {
        if (qx) launch_bwd_reduce<1, 1, -1>(grid, nhwc_smem(g, 1, 16), st, gq, x, nullptr, g, part, q_range, q_bits, ga);
        else launch_bwd_reduce<1, 0, -1>(grid, nhwc_smem(g, 1, 16), st, gq, x, nullptr, g, part, q_range, q_bits, ga);
    }
    QB_LAUNCH_CHECK();
    if (ca) k_rbn_bwd_pack_nhwc<true><<<(unsigned)((C * 32 + 127) / 128), 128, 0, st>>>(part, C, g.S, sg, sgx, *ca);
    else k_rbn_bwd_pack_nhwc<false><<<(unsigned)((C * 32 + 127) / 128), 128, 0, st>>>(part, C, g.S, sg, sgx, RbnCoefArg{});
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_rangebn_bwd_reduce_nhwc(const float* gq, const float* x, const uint8_t* x_codes, int64_t B, int64_t C,
                                          int64_t HW, double* sg, double* sgx, const float* q_range, int q_bits,
                                          const float* g_range, int g_bits,
                                          const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                                          const uint64_t* step_dev, uint8_t* gq_codes_out, const uint8_t* relu_bits, void* ws,
                                          size_t ws_bytes, void* stream) {
    return bwd_reduce_nhwc_impl(gq, x, x_codes, B, C, HW, sg, sgx, q_range, q_bits, g_range, g_bits, noise, stochastic, seed, offset,
                                step_dev, gq_codes_out, relu_bits, ws, ws_bytes, stream, nullptr);
}

/* qb_rangebn_bwd_reduce_nhwc followed by qb_rangebn_bwd_coeffs on the same stream in one call: the warp that folds a channel's
 * partial sums also derives its coefficients (same arithmetic).  Single-process use - across ranks the sums are exchanged
 * between the two calls. */
extern "C" int qb_rangebn_bwd_reduce_coeffs_nhwc(const float* gq, const float* x, const uint8_t* x_codes, int64_t B, int64_t C,
                                                 int64_t HW, double* sg, double* sgx, const float* q_range, int q_bits,
                                                 const float* g_range, int g_bits, const float* noise, int stochastic,
                                                 uint64_t seed, uint64_t offset, const uint64_t* step_dev, uint8_t* gq_codes_out,
                                                 const uint8_t* relu_bits, int num_chunks, double count, double chunk_len,
                                                 int training, const float* mean, const float* scale_raw, const float* qscale,
                                                 const float* qgamma, float* dgamma, float* dbeta, float* coef_a, float* coef_e,
                                                 void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(mean && scale_raw && qscale && coef_a && coef_e, "qb_rangebn_bwd_reduce_coeffs_nhwc: null pointer");
    QB_CHECK_ARG(!training || (num_chunks > 0 && count > 0 && chunk_len > 1), "qb_rangebn_bwd_reduce_coeffs_nhwc: bad chunking");
    const RbnCoefArg a{(double)num_chunks, count, training ? rbn_scale_fix(chunk_len) : 0.0, training, mean, scale_raw, qscale, qgamma,
                       dgamma, dbeta, coef_a, coef_e};
    return bwd_reduce_nhwc_impl(gq, x, x_codes, B, C, HW, sg, sgx, q_range, q_bits, g_range, g_bits, noise, stochastic, seed, offset,
                                step_dev, gq_codes_out, relu_bits, ws, ws_bytes, stream, &a);
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

// ---------------------------------------------------------------------------------------------------
// Fused backward apply (C % 8 == 0): one 32-byte octet = 8 adjacent channels of one row.
//   gq = Q_grad(g) recomputed in registers (same noise as the reduce pass: it is a function of the element index),
//   dx = (gq * qgamma_c) * qscale_c + a_c, then +e_c / -e_c where this row is the chunk's first arg-max / arg-min
//   (amax_kc / amin_kc are [chunk][channel] tables, one vector load each), and the global min / max of the FINAL dx
//   (the range UniformQuantizeGrad.backward of the preceding QConv2d needs, quantize.py:126,132) as a by-product.
// ---------------------------------------------------------------------------------------------------
struct BwdApplyArgs {
    const float *qgamma, *qscale, *coef_a, *coef_e;
    const int32_t *amax_kc, *amin_kc;
    uint32_t C;
    FastDiv dC, dL;          // row = e / C ; chunk = row / L
    FastDiv d_rl;            // rows a thread advances by per iteration (geometry RL)
    float* range_out;        // float[2] or nullptr
    float* pmin; float* pmax; unsigned int* counter;   // MinMaxWs-style partials (see fakequant_grad.cu)
};

template <int NOISE, bool RANGE>
__global__ void __launch_bounds__(NH_THREADS, 3)
k_rbn_bwd_apply_fused(const float* __restrict__ g, float* __restrict__ dx, NhwcGeom geo, BwdApplyArgs A, GradQuantArg ga) {
    // Column-persistent mapping (same geometry as the statistics kernel): a thread owns 8 adjacent channels for the whole
    // CTA lifetime, so the per-channel coefficients and the chunk's arg rows live in registers - the row loop touches
    // memory only for the data itself (the first version re-loaded 10 parameter vectors per octet and was L1-bound).
    __shared__ float sm[64];
    __shared__ bool last;
    GradQ GQ;
    uint64_t seed = 0;
    if (NOISE >= 0 || NOISE == -2) {
        GQ = make_gradq((double)__ldg(ga.range), (double)__ldg(ga.range + 1), ga.bits);
        seed = effective_seed(ga.noise);
    }
    float mn = QB_INF, mx = -QB_INF;
    const int ct = threadIdx.x % geo.CT, rl = threadIdx.x / geo.CT;
    const int64_t c0 = (int64_t)blockIdx.y * geo.tileC + (int64_t)ct * 8;
    const int k = blockIdx.x / geo.S, sgm = blockIdx.x % geo.S;
    const int64_t r_begin = (int64_t)k * geo.L + (int64_t)sgm * geo.seg_rows;
    int64_t r_end = r_begin + geo.seg_rows;
    if (r_end > (int64_t)(k + 1) * geo.L) r_end = (int64_t)(k + 1) * geo.L;
    if (rl < geo.RL && c0 < geo.C) {
        float sc[8], ca[8], gm[8], ce[8];
        int32_t am[8], an[8];
        // rows this thread visits: r0 + i*RL.  A 64-bit map marks the few i at which one of the 16 arg rows of its
        // channels falls, so the +-e fix-up costs one bit test per ROW instead of 16 compares per row
        const int64_t r0 = r_begin + rl;
        uint64_t hitmap = 0ull;
        bool hit_far = false;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            sc[j] = __ldg(A.qscale + c0 + j);
            ca[j] = __ldg(A.coef_a + c0 + j);
            gm[j] = A.qgamma ? __ldg(A.qgamma + c0 + j) : 1.0f;
            ce[j] = A.amax_kc ? __ldg(A.coef_e + c0 + j) : 0.0f;
            am[j] = A.amax_kc ? __ldg(A.amax_kc + (size_t)k * A.C + c0 + j) : -1;
            an[j] = A.amax_kc ? __ldg(A.amin_kc + (size_t)k * A.C + c0 + j) : -1;
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const int64_t d = (int64_t)(w ? an[j] : am[j]) - r0;
                if (A.amax_kc && d >= 0 && d < (int64_t)geo.seg_rows) {
                    const uint32_t i = fdiv((uint32_t)d, A.d_rl);
                    if ((int64_t)i * geo.RL == d) {
                        if (i < 64) hitmap |= 1ull << i;
                        else hit_far = true;
                    }
                }
            }
        }
        uint64_t sc2[4], gm2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sc2[j] = pack2(sc[2 * j], sc[2 * j + 1]);
            gm2[j] = pack2(gm[2 * j], gm[2 * j + 1]);
        }
        auto one_row = [&](const F8& raw, const uint2& pk, int64_t r, uint32_t i) {
            const int64_t e = r * geo.C + c0;
            float gq[8];
            if (NOISE == -2) {                            // 8 codes instead of 8 floats: dequantize only
                zp_deq8(pk, GQ.p, gq);
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) gq[j] = raw.v[j];
                if (NOISE >= 0) {
                    float u[8];
                    if (NOISE == NOISE_INJECT) {
                        const F8 w = ld256_stream(ga.noise.inject + e);
#pragma unroll
                        for (int j = 0; j < 8; ++j) u[j] = w.v[j];
                    } else if (NOISE == NOISE_PHILOX) {
                        philox_noise8(seed, (uint64_t)(e >> 3) + (ga.noise.offset >> 3), u);
                    }
                    grad_fq8<(NOISE < 0 ? 0 : NOISE)>(gq, u, GQ);
                }
            }
            F8 o;
#pragma unroll
            for (int j = 0; j < 4; ++j) {                 // (gq*gamma)*scale + a, every operation separately rounded: packed
                uint64_t t = pack2(gq[2 * j], gq[2 * j + 1]);      // multiplies, SCALAR adds (ptxas would contract a packed
                if (A.qgamma) t = mul2(t, gm2[j]);                 // mul + add pair into one FFMA2)
                float a, b;
                unpack2(mul2(t, sc2[j]), a, b);
                o.v[2 * j] = __fadd_rn(a, ca[2 * j]);
                o.v[2 * j + 1] = __fadd_rn(b, ca[2 * j + 1]);
            }
            if ((i < 64 ? ((hitmap >> i) & 1ull) != 0ull : hit_far)) {
                const int32_t row = (int32_t)r;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    if (am[j] == row) o.v[j] = __fadd_rn(o.v[j], ce[j]);
                    if (an[j] == row) o.v[j] = __fsub_rn(o.v[j], ce[j]);
                }
            }
            if (RANGE) {
#pragma unroll
                for (int j = 0; j < 8; j += 2) { mn = min3(mn, o.v[j], o.v[j + 1]); mx = max3(mx, o.v[j], o.v[j + 1]); }
            }
            st256(dx + e, o);                              // default policy: the next kernel re-reads dx (L2)
        };
        int64_t r = r_begin + rl;
        const int64_t step = geo.RL;
        uint32_t it = 0;
        if (NOISE == -2) {
            // code rows are only 8 bytes per thread: keep four of them in flight to cover the memory latency
            for (; r + 3 * step < r_end; r += 4 * step, it += 4) {
                uint2 pk[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) pk[u] = __ldcs(reinterpret_cast<const uint2*>(ga.codes + (r + u * step) * geo.C + c0));
#pragma unroll
                for (int u = 0; u < 4; ++u) one_row(F8{}, pk[u], r + u * step, it + u);
            }
        } else {
            for (; r + step < r_end; r += 2 * step, it += 2) {
                const F8 v0 = ld256_stream(g + r * geo.C + c0), v1 = ld256_stream(g + (r + step) * geo.C + c0);
                one_row(v0, make_uint2(0, 0), r, it);
                one_row(v1, make_uint2(0, 0), r + step, it + 1);
            }
        }
        for (; r < r_end; r += step, ++it) {
            F8 v0;
            uint2 p0 = make_uint2(0, 0);
            if (NOISE == -2) p0 = __ldcs(reinterpret_cast<const uint2*>(ga.codes + r * geo.C + c0));
            else v0 = ld256_stream(g + r * geo.C + c0);
            one_row(v0, p0, r, it);
        }
    }
    const unsigned int nblocks = gridDim.x * gridDim.y, bid = blockIdx.y * gridDim.x + blockIdx.x;
    if (!RANGE) return;
    mn = warp_min(mn);
    mx = warp_max(mx);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) { sm[w] = mn; sm[32 + w] = mx; }
    __syncthreads();
    if (w == 0) {
        mn = warp_min(l < NH_THREADS / 32 ? sm[l] : QB_INF);
        mx = warp_max(l < NH_THREADS / 32 ? sm[32 + l] : -QB_INF);
        if (l == 0) {
            A.pmin[bid] = mn;
            A.pmax[bid] = mx;
            __threadfence();
            last = (atomicAdd(A.counter, 1u) == nblocks - 1);
        }
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    mn = QB_INF;
    mx = -QB_INF;
    for (int i = threadIdx.x; i < (int)nblocks; i += NH_THREADS) {
        mn = fminf(mn, __ldcg(&A.pmin[i]));
        mx = fmaxf(mx, __ldcg(&A.pmax[i]));
    }
    mn = warp_min(mn);
    mx = warp_max(mx);
    __syncthreads();
    if (l == 0) { sm[w] = mn; sm[32 + w] = mx; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 1; q < NH_THREADS / 32; ++q) { mn = fminf(mn, sm[q]); mx = fmaxf(mx, sm[32 + q]); }
        A.range_out[0] = mn;
        A.range_out[1] = mx;
        *A.counter = 0u;
    }
}

template <int NOISE>
static void launch_bwd_apply_fused(dim3 grid, cudaStream_t st, const float* g, float* dx, const NhwcGeom& geo,
                                   const BwdApplyArgs& A, const GradQuantArg& ga) {
    if (A.range_out) k_rbn_bwd_apply_fused<NOISE, true><<<grid, NH_THREADS, 0, st>>>(g, dx, geo, A, ga);
    else k_rbn_bwd_apply_fused<NOISE, false><<<grid, NH_THREADS, 0, st>>>(g, dx, geo, A, ga);
}

/* Fused variant used when C % 8 == 0: g_range != NULL quantizes g on the fly (as in qb_rangebn_bwd_reduce_nhwc);
 * amax_kc / amin_kc are the [chunk][channel] arg tables of qb_rangebn_stats_nhwc; dx_range_out (float[2], optional)
 * receives min/max of dx; mm_ws: qb_minmax_ws_bytes() bytes, zero-filled once. */
extern "C" int qb_rangebn_bwd_apply_fused_nhwc(const float* g, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                                               const float* qscale, const float* qgamma, const float* coef_a,
                                               const float* coef_e, const int32_t* amax_kc, const int32_t* amin_kc,
                                               const float* g_range, int g_bits, const float* noise, int stochastic,
                                               uint64_t seed, uint64_t offset, const uint64_t* step_dev,
                                               const uint8_t* gq_codes, float* dx_range_out, void* mm_ws,
                                               size_t mm_ws_bytes, void* stream) {
    QB_CHECK_ARG((g || gq_codes) && dx && qscale && coef_a, "qb_rangebn_bwd_apply_fused_nhwc: null pointer");
    QB_CHECK_ARG(!gq_codes || (g_range && g_bits <= 8 && (reinterpret_cast<uintptr_t>(gq_codes) & 7) == 0),
                 "qb_rangebn_bwd_apply_fused_nhwc: codes input needs g_range, bits <= 8, 8-byte alignment");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_apply_fused_nhwc: bad shape");
    QB_CHECK_ARG(!g_range || (g_bits >= 1 && g_bits <= 24), "qb_rangebn_bwd_apply_fused_nhwc: g_bits=%d", g_bits);
    const int64_t R = B * HW, n = R * C;
    const uintptr_t al = reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(dx) | reinterpret_cast<uintptr_t>(noise);
    if (C % 8 != 0 || (al & 31) != 0 || (offset & 7) != 0 || n >= 4294967296LL) {
        qb_set_error("qb_rangebn_bwd_apply_fused_nhwc: needs C %% 8 == 0, 32-byte aligned tensors, < 2^32 elements");
        return QB_ERR_UNSUPPORTED;
    }
    const bool scatter = coef_e && amax_kc && amin_kc && num_chunks > 0;
    QB_CHECK_ARG(!scatter || R % num_chunks == 0, "qb_rangebn_bwd_apply_fused_nhwc: rows not divisible by chunks");
    QB_CHECK_ARG(!dx_range_out || (mm_ws && mm_ws_bytes >= (size_t)(2 * 4096 + 16) * 4),
                 "qb_rangebn_bwd_apply_fused_nhwc: range output needs the min/max workspace");
    BwdApplyArgs A;
    A.qgamma = qgamma; A.qscale = qscale; A.coef_a = coef_a; A.coef_e = coef_e;
    A.amax_kc = scatter ? amax_kc : nullptr; A.amin_kc = scatter ? amin_kc : nullptr;
    A.C = (uint32_t)C;
    A.dC = make_fastdiv((uint32_t)C);
    A.dL = make_fastdiv((uint32_t)(scatter ? R / num_chunks : 1));
    A.range_out = dx_range_out;
    float* wsf = static_cast<float*>(mm_ws);
    A.pmin = wsf; A.pmax = wsf ? wsf + 4096 : nullptr;
    A.counter = wsf ? reinterpret_cast<unsigned int*>(wsf + 8192) : nullptr;
    const int nm = gq_codes ? -2 : grad_noise_mode(g_range, noise, stochastic);
    GradQuantArg ga{g_range, g_bits, nm, NoiseArg{noise, seed, offset, step_dev}, const_cast<uint8_t*>(gq_codes)};
    // chunk-aligned row segments (a CTA never straddles two chunks, so its arg rows are loop invariant); enough CTAs
    // to fill the machine several times, at most 4096 partial slots for the range reduction
    const int Kc = scatter ? num_chunks : 1;
    NhwcGeom geo = make_geom<8>(R, C, Kc, 256);
    {
        const int ctiles = (int)((C + geo.tileC - 1) / geo.tileC);
        int sms = qb_sm_count();
        if (sms <= 0) sms = 148;
        int64_t want = (int64_t)(sms * 3 * 2) / ((int64_t)Kc * ctiles);      // two full waves at 3 resident CTAs per SM (80 registers)
        const int64_t max_by_rows = geo.L / ((int64_t)geo.RL * 2) > 0 ? geo.L / ((int64_t)geo.RL * 2) : 1;
        if (want > max_by_rows) want = max_by_rows;
        while ((int64_t)Kc * want * ctiles > 4096 && want > 1) --want;
        geo.S = (int)(want < 1 ? 1 : want);
        geo.seg_rows = (geo.L + geo.S - 1) / geo.S;
        QB_CHECK_ARG((int64_t)Kc * geo.S * ctiles <= 4096, "qb_rangebn_bwd_apply_fused_nhwc: too many chunk x channel tiles");
    }
    A.d_rl = make_fastdiv((uint32_t)geo.RL);
    dim3 grid((unsigned)(Kc * geo.S), (unsigned)((C + geo.tileC - 1) / geo.tileC));
    cudaStream_t st = qb_stream(stream);
    if (nm == -2) launch_bwd_apply_fused<-2>(grid, st, g, dx, geo, A, ga);
    else if (nm < 0) launch_bwd_apply_fused<-1>(grid, st, g, dx, geo, A, ga);
    else if (nm == NOISE_NONE) launch_bwd_apply_fused<NOISE_NONE>(grid, st, g, dx, geo, A, ga);
    else if (nm == NOISE_INJECT) launch_bwd_apply_fused<NOISE_INJECT>(grid, st, g, dx, geo, A, ga);
    else launch_bwd_apply_fused<NOISE_PHILOX>(grid, st, g, dx, geo, A, ga);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_rangebn_bwd_apply_nhwc(const float* gq, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                                         const float* qscale, const float* qgamma, const float* coef_a,
                                         const float* coef_e, const int32_t* amax, const int32_t* amin, void* stream) {
    QB_CHECK_ARG(gq && dx && qscale && coef_a, "qb_rangebn_bwd_apply_nhwc: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0 && HW > 0, "qb_rangebn_bwd_apply_nhwc: bad shape");
    cudaStream_t st = qb_stream(stream);
    int rc = launch_apply_nhwc<1>("qb_rangebn_bwd_apply_nhwc", gq, dx, B * HW, C, NhwcVecs{qgamma, qscale, coef_a, nullptr, nullptr, 0}, st);
    if (rc) return rc;
    if (coef_e && amax && amin && num_chunks > 0) {
        const int64_t t = C * num_chunks;
        k_rbn_bwd_scatter_nhwc<<<(unsigned)((t + 127) / 128), 128, 0, st>>>(dx, C, num_chunks, coef_e, amax, amin);
        QB_LAUNCH_CHECK();
    }
    return QB_OK;
}
