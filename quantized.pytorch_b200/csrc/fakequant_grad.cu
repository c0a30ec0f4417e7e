// fakequant_grad.cu - gradient fake-quantization (zero-point + stochastic branch) and global min/max.
//
// UniformQuantizeGrad.backward (models/modules/quantize.py:124-139): min/max of the whole gradient, then
// UniformQuantize.forward with enforce_true_zero=True, stochastic=True (:77-88, 92-98).
#include <cooperative_groups.h>

#include <type_traits>

#include "common.cuh"

namespace cg = cooperative_groups;

constexpr int GQ_THREADS = 256;
constexpr int GQ_U = 2;              // 32-byte octets in flight per thread in the quantizing sweep
constexpr int MM_U = 4;              // ... in the min/max sweep (no stores, few registers)
constexpr int MM_MAX_BLOCKS = 4096;  // partial slots in the workspace

// workspace layout: float partial_min[MM_MAX_BLOCKS], float partial_max[MM_MAX_BLOCKS], uint32 counter
struct MinMaxWs {
    float pmin[MM_MAX_BLOCKS];
    float pmax[MM_MAX_BLOCKS];
    unsigned int counter;
    unsigned int pad[15];
};

extern "C" size_t qb_minmax_ws_bytes(void) { return sizeof(MinMaxWs); }

__device__ __forceinline__ void block_reduce_minmax(float& mn, float& mx, float* sm) {
    mn = warp_min(mn);
    mx = warp_max(mx);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) {
        sm[w] = mn;
        sm[32 + w] = mx;
    }
    __syncthreads();
    const int nw = GQ_THREADS / 32;
    mn = warp_min(l < nw ? sm[l] : QB_INF);
    mx = warp_max(l < nw ? sm[32 + l] : -QB_INF);
}

__device__ __forceinline__ void octet_minmax(const F8& v, float& mn, float& mx) {
    const float a = fminf(fminf(v.v[0], v.v[1]), fminf(v.v[2], v.v[3])), b = fminf(fminf(v.v[4], v.v[5]), fminf(v.v[6], v.v[7]));
    const float c = fmaxf(fmaxf(v.v[0], v.v[1]), fmaxf(v.v[2], v.v[3])), d = fmaxf(fmaxf(v.v[4], v.v[5]), fmaxf(v.v[6], v.v[7]));
    mn = fminf(mn, fminf(a, b));
    mx = fmaxf(mx, fmaxf(c, d));
}

// per-thread min/max over a grid-strided sweep of n floats (256-bit body when 32-byte aligned, scalar tail).
// Default cache policy on purpose: the fused kernel wants these lines to stay in L2 for its second pass.
__device__ __forceinline__ void sweep_minmax(const float* __restrict__ x, int64_t n, bool vec, float& mn, float& mx) {
    const int64_t stride = (int64_t)gridDim.x * GQ_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x;
    if (vec) {
        const int64_t noct = n >> 3;
        int64_t i = tid;
        for (; i + (MM_U - 1) * stride < noct; i += MM_U * stride) {
            F8 v[MM_U];
#pragma unroll
            for (int u = 0; u < MM_U; ++u) v[u] = ld256_keep(x + 8 * (i + u * stride));
#pragma unroll
            for (int u = 0; u < MM_U; ++u) octet_minmax(v[u], mn, mx);
        }
        for (; i < noct; i += stride) octet_minmax(ld256_keep(x + 8 * i), mn, mx);
        if (tid < (n & 7)) {
            const float v = x[(noct << 3) + tid];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    } else {
        for (int64_t e = tid; e < n; e += stride) {
            const float v = x[e];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    }
}

// global min/max, one launch, last-block-done finalisation (deterministic; leaves the counter at 0)
__global__ void __launch_bounds__(GQ_THREADS)
k_minmax(const float* __restrict__ x, int64_t n, int vec, float* range_out, MinMaxWs* ws) {
    __shared__ float sm[64];
    __shared__ bool last;
    float mn = QB_INF, mx = -QB_INF;
    sweep_minmax(x, n, vec != 0, mn, mx);
    block_reduce_minmax(mn, mx, sm);
    if (threadIdx.x == 0) {
        ws->pmin[blockIdx.x] = mn;
        ws->pmax[blockIdx.x] = mx;
        __threadfence();
        unsigned int t = atomicAdd(&ws->counter, 1u);
        last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    mn = QB_INF;
    mx = -QB_INF;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += GQ_THREADS) {
        mn = fminf(mn, __ldcg(&ws->pmin[i]));
        mx = fmaxf(mx, __ldcg(&ws->pmax[i]));
    }
    block_reduce_minmax(mn, mx, sm);
    if (threadIdx.x == 0) {
        range_out[0] = mn;
        range_out[1] = mx;
        ws->counter = 0;
    }
}

static bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }

static int gq_grid(int64_t items, int per_sm) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (items + GQ_THREADS - 1) / GQ_THREADS;
    int64_t cap = (int64_t)sms * per_sm;
    if (cap > MM_MAX_BLOCKS) cap = MM_MAX_BLOCKS;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

extern "C" int qb_minmax(const float* x, int64_t n, float* range_out, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x && range_out && ws, "qb_minmax: null pointer");
    QB_CHECK_ARG(n > 0, "qb_minmax: n=%lld (min/max of an empty tensor is undefined, as in torch)", (long long)n);
    if (ws_bytes < sizeof(MinMaxWs)) {
        qb_set_error("qb_minmax: workspace %zu < %zu bytes", ws_bytes, sizeof(MinMaxWs));
        return QB_ERR_WORKSPACE;
    }
    const int vec = aligned32(x);
    int grid = gq_grid(vec ? (n >> 3) / MM_U : n / MM_U, 8);
    k_minmax<<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(x, n, vec, range_out, static_cast<MinMaxWs*>(ws));
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// ReLU backward + the range of its result in one pass (SURVEY section 8f N1 glue).  dx = out <= 0 ? 0 : g is what
// aten::threshold_backward computes for nn.ReLU (models/resnet_quantized.py:47-75); the gradient quantizer of the RangeBN
// / QConv2d that receives dx needs float(dx.min()), float(dx.max()) next (quantize.py:126,132), so the min/max sweep
// (a second 4 B/elem read) rides along: 12 B/elem for both instead of 12 + 4.
// ------------------------------------------------------------------------------------------------
// BITS: the ReLU output is given as the bit mask qb_rangebn_apply_relu_measure wrote (bit e%8 of byte e/8 = "out[e] is not
// <= 0"): 1/8 B/elem instead of 4, octets only.  WRITE = false: only the range of the masked gradient is produced (the
// RangeBN backward that consumes it applies the same mask on the fly), 4.125 B/elem in total.
template <bool BITS, bool WRITE>
__global__ void __launch_bounds__(GQ_THREADS)
k_relu_bwd_minmax(const float* __restrict__ g, const float* __restrict__ out, float* __restrict__ dx, int64_t n, int vec,
                  float* range_out, MinMaxWs* ws) {
    __shared__ float sm[64];
    __shared__ bool last;
    float mn = QB_INF, mx = -QB_INF;
    const int64_t stride = (int64_t)gridDim.x * GQ_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x;
    if (BITS) {
        const uint8_t* bits = reinterpret_cast<const uint8_t*>(out);
        const int64_t noct = n >> 3;
        constexpr int U = 4;
        int64_t i = tid;
        for (; i + (U - 1) * stride < noct; i += U * stride) {
            F8 a[U];
            uint32_t m[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                a[u] = ld256_stream(g + 8 * (i + u * stride));
                m[u] = __ldcs(bits + i + u * stride);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[u].v[j] = ((m[u] >> j) & 1u) ? a[u].v[j] : 0.0f;
                octet_minmax(a[u], mn, mx);
                if (WRITE) *reinterpret_cast<F8*>(dx + 8 * (i + u * stride)) = a[u];
            }
        }
        for (; i < noct; i += stride) {
            F8 a = ld256_stream(g + 8 * i);
            const uint32_t m = __ldcs(bits + i);
#pragma unroll
            for (int j = 0; j < 8; ++j) a.v[j] = ((m >> j) & 1u) ? a.v[j] : 0.0f;
            octet_minmax(a, mn, mx);
            if (WRITE) *reinterpret_cast<F8*>(dx + 8 * i) = a;
        }
    } else if (vec) {
        const int64_t noct = n >> 3;
        constexpr int U = 2;
        int64_t i = tid;
        for (; i + (U - 1) * stride < noct; i += U * stride) {
            F8 a[U], b[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                a[u] = ld256_stream(g + 8 * (i + u * stride));
                b[u] = ld256_stream(out + 8 * (i + u * stride));
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
#pragma unroll
                for (int j = 0; j < 8; ++j) a[u].v[j] = b[u].v[j] <= 0.0f ? 0.0f : a[u].v[j];
                octet_minmax(a[u], mn, mx);
                *reinterpret_cast<F8*>(dx + 8 * (i + u * stride)) = a[u];   // default policy: the consumer reads it next
            }
        }
        for (; i < noct; i += stride) {
            F8 a = ld256_stream(g + 8 * i);
            const F8 b = ld256_stream(out + 8 * i);
#pragma unroll
            for (int j = 0; j < 8; ++j) a.v[j] = b.v[j] <= 0.0f ? 0.0f : a.v[j];
            octet_minmax(a, mn, mx);
            *reinterpret_cast<F8*>(dx + 8 * i) = a;
        }
        if (tid < (n & 7)) {
            const int64_t e = (noct << 3) + tid;
            const float v = out[e] <= 0.0f ? 0.0f : g[e];
            dx[e] = v;
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    } else {
        for (int64_t e = tid; e < n; e += stride) {
            const float v = out[e] <= 0.0f ? 0.0f : g[e];
            dx[e] = v;
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    }
    block_reduce_minmax(mn, mx, sm);
    if (threadIdx.x == 0) {
        ws->pmin[blockIdx.x] = mn;
        ws->pmax[blockIdx.x] = mx;
        __threadfence();
        last = (atomicAdd(&ws->counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    mn = QB_INF;
    mx = -QB_INF;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += GQ_THREADS) {
        mn = fminf(mn, __ldcg(&ws->pmin[i]));
        mx = fmaxf(mx, __ldcg(&ws->pmax[i]));
    }
    block_reduce_minmax(mn, mx, sm);
    if (threadIdx.x == 0) {
        range_out[0] = mn;
        range_out[1] = mx;
        ws->counter = 0;
    }
}

extern "C" int qb_relu_bwd_minmax(const float* grad_out, const float* relu_out, float* dx, int64_t n, float* range_out, void* ws,
                                  size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(grad_out && relu_out && dx && range_out && ws, "qb_relu_bwd_minmax: null pointer");
    QB_CHECK_ARG(n > 0, "qb_relu_bwd_minmax: n=%lld", (long long)n);
    if (ws_bytes < sizeof(MinMaxWs)) {
        qb_set_error("qb_relu_bwd_minmax: workspace %zu < %zu bytes", ws_bytes, sizeof(MinMaxWs));
        return QB_ERR_WORKSPACE;
    }
    const int vec = aligned32(grad_out) && aligned32(relu_out) && aligned32(dx);
    const int grid = gq_grid(vec ? (n >> 3) / 2 : n / 2, 8);
    k_relu_bwd_minmax<false, true><<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(grad_out, relu_out, dx, n, vec, range_out,
                                                                              static_cast<MinMaxWs*>(ws));
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* qb_relu_bwd_minmax with the ReLU output given as the bit mask written by qb_rangebn_apply_relu_measure (bit e%8 of byte
 * e/8 set = gradient passes).  dx == NULL: only range_out = {min, max} of the masked gradient is produced (4.125 B/elem) -
 * for a consumer that applies the mask itself (qb_rangebn_bwd_reduce_nhwc's relu_bits).  n % 8 == 0, 32-byte aligned. */
extern "C" int qb_relu_bwd_minmax_bits(const float* grad_out, const uint8_t* relu_bits, float* dx, int64_t n, float* range_out,
                                       void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(grad_out && relu_bits && range_out && ws, "qb_relu_bwd_minmax_bits: null pointer");
    QB_CHECK_ARG(n > 0 && n % 8 == 0, "qb_relu_bwd_minmax_bits: n=%lld (a positive multiple of 8)", (long long)n);
    QB_CHECK_ARG(aligned32(grad_out) && (!dx || aligned32(dx)), "qb_relu_bwd_minmax_bits: tensors must be 32-byte aligned");
    if (ws_bytes < sizeof(MinMaxWs)) {
        qb_set_error("qb_relu_bwd_minmax_bits: workspace %zu < %zu bytes", ws_bytes, sizeof(MinMaxWs));
        return QB_ERR_WORKSPACE;
    }
    const int grid = gq_grid((n >> 3) / 4, 8);
    const float* bits = reinterpret_cast<const float*>(relu_bits);
    if (dx) k_relu_bwd_minmax<true, true><<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(grad_out, bits, dx, n, 1, range_out, static_cast<MinMaxWs*>(ws));
    else k_relu_bwd_minmax<true, false><<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(grad_out, bits, nullptr, n, 1, range_out, static_cast<MinMaxWs*>(ws));
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// zero-point stochastic quantizer (A2), range given
// ------------------------------------------------------------------------------------------------
template <int NOISE, bool WRITE_Y, bool WRITE_C>
__device__ __forceinline__ void grad_elem(const float* g, float* out, uint8_t* codes, int64_t e, const GradQ& Q,
                                          const NoiseArg& na, uint64_t seed) {
    float t = __fadd_rn(qb_div(g[e], Q.d), Q.p.zp);
    if (NOISE != NOISE_NONE) t = __fadd_rn(t, noise_elem<NOISE>(na, seed, e));
    const float q = qb_rne(fminf(fmaxf(t, 0.0f), Q.p.qmax), Q.p);
    if (WRITE_Y) out[e] = q_zp_deq(q, Q.p);
    if (WRITE_C) codes[e] = (uint8_t)q;
}

// one 32-byte octet: 8 quotients, zero point, noise, clamp, round, dequantize.  SMALL (qmax < 2^22): rounding and the
// integer code both come out of one magic-number add (no FRND / F2I on the conversion pipe).
template <int NOISE, bool WRITE_Y, bool WRITE_C, bool SMALL>
__device__ __forceinline__ void grad_octet(const F8& v, float* out, uint8_t* codes, int64_t oct, const GradQ& Q,
                                           const NoiseArg& na, uint64_t seed) {
    float u[8];
    if (NOISE == NOISE_INJECT) {
        const F8 w = ld256_stream(na.inject + 8 * oct);
#pragma unroll
        for (int j = 0; j < 8; ++j) u[j] = w.v[j];
    } else if (NOISE == NOISE_PHILOX) {
        philox_noise8(seed, (uint64_t)oct + (na.offset >> 3), u);
    }
    bool ok = Q.d.fast;
#pragma unroll
    for (int j = 0; j < 8; ++j) ok = ok && qb_div_ok(v.v[j]);
    float t[8];
    if (ok) {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = qb_div_fast(v.v[j], Q.d);
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = __fdiv_rn(v.v[j], Q.d.b);
    }
    F8 y;
    uint32_t c[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float s = __fadd_rn(t[j], Q.p.zp);
        if (NOISE != NOISE_NONE) s = __fadd_rn(s, u[j]);
        s = fminf(fmaxf(s, 0.0f), Q.p.qmax);
        float q;
        if (SMALL) {
            const float m = __fadd_rn(s, QB_MAGIC);
            c[j] = __float_as_uint(m);                       // low byte = the code (bits <= 8 whenever codes are written)
            q = __fsub_rn(m, QB_MAGIC);
        } else {
            q = rintf(s);
            c[j] = (uint32_t)q;
        }
        y.v[j] = q_zp_deq(q, Q.p);
    }
    if (WRITE_Y) st256_stream(out + 8 * oct, y);
    if (WRITE_C) {
        uint2 pk;
        pk.x = __byte_perm(__byte_perm(c[0], c[1], 0x0040), __byte_perm(c[2], c[3], 0x0040), 0x5410);
        pk.y = __byte_perm(__byte_perm(c[4], c[5], 0x0040), __byte_perm(c[6], c[7], 0x0040), 0x5410);
        reinterpret_cast<uint2*>(codes)[oct] = pk;
    }
}

// Body shared by the given-range kernel and phase 2 of the fused kernel.  REVERSE sweeps the tensor from
// its end so that the fused kernel's second read starts with the lines its first read left in L2.
template <int NOISE, bool WRITE_Y, bool WRITE_C, bool REVERSE, bool SMALL>
__device__ __forceinline__ void sweep_grad_impl(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, bool vec,
                                                const GradQ& Q, const NoiseArg& na) {
    const uint64_t seed = effective_seed(na);
    const int64_t stride = (int64_t)gridDim.x * GQ_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x;
    if (vec) {
        const int64_t noct = n >> 3;
        int64_t i = tid;
        for (; i + (GQ_U - 1) * stride < noct; i += GQ_U * stride) {
            F8 v[GQ_U];
            int64_t od[GQ_U];
#pragma unroll
            for (int u = 0; u < GQ_U; ++u) {
                od[u] = REVERSE ? noct - 1 - (i + u * stride) : i + u * stride;
                v[u] = ld256_stream(g + 8 * od[u]);
            }
#pragma unroll
            for (int u = 0; u < GQ_U; ++u) grad_octet<NOISE, WRITE_Y, WRITE_C, SMALL>(v[u], out, codes, od[u], Q, na, seed);
        }
        for (; i < noct; i += stride) {
            const int64_t od = REVERSE ? noct - 1 - i : i;
            grad_octet<NOISE, WRITE_Y, WRITE_C, SMALL>(ld256_stream(g + 8 * od), out, codes, od, Q, na, seed);
        }
        if (tid < (n & 7)) grad_elem<NOISE, WRITE_Y, WRITE_C>(g, out, codes, (noct << 3) + tid, Q, na, seed);
    } else {
        for (int64_t e = tid; e < n; e += stride) grad_elem<NOISE, WRITE_Y, WRITE_C>(g, out, codes, e, Q, na, seed);
    }
}
template <int NOISE, bool WRITE_Y, bool WRITE_C, bool REVERSE>
__device__ __forceinline__ void sweep_grad(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, bool vec,
                                           const GradQ& Q, const NoiseArg& na) {
    if (Q.p.qmax < 4194304.0f)
        sweep_grad_impl<NOISE, WRITE_Y, WRITE_C, REVERSE, true>(g, out, codes, n, vec, Q, na);
    else
        sweep_grad_impl<NOISE, WRITE_Y, WRITE_C, REVERSE, false>(g, out, codes, n, vec, Q, na);
}

template <int NOISE, bool WRITE_Y, bool WRITE_C>
__global__ void __launch_bounds__(GQ_THREADS, 4)
k_fq_grad(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, int vec, int bits, const float* range_dev,
          double lo, double hi, NoiseArg na) {
    if (range_dev) {
        lo = (double)__ldg(range_dev);
        hi = (double)__ldg(range_dev + 1);
    }
    const GradQ Q = make_gradq(lo, hi, bits);
    sweep_grad<NOISE, WRITE_Y, WRITE_C, false>(g, out, codes, n, vec != 0, Q, na);
}

// fused: phase 1 min/max, grid sync, phase 2 quantize (cooperative launch: grid <= co-resident CTAs)
template <int NOISE, bool WRITE_Y, bool WRITE_C>
__global__ void __launch_bounds__(GQ_THREADS, 4)
k_fq_grad_fused(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, int vec, int bits, float* range_out,
                NoiseArg na, MinMaxWs* ws) {
    __shared__ float sm[64];
    float mn = QB_INF, mx = -QB_INF;
    sweep_minmax(g, n, vec != 0, mn, mx);
    block_reduce_minmax(mn, mx, sm);
    if (threadIdx.x == 0) {
        ws->pmin[blockIdx.x] = mn;
        ws->pmax[blockIdx.x] = mx;
    }
    cg::this_grid().sync();
    mn = QB_INF;
    mx = -QB_INF;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += GQ_THREADS) {
        mn = fminf(mn, __ldcg(&ws->pmin[i]));
        mx = fmaxf(mx, __ldcg(&ws->pmax[i]));
    }
    block_reduce_minmax(mn, mx, sm);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        range_out[0] = mn;
        range_out[1] = mx;
    }
    const GradQ Q = make_gradq((double)mn, (double)mx, bits);
    sweep_grad<NOISE, WRITE_Y, WRITE_C, true>(g, out, codes, n, vec != 0, Q, na);
}

template <typename F>
static int dispatch_grad(int noise_mode, bool wy, bool wc, F&& f) {
#define QB_CASE(NM)                                                        \
    if (noise_mode == NM) {                                                \
        if (wy && wc) return f(std::integral_constant<int, NM>{}, std::true_type{}, std::true_type{});   \
        if (wy) return f(std::integral_constant<int, NM>{}, std::true_type{}, std::false_type{});        \
        return f(std::integral_constant<int, NM>{}, std::false_type{}, std::true_type{});                \
    }
    QB_CASE(NOISE_NONE)
    QB_CASE(NOISE_INJECT)
    QB_CASE(NOISE_PHILOX)
#undef QB_CASE
    return QB_ERR_ARG;
}

static int check_grad_args(const char* who, const float* g, float* out, uint8_t* codes, int64_t n, int bits) {
    QB_CHECK_ARG(n >= 0, "%s: n=%lld", who, (long long)n);
    QB_CHECK_ARG(n == 0 || (g && (out || codes)), "%s: null g or no output", who);
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "%s: bits=%d outside 1..24", who, bits);
    QB_CHECK_ARG(!codes || bits <= 8, "%s: uint8 codes need bits<=8", who);
    return QB_OK;
}

static bool grad_vec_ok(const float* g, const float* out, const uint8_t* codes, const float* noise, uint64_t offset) {
    return aligned32(g) && (!out || aligned32(out)) && (!noise || aligned32(noise)) &&
           (!codes || (reinterpret_cast<uintptr_t>(codes) & 7) == 0) && (offset & 7) == 0;
}

extern "C" int qb_fakequant_grad(const float* g, float* out, uint8_t* codes, int64_t n, int bits, const float* range_dev,
                                 double lo, double hi, const float* noise, int stochastic, uint64_t seed,
                                 uint64_t offset, const uint64_t* step_dev, void* stream) {
    int rc = check_grad_args("qb_fakequant_grad", g, out, codes, n, bits);
    if (rc) return rc;
    if (n == 0) return QB_OK;
    const int nm = !stochastic ? NOISE_NONE : (noise ? NOISE_INJECT : NOISE_PHILOX);
    NoiseArg na{noise, seed, offset, step_dev};
    const int vec = grad_vec_ok(g, out, codes, noise, offset);
    const int grid = gq_grid(vec ? (n >> 3) : n, 8);
    cudaStream_t st = qb_stream(stream);
    rc = dispatch_grad(nm, out != nullptr, codes != nullptr, [&](auto NM, auto WY, auto WC) {
        k_fq_grad<decltype(NM)::value, decltype(WY)::value, decltype(WC)::value>
            <<<grid, GQ_THREADS, 0, st>>>(g, out, codes, n, vec, bits, range_dev, lo, hi, na);
        return (int)QB_OK;
    });
    if (rc) return rc;
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_fakequant_grad_fused(const float* g, float* out, uint8_t* codes, int64_t n, int bits, float* range_out,
                                       const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                                       const uint64_t* step_dev, void* ws, size_t ws_bytes, void* stream) {
    int rc = check_grad_args("qb_fakequant_grad_fused", g, out, codes, n, bits);
    if (rc) return rc;
    QB_CHECK_ARG(n > 0, "qb_fakequant_grad_fused: empty tensor has no range");
    QB_CHECK_ARG(range_out && ws, "qb_fakequant_grad_fused: null range_out/ws");
    if (ws_bytes < sizeof(MinMaxWs)) {
        qb_set_error("qb_fakequant_grad_fused: workspace %zu < %zu bytes", ws_bytes, sizeof(MinMaxWs));
        return QB_ERR_WORKSPACE;
    }
    const int nm = !stochastic ? NOISE_NONE : (noise ? NOISE_INJECT : NOISE_PHILOX);
    NoiseArg na{noise, seed, offset, step_dev};
    int vec = grad_vec_ok(g, out, codes, noise, offset);
    cudaStream_t st = qb_stream(stream);
    MinMaxWs* w = static_cast<MinMaxWs*>(ws);
    return dispatch_grad(nm, out != nullptr, codes != nullptr, [&](auto NM, auto WY, auto WC) -> int {
        auto kern = k_fq_grad_fused<decltype(NM)::value, decltype(WY)::value, decltype(WC)::value>;
        int per_sm = 0;
        QB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, GQ_THREADS, 0));
        if (per_sm < 1) {
            qb_set_error("qb_fakequant_grad_fused: kernel does not fit on an SM");
            return QB_ERR_CUDA;
        }
        int grid = gq_grid(vec ? (n >> 3) : n, per_sm < 8 ? per_sm : 8);
        void* args[] = {(void*)&g, (void*)&out, (void*)&codes, (void*)&n, (void*)&vec, (void*)&bits,
                        (void*)&range_out, (void*)&na, (void*)&w};
        QB_CUDA(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(GQ_THREADS), args, 0, st));
        return (int)QB_OK;
    });
}

// ------------------------------------------------------------------------------------------------
// min-anchored branch WITH stochastic rounding (quantize.py:90-100 with enforce_true_zero=False, stochastic=True): what
// quantize(x, num_chunks=<truthy>) reaches through the reference's positional shuffle (:159-160).  No model of the
// reference calls it, so this is a plain grid-stride kernel: q = rne(clamp((x - lo)/s + u, 0, qmax)), y = q*s + lo.
// ------------------------------------------------------------------------------------------------
template <int NOISE>
__global__ void __launch_bounds__(GQ_THREADS)
k_fq_affine_stochastic(const float* __restrict__ x, float* __restrict__ y, int64_t n, int bits, const float* __restrict__ range_dev,
                       double lo, double hi, int scale_mode, NoiseArg na) {
    if (range_dev) {
        lo = (double)__ldg(range_dev);
        hi = (double)__ldg(range_dev + 1);
    }
    const QParams p = make_qparams(lo, hi, bits, scale_mode);
    const Divisor d = make_divisor(p.scale);
    const uint64_t seed = effective_seed(na);
    for (int64_t e = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x; e < n; e += (int64_t)gridDim.x * GQ_THREADS) {
        float t = qb_div(__fsub_rn(x[e], p.lo), d);
        t = __fadd_rn(t, noise_elem<NOISE>(na, seed, e));
        const float q = qb_rne(fminf(fmaxf(t, 0.0f), p.qmax), p);
        y[e] = q_affine_deq(q, p);
    }
}

extern "C" int qb_fakequant_affine_stochastic(const float* x, float* y, int64_t n, int bits, const float* range_dev, double lo,
                                              double hi, int scale_mode, const float* noise, uint64_t seed, uint64_t offset,
                                              const uint64_t* step_dev, void* stream) {
    QB_CHECK_ARG(n >= 0, "qb_fakequant_affine_stochastic: n=%lld", (long long)n);
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(x && y, "qb_fakequant_affine_stochastic: null pointer");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_fakequant_affine_stochastic: bits=%d outside 1..24", bits);
    QB_CHECK_ARG(scale_mode == QB_SCALE_DOUBLE || scale_mode == QB_SCALE_FLOAT, "qb_fakequant_affine_stochastic: scale_mode=%d", scale_mode);
    NoiseArg na{noise, seed, offset, step_dev};
    const int grid = gq_grid(n, 8);
    if (noise)
        k_fq_affine_stochastic<NOISE_INJECT><<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(x, y, n, bits, range_dev, lo, hi, scale_mode, na);
    else
        k_fq_affine_stochastic<NOISE_PHILOX><<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(x, y, n, bits, range_dev, lo, hi, scale_mode, na);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
