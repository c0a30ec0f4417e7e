// fakequant_grad.cu - gradient fake-quantization (zero-point + stochastic branch) and global min/max.
//
// UniformQuantizeGrad.backward (models/modules/quantize.py:124-139): min/max of the whole gradient, then
// UniformQuantize.forward with enforce_true_zero=True, stochastic=True (:77-88, 92-98).
#include <cooperative_groups.h>

#include <type_traits>

#include "common.cuh"

namespace cg = cooperative_groups;

constexpr int GQ_THREADS = 256;
constexpr int GQ_UNROLL = 4;
constexpr int MM_MAX_BLOCKS = 4096;  // partial slots in the workspace

// workspace layout: float partial_min[MM_MAX_BLOCKS], float partial_max[MM_MAX_BLOCKS], uint32 counter
struct MinMaxWs {
    float pmin[MM_MAX_BLOCKS];
    float pmax[MM_MAX_BLOCKS];
    unsigned int counter;
    unsigned int pad[15];
};

extern "C" size_t qb_minmax_ws_bytes(void) { return sizeof(MinMaxWs); }

enum { NOISE_NONE = 0, NOISE_INJECT = 1, NOISE_PHILOX = 2 };

struct NoiseArg {
    const float* inject;       // NOISE_INJECT
    uint64_t seed, offset;     // NOISE_PHILOX: element e uses lane (e+offset)&3 of philox(seed', (e+offset)>>2)
    const uint64_t* step_dev;  // optional device step counter mixed into the seed (CUDA-graph replays)
};

__device__ __forceinline__ uint64_t effective_seed(const NoiseArg& a) {
    uint64_t s = a.seed;
    if (a.step_dev) s += __ldg(a.step_dev) * 0x9E3779B97F4A7C15ull;
    return s;
}

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

// per-thread min/max over a grid-strided sweep of n floats (vector body when aligned, scalar tail)
template <bool STREAM>
__device__ __forceinline__ void sweep_minmax(const float* __restrict__ x, int64_t n, bool vec, float& mn, float& mx) {
    const int64_t stride = (int64_t)gridDim.x * GQ_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x;
    if (vec) {
        const int64_t nquads = n >> 2;
        int64_t i = tid;
        for (; i + (GQ_UNROLL - 1) * stride < nquads; i += GQ_UNROLL * stride) {
            float4 v[GQ_UNROLL];
#pragma unroll
            for (int u = 0; u < GQ_UNROLL; ++u) {
                const float4* p = reinterpret_cast<const float4*>(x) + (i + u * stride);
                v[u] = STREAM ? __ldcs(p) : __ldg(p);
            }
#pragma unroll
            for (int u = 0; u < GQ_UNROLL; ++u) {
                mn = fminf(fminf(mn, fminf(v[u].x, v[u].y)), fminf(v[u].z, v[u].w));
                mx = fmaxf(fmaxf(mx, fmaxf(v[u].x, v[u].y)), fmaxf(v[u].z, v[u].w));
            }
        }
        for (; i < nquads; i += stride) {
            float4 v = __ldg(reinterpret_cast<const float4*>(x) + i);
            mn = fminf(fminf(mn, fminf(v.x, v.y)), fminf(v.z, v.w));
            mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
        }
        if (tid < (n & 3)) {
            float v = x[(nquads << 2) + tid];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
    } else {
        for (int64_t e = tid; e < n; e += stride) {
            float v = x[e];
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
    sweep_minmax<false>(x, n, vec != 0, mn, mx);
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

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

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
    const int vec = aligned16(x);
    int grid = gq_grid(vec ? (n >> 2) / GQ_UNROLL : n / GQ_UNROLL, 8);
    k_minmax<<<grid, GQ_THREADS, 0, qb_stream(stream)>>>(x, n, vec, range_out, static_cast<MinMaxWs*>(ws));
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// zero-point stochastic quantizer (A2), range given
// ------------------------------------------------------------------------------------------------
template <int NOISE>
__device__ __forceinline__ float4 noise_quad(const NoiseArg& a, uint64_t seed, int64_t quad) {
    if (NOISE == NOISE_INJECT) return __ldcs(reinterpret_cast<const float4*>(a.inject) + quad);
    if (NOISE == NOISE_PHILOX) return philox_noise4(seed, (uint64_t)quad + (a.offset >> 2));
    return make_float4(0.f, 0.f, 0.f, 0.f);
}
template <int NOISE>
__device__ __forceinline__ float noise_elem(const NoiseArg& a, uint64_t seed, int64_t e) {
    if (NOISE == NOISE_INJECT) return a.inject[e];
    if (NOISE == NOISE_PHILOX) {
        uint64_t ge = (uint64_t)e + a.offset;
        float4 r = philox_noise4(seed, ge >> 2);
        int l = (int)(ge & 3);
        return l == 0 ? r.x : l == 1 ? r.y : l == 2 ? r.z : r.w;
    }
    return 0.f;
}

template <int NOISE, bool WRITE_Y, bool WRITE_C>
__device__ __forceinline__ void grad_quad(float4 v, float4 u, float* out, uint8_t* codes, int64_t quad, const QParams& p) {
    float4 q = make_float4(q_zp_code(v.x, u.x, p), q_zp_code(v.y, u.y, p), q_zp_code(v.z, u.z, p), q_zp_code(v.w, u.w, p));
    if (WRITE_Y)
        st_stream4(out + 4 * quad, make_float4(q_zp_deq(q.x, p), q_zp_deq(q.y, p), q_zp_deq(q.z, p), q_zp_deq(q.w, p)));
    if (WRITE_C)
        reinterpret_cast<uchar4*>(codes)[quad] =
            make_uchar4((unsigned char)q.x, (unsigned char)q.y, (unsigned char)q.z, (unsigned char)q.w);
}

// Body shared by the given-range kernel and phase 2 of the fused kernel.  REVERSE sweeps the tensor from
// its end so that the fused kernel's second read starts with the lines its first read left in L2.
template <int NOISE, bool WRITE_Y, bool WRITE_C, bool REVERSE>
__device__ __forceinline__ void sweep_grad(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, bool vec,
                                           const QParams& p, const NoiseArg& na) {
    const uint64_t seed = effective_seed(na);
    const int64_t stride = (int64_t)gridDim.x * GQ_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * GQ_THREADS + threadIdx.x;
    if (vec) {
        // Philox keeps ~40 live registers per quad in flight: 2 quads keep the kernel at 3 CTAs/SM, 4 would halve occupancy
        constexpr int U = (NOISE == NOISE_PHILOX) ? 2 : GQ_UNROLL;
        const int64_t nquads = n >> 2;
        int64_t i = tid;
        for (; i + (U - 1) * stride < nquads; i += U * stride) {
            float4 v[U];
            int64_t qd[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                qd[u] = REVERSE ? nquads - 1 - (i + u * stride) : i + u * stride;
                v[u] = __ldcs(reinterpret_cast<const float4*>(g) + qd[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                grad_quad<NOISE, WRITE_Y, WRITE_C>(v[u], noise_quad<NOISE>(na, seed, qd[u]), out, codes, qd[u], p);
        }
        for (; i < nquads; i += stride) {
            int64_t qd = REVERSE ? nquads - 1 - i : i;
            grad_quad<NOISE, WRITE_Y, WRITE_C>(__ldcs(reinterpret_cast<const float4*>(g) + qd),
                                               noise_quad<NOISE>(na, seed, qd), out, codes, qd, p);
        }
        if (tid < (n & 3)) {
            int64_t e = (nquads << 2) + tid;
            float q = q_zp_code(g[e], noise_elem<NOISE>(na, seed, e), p);
            if (WRITE_Y) out[e] = q_zp_deq(q, p);
            if (WRITE_C) codes[e] = (uint8_t)q;
        }
    } else {
        for (int64_t e = tid; e < n; e += stride) {
            float q = q_zp_code(g[e], noise_elem<NOISE>(na, seed, e), p);
            if (WRITE_Y) out[e] = q_zp_deq(q, p);
            if (WRITE_C) codes[e] = (uint8_t)q;
        }
    }
}

template <int NOISE, bool WRITE_Y, bool WRITE_C>
__global__ void __launch_bounds__(GQ_THREADS, 3)
k_fq_grad(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, int vec, int bits, const float* range_dev,
          double lo, double hi, NoiseArg na) {
    if (range_dev) {
        lo = (double)__ldg(range_dev);
        hi = (double)__ldg(range_dev + 1);
    }
    const QParams p = make_qparams_d(lo, hi, bits);
    sweep_grad<NOISE, WRITE_Y, WRITE_C, false>(g, out, codes, n, vec != 0, p, na);
}

// fused: phase 1 min/max, grid sync, phase 2 quantize (cooperative launch: grid <= co-resident CTAs)
template <int NOISE, bool WRITE_Y, bool WRITE_C>
__global__ void __launch_bounds__(GQ_THREADS, 3)
k_fq_grad_fused(const float* __restrict__ g, float* out, uint8_t* codes, int64_t n, int vec, int bits, float* range_out,
                NoiseArg na, MinMaxWs* ws) {
    __shared__ float sm[64];
    float mn = QB_INF, mx = -QB_INF;
    sweep_minmax<false>(g, n, vec != 0, mn, mx);
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
    const QParams p = make_qparams_d((double)mn, (double)mx, bits);
    sweep_grad<NOISE, WRITE_Y, WRITE_C, true>(g, out, codes, n, vec != 0, p, na);
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
    return aligned16(g) && (!out || aligned16(out)) && (!noise || aligned16(noise)) &&
           (!codes || (reinterpret_cast<uintptr_t>(codes) & 3) == 0) && (offset & 3) == 0;
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
    const int grid = gq_grid(vec ? (n >> 2) : n, 8);
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
        int grid = gq_grid(vec ? (n >> 2) : n, per_sm < 8 ? per_sm : 8);
        void* args[] = {(void*)&g, (void*)&out, (void*)&codes, (void*)&n, (void*)&vec, (void*)&bits,
                        (void*)&range_out, (void*)&na, (void*)&w};
        QB_CUDA(cudaLaunchCooperativeKernel((const void*)kern, dim3(grid), dim3(GQ_THREADS), args, 0, st));
        return (int)QB_OK;
    });
}
