// fakequant.cu - elementwise fake-quantization kernels (HBM-bound: 8 B/elem given range, 12 B/elem fused).
//
// Replaces the clone + 8 in-place ATen passes of UniformQuantize.forward
// (models/modules/quantize.py:66-100) by one read + one write with 128-bit coalesced accesses,
// and UniformQuantizeGrad.backward's two reductions + host syncs + uniform_ + 8 passes (:124-139) by a
// single cooperative launch (min/max -> grid sync -> quantize with in-register Philox noise).
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

constexpr int FQ_THREADS = 256;
constexpr int FQ_U = 2;      // independent 256-bit loads in flight per thread

struct RangeArg {
    const float* dev;  // device {min,max} or nullptr
    double lo, hi;     // host range when dev == nullptr
};

__device__ __forceinline__ QParams load_qparams(const RangeArg& r, int bits, int mode) {
    double lo = r.lo, hi = r.hi;
    if (r.dev) {
        lo = (double)__ldg(r.dev);
        hi = (double)__ldg(r.dev + 1);
    }
    return make_qparams(lo, hi, bits, mode);
}

// ------------------------------------------------------------------------------------------------
// min-anchored affine (A1): thread = one 32-byte octet per step, FQ_U octets in flight
// ------------------------------------------------------------------------------------------------
// returns the sum of the octet's 8 integer codes (dead code unless the caller uses it: k_fq_affine_rowsum)
template <bool WRITE_Y, bool WRITE_C, bool SMALL, bool WRITE_H = false>
__device__ __forceinline__ uint32_t affine_octet(const F8& v, float* y, uint8_t* codes, int64_t oct, const QParams& p,
                                                 const Divisor& d, uint16_t* codes16 = nullptr) {
    float a[8];
    bool ok = d.fast;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        a[j] = __fsub_rn(v.v[j], p.lo);
        ok = ok && qb_div_ok(a[j]);
    }
    float t[8];
    if (ok) {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = qb_div_fast(a[j], d);
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = __fdiv_rn(a[j], d.b);
    }
    F8 o;
    uint32_t c[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float s = fminf(fmaxf(t[j], 0.0f), p.qmax);
        float q;
        if (SMALL) {
            const float m = __fadd_rn(s, QB_MAGIC);
            c[j] = __float_as_uint(m);
            q = __fsub_rn(m, QB_MAGIC);
        } else {
            q = rintf(s);
            c[j] = (uint32_t)q;
        }
        o.v[j] = q_affine_deq(q, p);
    }
    if (WRITE_Y) st256_stream(y + 8 * oct, o);
    if (WRITE_C) {
        uint2 pk;
        pk.x = __byte_perm(__byte_perm(c[0], c[1], 0x0040), __byte_perm(c[2], c[3], 0x0040), 0x5410);
        pk.y = __byte_perm(__byte_perm(c[4], c[5], 0x0040), __byte_perm(c[6], c[7], 0x0040), 0x5410);
        reinterpret_cast<uint2*>(codes)[oct] = pk;
    }
    if (WRITE_H) {   // the codes as bf16 (exact for bits <= 8): the upper half of (float)code
        uint32_t f[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = __float_as_uint(SMALL ? __fsub_rn(__uint_as_float(c[j]), QB_MAGIC) : (float)c[j]);
        uint4 o;
        o.x = __byte_perm(f[0], f[1], 0x7632);
        o.y = __byte_perm(f[2], f[3], 0x7632);
        o.z = __byte_perm(f[4], f[5], 0x7632);
        o.w = __byte_perm(f[6], f[7], 0x7632);
        reinterpret_cast<uint4*>(codes16)[oct] = o;
    }
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) sum += SMALL ? (c[j] & 0xffffu) : c[j];      // SMALL: the code sits in the low mantissa bits
    return sum;
}

template <bool WRITE_Y, bool WRITE_C, bool SMALL, bool WRITE_H = false>
__device__ __forceinline__ void affine_sweep(const float* __restrict__ x, float* y, uint8_t* codes, int64_t n,
                                             const QParams& p, const Divisor& d, uint16_t* codes16 = nullptr) {
    const int64_t noct = n >> 3;
    const int64_t stride = (int64_t)gridDim.x * FQ_THREADS;
    int64_t i = (int64_t)blockIdx.x * FQ_THREADS + threadIdx.x;
    for (; i + (FQ_U - 1) * stride < noct; i += FQ_U * stride) {
        F8 v[FQ_U];
#pragma unroll
        for (int u = 0; u < FQ_U; ++u) v[u] = ld256_stream(x + 8 * (i + u * stride));
#pragma unroll
        for (int u = 0; u < FQ_U; ++u) affine_octet<WRITE_Y, WRITE_C, SMALL, WRITE_H>(v[u], y, codes, i + u * stride, p, d, codes16);
    }
    for (; i < noct; i += stride) affine_octet<WRITE_Y, WRITE_C, SMALL, WRITE_H>(ld256_stream(x + 8 * i), y, codes, i, p, d, codes16);
    // n % 8 tail
    if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
        const int64_t e = (noct << 3) + threadIdx.x;
        const float q = q_affine_code(x[e], p, d);
        if (WRITE_Y) y[e] = q_affine_deq(q, p);
        if (WRITE_C) codes[e] = (uint8_t)q;
        if (WRITE_H) codes16[e] = (uint16_t)(__float_as_uint(q) >> 16);
    }
}

template <bool WRITE_Y, bool WRITE_C, bool WRITE_H = false>
__global__ void __launch_bounds__(FQ_THREADS, 4)
k_fq_affine_v8(const float* __restrict__ x, float* y, uint8_t* codes, int64_t n, int bits, RangeArg r, int mode,
               uint16_t* codes16 = nullptr) {
    const QParams p = load_qparams(r, bits, mode);
    const Divisor d = make_divisor(p.scale);
    if (p.qmax < 4194304.0f)
        affine_sweep<WRITE_Y, WRITE_C, true, WRITE_H>(x, y, codes, n, p, d, codes16);
    else
        affine_sweep<WRITE_Y, WRITE_C, false, WRITE_H>(x, y, codes, n, p, d, codes16);
}

// The activation quantizer of a tensor-core contraction with the per-pixel (per-row) code sums the zero-point correction of
// the int8 kernels needs (qb_rowsum_u8) taken from the registers that hold the codes: a row of `cols` = 8 * lpp codes sits in
// lpp adjacent lanes (lpp = 2 .. 32, a power of two; the octet index of lane 0 is a multiple of 32), so a xor-shuffle tree
// over lpp lanes yields the row sum - no second pass over the codes, one launch less per layer.  Loop bounds are warp-uniform.
template <bool WRITE_H>
__global__ void __launch_bounds__(FQ_THREADS, 4)
k_fq_affine_rowsum(const float* __restrict__ x, uint8_t* codes, uint16_t* codes16, int64_t n, int bits, RangeArg r, int mode, int lpp,
                   int lpp_log2, int32_t* __restrict__ rowsum) {
    const QParams p = load_qparams(r, bits, mode);
    const Divisor d = make_divisor(p.scale);
    const int64_t noct = n >> 3, stride = (int64_t)gridDim.x * FQ_THREADS;
    const int lane = threadIdx.x & 31;
    for (int64_t iw = (int64_t)blockIdx.x * FQ_THREADS + (threadIdx.x - lane); iw < noct; iw += FQ_U * stride) {
        F8 v[FQ_U];
        bool ok[FQ_U];
#pragma unroll
        for (int u = 0; u < FQ_U; ++u) {
            const int64_t i = iw + u * stride + lane;
            ok[u] = i < noct;
            if (ok[u]) v[u] = ld256_stream(x + 8 * i);
        }
#pragma unroll
        for (int u = 0; u < FQ_U; ++u) {
            if (iw + u * stride >= noct) break;                                  // warp-uniform
            const int64_t i = iw + u * stride + lane;
            uint32_t sum = 0;
            if (ok[u]) sum = affine_octet<false, true, true, WRITE_H>(v[u], nullptr, codes, i, p, d, codes16);
            for (int o = 1; o < lpp; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (ok[u] && (lane & (lpp - 1)) == 0) rowsum[i >> lpp_log2] = (int32_t)sum;
        }
    }
}

// scalar path for pointers that are not 32-byte aligned (views into larger tensors)
__global__ void __launch_bounds__(FQ_THREADS)
k_fq_affine_scalar(const float* __restrict__ x, float* y, uint8_t* codes, int64_t n, int bits, RangeArg r, int mode,
                   uint16_t* codes16 = nullptr) {
    const QParams p = load_qparams(r, bits, mode);
    for (int64_t e = (int64_t)blockIdx.x * FQ_THREADS + threadIdx.x; e < n; e += (int64_t)gridDim.x * FQ_THREADS) {
        float q = q_affine_code(x[e], p);
        if (y) y[e] = q_affine_deq(q, p);
        if (codes) codes[e] = (uint8_t)q;
        if (codes16) codes16[e] = (uint16_t)(__float_as_uint(q) >> 16);
    }
}

static int fq_grid(int64_t work_items, int per_sm) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = (work_items + FQ_THREADS - 1) / FQ_THREADS;
    int64_t cap = (int64_t)sms * per_sm;
    if (want < 1) want = 1;
    return (int)(want < cap ? want : cap);
}

static bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

extern "C" int qb_fakequant_affine(const float* x, float* y, uint8_t* codes, int64_t n, int bits,
                                   const float* range_dev, double lo, double hi, int scale_mode, void* stream) {
    QB_CHECK_ARG(n >= 0, "qb_fakequant_affine: n=%lld", (long long)n);
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(x && (y || codes), "qb_fakequant_affine: null x or no output");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_fakequant_affine: bits=%d outside 1..24", bits);
    QB_CHECK_ARG(!codes || bits <= 8, "qb_fakequant_affine: uint8 codes need bits<=8 (bits=%d)", bits);
    QB_CHECK_ARG(scale_mode == QB_SCALE_DOUBLE || scale_mode == QB_SCALE_FLOAT, "qb_fakequant_affine: scale_mode=%d",
                 scale_mode);
    RangeArg r{range_dev, lo, hi};
    cudaStream_t st = qb_stream(stream);
    const bool vec = aligned(x, 32) && (!y || aligned(y, 32)) && (!codes || aligned(codes, 8));
    if (vec) {
        int grid = fq_grid(n >> 3, 4);
        if (y && codes)
            k_fq_affine_v8<true, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode);
        else if (y)
            k_fq_affine_v8<true, false><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode);
        else
            k_fq_affine_v8<false, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode);
    } else {
        k_fq_affine_scalar<<<fq_grid(n, 8), FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode);
    }
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// First layer of a network: the image batch arrives NCHW (torch's default), the convolution kernels read NHWC codes.  The
// quantizer itself does the transposition - 3 coalesced plane reads, one packed 12-byte write per 4 pixels - instead of a
// separate NCHW -> NHWC copy of the fp32 batch (154 MB read + written per ResNet-18 step).  Same arithmetic as
// k_fq_affine_v8 (device range, QB_SCALE_DOUBLE); codes only.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FQ_THREADS)
k_fq_affine_nchw3(const float* __restrict__ x, uint8_t* __restrict__ codes, int64_t B, int64_t HW, int bits, const float* range_dev) {
    const AffineQ aq = load_affineq(range_dev, bits);
    const int64_t quads = HW >> 2, total = B * quads;
    for (int64_t i = (int64_t)blockIdx.x * FQ_THREADS + threadIdx.x; i < total; i += (int64_t)gridDim.x * FQ_THREADS) {
        const int64_t b = i / quads, q = i - b * quads;
        const float* src = x + b * 3 * HW + 4 * q;
        float v[12];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float4 f = __ldcs(reinterpret_cast<const float4*>(src + c * HW));
            v[c] = f.x; v[3 + c] = f.y; v[6 + c] = f.z; v[9 + c] = f.w;          // pixel-major: v[3*pixel + channel]
        }
        uint32_t cd[12];
        affine_fq_vec<12>(v, aq, cd);
        uint32_t* dst = reinterpret_cast<uint32_t*>(codes + (b * HW + 4 * q) * 3);
        dst[0] = cd[0] | (cd[1] << 8) | (cd[2] << 16) | (cd[3] << 24);
        dst[1] = cd[4] | (cd[5] << 8) | (cd[6] << 16) | (cd[7] << 24);
        dst[2] = cd[8] | (cd[9] << 8) | (cd[10] << 16) | (cd[11] << 24);
    }
}

/* uint8 codes [B, H*W, 3] (channels-last) of the min-anchored quantizer applied to an fp32 NCHW batch [B, 3, H*W]: the
 * activation quantizer of a network's first QConv2d (quantize.py:339-343 on models/resnet_quantized.py:147 /
 * mobilenet_quantized.py:66) fused with the layout change the NHWC kernels need.  H*W % 4 == 0, 16-byte aligned x, bits <= 8. */
extern "C" int qb_fakequant_affine_nchw3(const float* x, uint8_t* codes, int64_t B, int64_t HW, int bits, const float* range_dev,
                                         void* stream) {
    QB_CHECK_ARG(x && codes && range_dev, "qb_fakequant_affine_nchw3: null pointer");
    QB_CHECK_ARG(B > 0 && HW > 0 && HW % 4 == 0, "qb_fakequant_affine_nchw3: B=%lld H*W=%lld (H*W %% 4 == 0 required)", (long long)B,
                 (long long)HW);
    QB_CHECK_ARG(bits >= 1 && bits <= 8, "qb_fakequant_affine_nchw3: bits=%d", bits);
    QB_CHECK_ARG(aligned(x, 16) && aligned(codes, 4), "qb_fakequant_affine_nchw3: misaligned tensor");
    k_fq_affine_nchw3<<<fq_grid(B * (HW >> 2), 8), FQ_THREADS, 0, qb_stream(stream)>>>(x, codes, B, HW, bits, range_dev);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* As qb_fakequant_affine, plus codes16: the integer codes as bf16 (bits <= 8; exact) - the x operand of qb_conv_wgrad. */
extern "C" int qb_fakequant_affine16(const float* x, float* y, uint8_t* codes, uint16_t* codes16, int64_t n, int bits,
                                     const float* range_dev, double lo, double hi, int scale_mode, void* stream) {
    if (!codes16) return qb_fakequant_affine(x, y, codes, n, bits, range_dev, lo, hi, scale_mode, stream);
    QB_CHECK_ARG(n >= 0, "qb_fakequant_affine16: n=%lld", (long long)n);
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(x, "qb_fakequant_affine16: null x");
    QB_CHECK_ARG(bits >= 1 && bits <= 8, "qb_fakequant_affine16: bf16 codes need bits<=8 (bits=%d)", bits);
    QB_CHECK_ARG(scale_mode == QB_SCALE_DOUBLE || scale_mode == QB_SCALE_FLOAT, "qb_fakequant_affine16: scale_mode=%d", scale_mode);
    RangeArg r{range_dev, lo, hi};
    cudaStream_t st = qb_stream(stream);
    const bool vec = aligned(x, 32) && (!y || aligned(y, 32)) && (!codes || aligned(codes, 8)) && aligned(codes16, 16);
    if (vec) {
        int grid = fq_grid(n >> 3, 4);
        if (y && codes) k_fq_affine_v8<true, true, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode, codes16);
        else if (y) k_fq_affine_v8<true, false, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode, codes16);
        else if (codes) k_fq_affine_v8<false, true, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode, codes16);
        else k_fq_affine_v8<false, false, true><<<grid, FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode, codes16);
    } else {
        k_fq_affine_scalar<<<fq_grid(n, 8), FQ_THREADS, 0, st>>>(x, y, codes, n, bits, r, scale_mode, codes16);
    }
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_fakequant_affine_rowsum_supported(int64_t cols, int bits) {
    return bits >= 1 && bits <= 8 && cols >= 16 && cols <= 256 && (cols & (cols - 1)) == 0;
}

/* Codes (uint8, and optionally as bf16 in codes16) of the min-anchored quantizer over x [rows, cols] with a device range, plus
 * rowsum[r] = SUM_c codes[r][c] - qb_fakequant_affine16(y = NULL) + qb_rowsum_u8 in one pass.  cols a power of two in 16..256. */
extern "C" int qb_fakequant_affine_rowsum(const float* x, uint8_t* codes, uint16_t* codes16, int64_t rows, int64_t cols, int bits,
                                          const float* range_dev, int32_t* rowsum, void* stream) {
    QB_CHECK_ARG(x && codes && range_dev && rowsum, "qb_fakequant_affine_rowsum: null pointer");
    QB_CHECK_ARG(rows > 0 && qb_fakequant_affine_rowsum_supported(cols, bits),
                 "qb_fakequant_affine_rowsum: rows=%lld cols=%lld bits=%d (cols a power of two in 16..256, bits <= 8)", (long long)rows,
                 (long long)cols, bits);
    QB_CHECK_ARG(aligned(x, 32) && aligned(codes, 8) && (!codes16 || aligned(codes16, 16)), "qb_fakequant_affine_rowsum: misaligned tensor");
    const int lpp = (int)(cols >> 3);
    int lg = 0;
    while ((1 << lg) < lpp) ++lg;
    const int64_t n = rows * cols;
    RangeArg r{range_dev, 0.0, 0.0};
    const int grid = fq_grid(n >> 3, 4);
    cudaStream_t st = qb_stream(stream);
    if (codes16) k_fq_affine_rowsum<true><<<grid, FQ_THREADS, 0, st>>>(x, codes, codes16, n, bits, r, QB_SCALE_DOUBLE, lpp, lg, rowsum);
    else k_fq_affine_rowsum<false><<<grid, FQ_THREADS, 0, st>>>(x, codes, codes16, n, bits, r, QB_SCALE_DOUBLE, lpp, lg, rowsum);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// per-row (per-output-channel) weights: one CTA per row, row staged in registers/shared memory
// quantize.py:318-323 (row min/max over flatten(1)) + :332-334 (tensor-range quantize)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float2 block_minmax(float mn, float mx, float* sm /* 2*32 floats */) {
    mn = warp_min(mn);
    mx = warp_max(mx);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
    __syncthreads();  // protect sm reuse
    if (l == 0) {
        sm[w] = mn;
        sm[32 + w] = mx;
    }
    __syncthreads();
    mn = l < nw ? sm[l] : QB_INF;
    mx = l < nw ? sm[32 + l] : -QB_INF;
    mn = warp_min(mn);
    mx = warp_max(mx);
    return make_float2(mn, mx);
}

constexpr int ROW_THREADS = 256;

__global__ void __launch_bounds__(ROW_THREADS)
k_fq_rows(const float* __restrict__ w, float* wq, uint8_t* codes, int64_t cols, int bits, float* row_min,
          float* row_max, int compute_range) {
    __shared__ float sm[64];
    const int64_t row = blockIdx.x;
    const float* src = w + row * cols;
    float lo, hi;
    if (compute_range) {
        float mn = QB_INF, mx = -QB_INF;
        for (int64_t j = threadIdx.x; j < cols; j += ROW_THREADS) {
            float v = src[j];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
        float2 r = block_minmax(mn, mx, sm);
        lo = r.x;
        hi = r.y;
        if (threadIdx.x == 0) {
            row_min[row] = lo;
            row_max[row] = hi;
        }
    } else {
        lo = row_min[row];
        hi = row_max[row];
    }
    const QParams p = make_qparams_f(lo, hi, bits);
    for (int64_t j = threadIdx.x; j < cols; j += ROW_THREADS) {   // second read hits L1/L2 (rows <= 18 KB)
        float q = q_affine_code(src[j], p);
        if (wq) wq[row * cols + j] = q_affine_deq(q, p);
        if (codes) codes[row * cols + j] = (uint8_t)q;
    }
}

// Long rows (cols >= 8192: per-channel ranges of activation-sized tensors, BASELINE configs[4]): one CTA per row leaves most
// SMs idle (64 rows x 65 536: 558 GB/s).  Here a row is split over a thread-block CLUSTER of FQ_ROW_CL CTAs: each CTA reduces
// its segment, the CL partial extrema meet through distributed shared memory, then each CTA quantizes its own segment, which
// is still in L1/L2.  Same arithmetic as k_fq_rows (min / max are order independent), 128-bit accesses when the row pitch allows.
constexpr int FQ_ROW_CL = 8;

__global__ void __cluster_dims__(FQ_ROW_CL, 1, 1) __launch_bounds__(ROW_THREADS)
k_fq_rows_cluster(const float* __restrict__ w, float* wq, uint8_t* codes, int64_t cols, int bits, float* row_min, float* row_max,
                  int compute_range, int vec4) {
    __shared__ float sm[64];
    __shared__ float s_part[2];
    const int64_t row = blockIdx.x / FQ_ROW_CL;
    uint32_t part;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(part));
    int64_t seg = (cols + FQ_ROW_CL - 1) / FQ_ROW_CL;
    seg = (seg + 3) & ~(int64_t)3;
    const int64_t j0 = (int64_t)part * seg < cols ? (int64_t)part * seg : cols;
    const int64_t j1 = j0 + seg < cols ? j0 + seg : cols;
    const float* src = w + row * cols;
    float lo, hi;
    if (compute_range) {
        float mn = QB_INF, mx = -QB_INF;
        if (vec4) {
            for (int64_t j = j0 + 4 * (int64_t)threadIdx.x; j + 3 < j1; j += 4 * ROW_THREADS) {
                const float4 v = *reinterpret_cast<const float4*>(src + j);
                mn = fminf(fminf(mn, fminf(v.x, v.y)), fminf(v.z, v.w));
                mx = fmaxf(fmaxf(mx, fmaxf(v.x, v.y)), fmaxf(v.z, v.w));
            }
            for (int64_t j = j0 + ((j1 - j0) & ~(int64_t)3) + threadIdx.x; j < j1; j += ROW_THREADS) {
                mn = fminf(mn, src[j]);
                mx = fmaxf(mx, src[j]);
            }
        } else {
            for (int64_t j = j0 + threadIdx.x; j < j1; j += ROW_THREADS) {
                mn = fminf(mn, src[j]);
                mx = fmaxf(mx, src[j]);
            }
        }
        const float2 r = block_minmax(mn, mx, sm);
        if (threadIdx.x == 0) { s_part[0] = r.x; s_part[1] = r.y; }
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        lo = QB_INF; hi = -QB_INF;
        const uint32_t local = (uint32_t)__cvta_generic_to_shared(s_part);
#pragma unroll
        for (uint32_t c = 0; c < FQ_ROW_CL; ++c) {
            uint32_t remote;
            asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(c));
            float a, b;
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(a) : "r"(remote));
            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(b) : "r"(remote + 4));
            lo = fminf(lo, a);
            hi = fmaxf(hi, b);
        }
        // nobody leaves (and frees its shared memory) before every CTA of the cluster has read the partial extrema
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (part == 0 && threadIdx.x == 0) {
            row_min[row] = lo;
            row_max[row] = hi;
        }
    } else {
        lo = row_min[row];
        hi = row_max[row];
    }
    const QParams p = make_qparams_f(lo, hi, bits);
    const Divisor d = make_divisor(p.scale);
    if (vec4) {
        for (int64_t j = j0 + 4 * (int64_t)threadIdx.x; j + 3 < j1; j += 4 * ROW_THREADS) {
            const float4 v = *reinterpret_cast<const float4*>(src + j);
            const float q0 = q_affine_code(v.x, p, d), q1 = q_affine_code(v.y, p, d), q2 = q_affine_code(v.z, p, d),
                        q3 = q_affine_code(v.w, p, d);
            if (wq)
                *reinterpret_cast<float4*>(wq + row * cols + j) =
                    make_float4(q_affine_deq(q0, p), q_affine_deq(q1, p), q_affine_deq(q2, p), q_affine_deq(q3, p));
            if (codes)
                *reinterpret_cast<uint32_t*>(codes + row * cols + j) =
                    (uint32_t)q0 | ((uint32_t)q1 << 8) | ((uint32_t)q2 << 16) | ((uint32_t)q3 << 24);
        }
        for (int64_t j = j0 + ((j1 - j0) & ~(int64_t)3) + threadIdx.x; j < j1; j += ROW_THREADS) {
            const float q = q_affine_code(src[j], p, d);
            if (wq) wq[row * cols + j] = q_affine_deq(q, p);
            if (codes) codes[row * cols + j] = (uint8_t)q;
        }
    } else {
        for (int64_t j = j0 + threadIdx.x; j < j1; j += ROW_THREADS) {
            const float q = q_affine_code(src[j], p, d);
            if (wq) wq[row * cols + j] = q_affine_deq(q, p);
            if (codes) codes[row * cols + j] = (uint8_t)q;
        }
    }
}

extern "C" int qb_fakequant_rows(const float* w, float* wq, uint8_t* codes, int64_t rows, int64_t cols, int bits,
                                 float* row_min, float* row_max, int compute_range, void* stream) {
    QB_CHECK_ARG(rows >= 0 && cols >= 0, "qb_fakequant_rows: rows=%lld cols=%lld", (long long)rows, (long long)cols);
    if (rows == 0 || cols == 0) return QB_OK;
    QB_CHECK_ARG(w && (wq || codes) && row_min && row_max, "qb_fakequant_rows: null pointer");
    QB_CHECK_ARG(bits >= 1 && bits <= 24, "qb_fakequant_rows: bits=%d outside 1..24", bits);
    QB_CHECK_ARG(!codes || bits <= 8, "qb_fakequant_rows: uint8 codes need bits<=8");
    QB_CHECK_ARG(rows <= 2147483647LL, "qb_fakequant_rows: too many rows");
    if (cols >= 8192 && rows * FQ_ROW_CL <= 2147483647LL) {
        const int vec4 = cols % 4 == 0 && aligned(w, 16) && (!wq || aligned(wq, 16)) && (!codes || aligned(codes, 4));
        k_fq_rows_cluster<<<(unsigned)(rows * FQ_ROW_CL), ROW_THREADS, 0, qb_stream(stream)>>>(w, wq, codes, cols, bits, row_min,
                                                                                            row_max, compute_range, vec4);
        QB_LAUNCH_CHECK();
        return QB_OK;
    }
    k_fq_rows<<<(unsigned)rows, ROW_THREADS, 0, qb_stream(stream)>>>(w, wq, codes, cols, bits, row_min, row_max,
                                                                    compute_range);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// self-test hook: qb_div (the FMA sequence every quantizer uses) next to __fdiv_rn on the same operands
// ------------------------------------------------------------------------------------------------
__global__ void k_selftest_div(const float* __restrict__ a, float b, float* out_fast, float* out_ieee, int64_t n) {
    const Divisor d = make_divisor(b);
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
        out_fast[e] = qb_div(a[e], d);
        out_ieee[e] = __fdiv_rn(a[e], b);
    }
}

extern "C" int qb_selftest_div(const float* a, float b, float* out_fast, float* out_ieee, int64_t n, void* stream) {
    QB_CHECK_ARG(a && out_fast && out_ieee && n > 0, "qb_selftest_div: null pointer or n<=0");
    k_selftest_div<<<fq_grid(n, 8), FQ_THREADS, 0, qb_stream(stream)>>>(a, b, out_fast, out_ieee, n);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
