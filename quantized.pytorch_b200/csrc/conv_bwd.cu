// conv_bwd.cu - the dense BACKWARD contractions of QConv2d / QLinear on tcgen05 (kind::f16, bf16 operands, fp32 TMEM
// accumulators), fed by TMA from integer codes.
//
// Reference: conv2d_biprec / linear_biprec (models/modules/quantize.py:142-156) and the non-biprec branch (:342-346):
//     dX = conv_bwd_data  (Q(g), w_hat)          Q(g) = (q_g - zp) * s_g, q_g in [0, 2^b - 1]      (quantized branch)
//     dW = conv_bwd_filter(g,    x_hat)          x_hat = q_a * s_a + l_a inside the image, 0 in the padding
// Neither is an integer GEMM (dX: the per-output-channel weight scale s_w[n] sits INSIDE the contraction over n; dW: g is
// the unquantized fp32 gradient), but one operand of each IS a small integer and therefore exact in bf16:
//     dX[m,c] = s_g * SUM_{tap,n} (q_g - zp)[m',n] * w_hat[n,c,tap]       (q_g - zp) exact, w_hat = hi + lo (two bf16 planes)
//     dW[n,c,tap] = s_a * SUM_m g[m,n] * q_a[m',c] + l_a * SUM_{m: tap in bounds} g[m,n]     q_a exact, g = hi + lo
// so each is TWO bf16 MMAs per K step with a 2^-17 relative operand error (TF32 would be one MMA at half rate with 2^-11).
// Operands are 2 bytes (codes) or 2 x 2 bytes (splits) per element instead of the 4-byte fp32 tensors cuDNN reads.
//
//   qb_bwd_pack    one pass over g: gradient quantizer (zero-point, stochastic; quantize.py:124-139) -> bf16 (q_g - zp),
//                  and the bf16 hi / lo split of the raw g
//   qb_wt_pack     w_hat fp32 [Cout][Cin][taps] -> bf16 hi / lo planes [Cin][taps][Cout] (K-major rows for the dgrad B operand)
//   qb_conv_dgrad  implicit GEMM over output-pixel tiles of dX, same skeleton as conv_i8.cu (4-D TMA boxes of the NHWC
//                  operand at the tap-shifted coordinate, zero fill = the padding); stride-s convolutions are s*s parity
//                  classes of dX, each a unit-stride problem over a subset of the taps writing a strided view of dX
//   qb_conv_wgrad  D[(tap,c)][n] = SUM_pixels X^T G with BOTH operands MN-major (pixels = K are the slow index of an NHWC
//                  tensor): TMA boxes {channels, pixels} land as [pixel][channel] rows, which IS the canonical MN-major
//                  swizzled layout; split-K over CTAs, accumulators stay in TMEM for the whole pixel slice, deterministic
//                  reduction of the per-slice partials
#include <cuda_bf16.h>
#include <stdlib.h>

#include "tc_common.cuh"

// ------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t bf16x2_rn(float a, float b) {        // {lo half = a, hi half = b}
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
__device__ __forceinline__ float bf16_lo_f(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16_hi_f(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }

// g = hi + lo (+ O(2^-17 |g|)):  hi = bf16(g), lo = bf16(g - hi)   (g - hi is exact in fp32)
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = bf16x2_rn(a, b);
    lo = bf16x2_rn(__fsub_rn(a, bf16_lo_f(hi)), __fsub_rn(b, bf16_hi_f(hi)));
}

// smem descriptor, K-major rows of `rb` bytes (rb-byte swizzle): 8-row groups are 8*rb bytes apart
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t addr, int rb) {
    const uint64_t layout = rb == 128 ? 2 : (rb == 64 ? 4 : 6);
    return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)((8 * rb) >> 4) << 32) | ((uint64_t)1 << 46) |
           (layout << 61);
}
// smem descriptor, MN-major: rows = K index, each row holds `rb` bytes of consecutive M (or N) elements, rb-byte swizzle.
// Canonical form ((T,rb/16,m),(8,k)) : ((1,T,LBO),(rb/16*T,SBO)): `lbo` = bytes between blocks of rb/2 elements along M/N,
// SBO = bytes between groups of 8 K rows (dense: 8*rb).
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t addr, int rb, uint32_t lbo) {
    const uint64_t layout = rb == 128 ? 2 : (rb == 64 ? 4 : 6);
    return (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((8 * rb) >> 4) << 32) |
           ((uint64_t)1 << 46) | (layout << 61);
}
// kind::f16 instruction descriptor: fp32 accumulate, bf16 x bf16, optional MN-major operands
__host__ __device__ constexpr uint32_t idesc_bf16(uint32_t M, uint32_t N, uint32_t a_mn, uint32_t b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// ------------------------------------------------------------------------------------------------
// qb_bwd_pack: gradient quantizer -> bf16 integer codes, and / or bf16 hi-lo split of the raw tensor
// ------------------------------------------------------------------------------------------------
constexpr int BP_THREADS = 256;

template <int NOISE, bool WQ, bool WS>
__global__ void __launch_bounds__(BP_THREADS, 4)
k_bwd_pack(const float* __restrict__ g, uint16_t* __restrict__ gq, uint16_t* __restrict__ ghi, uint16_t* __restrict__ glo, int64_t n,
           int vec, int bits, const float* __restrict__ range_dev, NoiseArg na) {
    GradQ Q;
    if (WQ) Q = make_gradq((double)__ldg(range_dev), (double)__ldg(range_dev + 1), bits);
    const uint64_t seed = effective_seed(na);
    const int64_t stride = (int64_t)gridDim.x * BP_THREADS;
    const int64_t tid = (int64_t)blockIdx.x * BP_THREADS + threadIdx.x;
    const int64_t noct = vec ? (n >> 3) : 0;
    for (int64_t i = tid; i < noct; i += stride) {
        const F8 v = ld256_stream(g + 8 * i);
        if (WS) {
            uint4 h, l;
            split2(v.v[0], v.v[1], h.x, l.x);
            split2(v.v[2], v.v[3], h.y, l.y);
            split2(v.v[4], v.v[5], h.z, l.z);
            split2(v.v[6], v.v[7], h.w, l.w);
            reinterpret_cast<uint4*>(ghi)[i] = h;
            reinterpret_cast<uint4*>(glo)[i] = l;
        }
        if (WQ) {
            float u[8];
            if (NOISE == NOISE_INJECT) {
                const F8 w = ld256_stream(na.inject + 8 * i);
#pragma unroll
                for (int j = 0; j < 8; ++j) u[j] = w.v[j];
            } else if (NOISE == NOISE_PHILOX) {
                philox_noise8(seed, (uint64_t)i + (na.offset >> 3), u);
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
            uint32_t c[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float s = __fadd_rn(t[j], Q.p.zp);
                if (NOISE != NOISE_NONE) s = __fadd_rn(s, u[j]);
                s = fminf(fmaxf(s, 0.0f), Q.p.qmax);
                const float q = __fsub_rn(__fadd_rn(s, QB_MAGIC), QB_MAGIC);        // RNE, qmax <= 255 here
                c[j] = __float_as_uint(__fsub_rn(q, Q.p.zp));                       // |q - zp| <= 255: exact in bf16
            }
            uint4 o;
            o.x = __byte_perm(c[0], c[1], 0x7632);
            o.y = __byte_perm(c[2], c[3], 0x7632);
            o.z = __byte_perm(c[4], c[5], 0x7632);
            o.w = __byte_perm(c[6], c[7], 0x7632);
            reinterpret_cast<uint4*>(gq)[i] = o;
        }
    }
    // scalar tail (or everything when the pointers are not aligned)
    for (int64_t e = (noct << 3) + tid; e < n; e += stride) {
        const float v = g[e];
        if (WS) {
            uint32_t h, l;
            split2(v, 0.f, h, l);
            ghi[e] = (uint16_t)h;
            glo[e] = (uint16_t)l;
        }
        if (WQ) {
            float t = __fadd_rn(qb_div(v, Q.d), Q.p.zp);
            if (NOISE != NOISE_NONE) t = __fadd_rn(t, noise_elem<NOISE>(na, seed, e));
            const float q = qb_rne(fminf(fmaxf(t, 0.0f), Q.p.qmax), Q.p);
            gq[e] = (uint16_t)(__float_as_uint(__fsub_rn(q, Q.p.zp)) >> 16);
        }
    }
}

extern "C" int qb_bwd_pack(const float* g, int64_t n, uint16_t* gq16, uint16_t* g_hi, uint16_t* g_lo, int bits,
                           const float* range_dev, const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                           const uint64_t* step_dev, void* stream) {
    QB_CHECK_ARG(n >= 0, "qb_bwd_pack: n=%lld", (long long)n);
    if (n == 0) return QB_OK;
    QB_CHECK_ARG(g && (gq16 || (g_hi && g_lo)), "qb_bwd_pack: null g or no output");
    QB_CHECK_ARG((g_hi == nullptr) == (g_lo == nullptr), "qb_bwd_pack: g_hi and g_lo come together");
    QB_CHECK_ARG(!gq16 || (range_dev && bits >= 1 && bits <= 8), "qb_bwd_pack: bf16 codes need a device range and bits <= 8");
    const int nm = !gq16 || !stochastic ? NOISE_NONE : (noise ? NOISE_INJECT : NOISE_PHILOX);
    NoiseArg na{noise, seed, offset, step_dev};
    auto al = [](const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; };
    const int vec = al(g, 32) && (!gq16 || al(gq16, 16)) && (!g_hi || (al(g_hi, 16) && al(g_lo, 16))) && (!noise || al(noise, 32)) &&
                    (offset & 7) == 0;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int64_t want = ((vec ? (n >> 3) : n) + BP_THREADS - 1) / BP_THREADS;
    if (want < 1) want = 1;
    const int grid = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    cudaStream_t st = qb_stream(stream);
#define QB_PACK(NM, WQ, WS) k_bwd_pack<NM, WQ, WS><<<grid, BP_THREADS, 0, st>>>(g, gq16, g_hi, g_lo, n, vec, bits, range_dev, na)
    const bool wq = gq16 != nullptr, wsplit = g_hi != nullptr;
    if (!wq) QB_PACK(NOISE_NONE, false, true);
    else if (nm == NOISE_NONE) { if (wsplit) QB_PACK(NOISE_NONE, true, true); else QB_PACK(NOISE_NONE, true, false); }
    else if (nm == NOISE_INJECT) { if (wsplit) QB_PACK(NOISE_INJECT, true, true); else QB_PACK(NOISE_INJECT, true, false); }
    else { if (wsplit) QB_PACK(NOISE_PHILOX, true, true); else QB_PACK(NOISE_PHILOX, true, false); }
#undef QB_PACK
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// qb_wt_pack: w_hat fp32 [Cout][Cin][taps] -> bf16 planes wt[plane][Cin][taps][Cout] (hi, lo)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_wt_pack(const float* __restrict__ wq, uint16_t* __restrict__ wt, int Cout, int Cin, int taps) {
    const int64_t total = (int64_t)Cout * Cin * taps;
    for (int64_t o = (int64_t)blockIdx.x * 256 + threadIdx.x; o < total; o += (int64_t)gridDim.x * 256) {
        const int n = (int)(o % Cout);
        const int64_t ct = o / Cout;                                       // c * taps + tap
        const float v = __ldg(wq + (int64_t)n * Cin * taps + ct);
        uint32_t h, l;
        split2(v, 0.f, h, l);
        wt[o] = (uint16_t)h;
        wt[total + o] = (uint16_t)l;
    }
}

extern "C" int qb_wt_pack(const float* w_hat, uint16_t* wt, int64_t Cout, int64_t Cin, int taps, void* stream) {
    QB_CHECK_ARG(w_hat && wt && Cout > 0 && Cin > 0 && taps > 0, "qb_wt_pack: bad argument");
    const int64_t total = Cout * Cin * taps;
    QB_CHECK_ARG(total < 2147483647LL, "qb_wt_pack: weight too large");
    int blocks = (int)((total + 255) / 256);
    if (blocks > 2048) blocks = 2048;
    k_wt_pack<<<blocks, 256, 0, qb_stream(stream)>>>(w_hat, wt, (int)Cout, (int)Cin, taps);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// dgrad
// ------------------------------------------------------------------------------------------------
constexpr int DG_THREADS = 320;      // TMA warp, MMA warp, 2 x 4 epilogue warps
constexpr int DG_MAX_STAGES = 8;
constexpr int DG_MAX_TAPS = 9;
constexpr int DG_SLAB_BYTES = 128 * 128;
constexpr int DG_STAGING_BYTES = 4 * DG_SLAB_BYTES;

struct DgradGeom {
    int B, Hc, Wc;              // output pixel grid of THIS parity class of dX
    int Cn, Ck;                 // GEMM N = conv input channels, GEMM K (per tap) = conv output channels
    int TW, TH, TN, tiles_w, tiles_h, tiles_b;
    int BK;                     // K elements per stage: 64 / 32 / 16  (128 / 64 / 32-byte rows)
    int stages, ntaps, a_planes, b_planes;
    int dh[DG_MAX_TAPS], dw[DG_MAX_TAPS], wtap[DG_MAX_TAPS];   // operand pixel = class pixel + (dh, dw); weight tap index
    FastDiv d_tiles_m, d_tiles_w, d_tiles_h;
};

__device__ __forceinline__ void dg_decode_tile(int t, const DgradGeom& g, int tiles_m, int& tn, int& b0, int& h0, int& w0) {
    tn = (int)fdiv((uint32_t)t, g.d_tiles_m);
    uint32_t tm = (uint32_t)(t - tn * tiles_m);
    uint32_t q = fdiv(tm, g.d_tiles_w);
    w0 = (int)(tm - q * (uint32_t)g.tiles_w) * g.TW;
    tm = q;
    q = fdiv(tm, g.d_tiles_h);
    h0 = (int)(tm - q * (uint32_t)g.tiles_h) * g.TH;
    b0 = (int)q * g.TN;
}

__device__ __forceinline__ void dg_epi_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

template <int BN>
__global__ void __launch_bounds__(DG_THREADS, 1)
k_conv_dgrad(const __grid_constant__ CUtensorMap tmap_a0, const __grid_constant__ CUtensorMap tmap_a1,
             const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_o, DgradGeom g,
             const float* __restrict__ scale_range, int scale_bits, float scale_mul) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int rb = g.BK * 2;                                     // bytes per operand row
    const uint32_t a_bytes = 128u * rb, b_bytes = (uint32_t)BN * rb;
    const uint32_t stage_bytes = g.a_planes * a_bytes + g.b_planes * b_bytes;
    float* staging = reinterpret_cast<float*>(smem + (size_t)g.stages * stage_bytes);       // 1024-byte aligned
    uint8_t* tail = reinterpret_cast<uint8_t*>(staging) + DG_STAGING_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
    uint64_t* empty_bar = full_bar + DG_MAX_STAGES;
    uint64_t* tfull_bar = empty_bar + DG_MAX_STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int kblocks = (g.Ck + g.BK - 1) / g.BK;
    const int num_kb = g.ntaps * kblocks;
    const int tiles_m = g.tiles_w * g.tiles_h * g.tiles_b;
    const int tiles_n = (g.Cn + BN - 1) / BN;
    const int total_tiles = tiles_m * tiles_n;
    constexpr uint32_t ACC = BN < 32 ? 32 : BN;                  // TMEM columns per accumulator buffer
    constexpr uint32_t TMEM_COLS = 2 * ACC;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a0);
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_o);
        for (int s = 0; s < g.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 8);
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            RingPos rp;
            const uint32_t a_box = (uint32_t)(g.TN * g.TH * g.TW) * rb;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                int tn, b0, h0, w0;
                dg_decode_tile(t, g, tiles_m, tn, b0, h0, w0);
                const int n0 = tn * BN;
                for (int tp = 0; tp < g.ntaps; ++tp) {
                    for (int kb = 0; kb < kblocks; ++kb, rp.next(g.stages)) {
                        const int s = rp.s;
                        const uint32_t ph = rp.ph;
                        mbar_wait_trap(&empty_bar[s], ph ^ 1);
                        mbar_arrive_expect_tx(&full_bar[s], g.a_planes * a_box + g.b_planes * b_bytes);
                        uint8_t* sa = smem + (size_t)s * stage_bytes;
                        tma_load_4d(sa, &tmap_a0, &full_bar[s], kb * g.BK, w0 + g.dw[tp], h0 + g.dh[tp], b0);
                        if (g.a_planes > 1)
                            tma_load_4d(sa + a_bytes, &tmap_a1, &full_bar[s], kb * g.BK, w0 + g.dw[tp], h0 + g.dh[tp], b0);
                        uint8_t* sb = sa + g.a_planes * a_bytes;
                        for (int p = 0; p < g.b_planes; ++p)
                            tma_load_3d(sb + (size_t)p * b_bytes, &tmap_w, &full_bar[s], kb * g.BK, g.wtap[tp], p * g.Cn + n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = idesc_bf16(128, BN, 0, 0);
        const int ksteps = g.BK / 16;
        RingPos rp;
        int j = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            const int buf = j & 1;
            mbar_wait_trap(&tempty_bar[buf], ((j >> 1) & 1) ^ 1);
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem_base + (uint32_t)buf * ACC;
            for (int kb = 0; kb < num_kb; ++kb, rp.next(g.stages)) {
                const int s = rp.s;
                const uint32_t ph = rp.ph;
                mbar_wait_trap(&full_bar[s], ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
                    const uint64_t a_desc = desc_kmajor(a_addr, rb), b_desc = desc_kmajor(a_addr + g.a_planes * a_bytes, rb);
                    uint32_t acc = kb != 0;
                    for (int k = 0; k < ksteps; ++k) {
                        // hi*hi, hi*lo, lo*hi  (lo*lo is below the split error)
                        for (int pa = 0; pa < g.a_planes; ++pa)
                            for (int pb = 0; pb < g.b_planes; ++pb) {
                                if (pa == 1 && pb == 1) continue;
                                umma_f16(d_tmem, desc_add(a_desc, pa * a_bytes + k * 32), desc_add(b_desc, pb * b_bytes + k * 32), idesc, acc);
                                acc = 1;
                            }
                    }
                    umma_commit(&empty_bar[s]);
                    if (kb == num_kb - 1) umma_commit(&tfull_bar[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue: 2 groups x 4 warps; group `grp` owns the 32-column slabs cs with cs % 2 == grp =====
        const int e_all = threadIdx.x - 64;
        const int grp = e_all >> 7, et = e_all & 127;
        const int sub = warp & 3;
        float scale = scale_mul;
        if (scale_range) scale *= make_qparams_d((double)__ldg(scale_range), (double)__ldg(scale_range + 1), scale_bits).scale;
        const int i = sub * 32 + lane;
        uint8_t* my_staging = reinterpret_cast<uint8_t*>(staging) + grp * 2 * DG_SLAB_BYTES;
        const uint32_t stage_u32 = smem_u32(my_staging);
        const uint32_t row_off = (uint32_t)i * 128u, sw = (uint32_t)(i & 7);
        const int bar_id = 1 + grp;
        int j = 0;
        uint32_t slab = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            int tn, b0, h0, w0;
            dg_decode_tile(t, g, tiles_m, tn, b0, h0, w0);
            const int n0 = tn * BN;
            const int ncols = (g.Cn - n0 < BN) ? g.Cn - n0 : BN;
            const int nslabs = (ncols + 31) >> 5;
            const int last_cs = ((nslabs - 1 - grp) & ~1) + grp;
            const int buf = j & 1;
            mbar_wait_trap(&tfull_bar[buf], (j >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t)buf * ACC + ((uint32_t)(sub * 32) << 16);
            if (nslabs <= grp) {
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty_bar[buf]);
            }
#pragma unroll 1
            for (int cs = grp; cs < nslabs; cs += 2, ++slab) {
                uint32_t r[32];
                tmem_ld_32x32(t_addr + (uint32_t)(cs * 32), r);
                tmem_ld_wait();
                if (cs == last_cs) {
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                const uint32_t dst = stage_u32 + (slab & 1u) * (uint32_t)DG_SLAB_BYTES + row_off;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float v0 = __uint_as_float(r[4 * q]) * scale, v1 = __uint_as_float(r[4 * q + 1]) * scale,
                                v2 = __uint_as_float(r[4 * q + 2]) * scale, v3 = __uint_as_float(r[4 * q + 3]) * scale;
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)q ^ sw) << 4)), "f"(v0), "f"(v1),
                                 "f"(v2), "f"(v3) : "memory");
                }
                fence_proxy_async();
                if (et == 0) tma_store_wait_read<0>();
                dg_epi_bar(bar_id);
                if (et == 0) {
                    tma_store_4d(&tmap_o, my_staging + (slab & 1u) * DG_SLAB_BYTES, n0 + cs * 32, w0, h0, b0);
                    tma_store_commit();
                }
            }
        }
        if (et == 0) tma_store_wait<0>();
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <int BN>
static int dgrad_launch(const CUtensorMap& ta0, const CUtensorMap& ta1, const CUtensorMap& tw, const CUtensorMap& to, DgradGeom& g,
                        const float* scale_range, int scale_bits, float scale_mul, cudaStream_t st) {
    const size_t stage = (size_t)g.a_planes * 128 * g.BK * 2 + (size_t)g.b_planes * BN * g.BK * 2;
    const size_t fixed = DG_STAGING_BYTES + 1024 + 512;
    int s = (int)((225 * 1024 - fixed) / stage);
    g.stages = s > DG_MAX_STAGES ? DG_MAX_STAGES : s;
    if (g.stages < 2) {
        qb_set_error("qb_conv_dgrad: a pipeline stage of %zu bytes does not fit twice", stage);
        return QB_ERR_UNSUPPORTED;
    }
    const size_t smem = (size_t)g.stages * stage + fixed;
    QB_CUDA(cudaFuncSetAttribute(k_conv_dgrad<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t total = (int64_t)g.tiles_w * g.tiles_h * g.tiles_b * ((g.Cn + BN - 1) / BN);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const unsigned grid = (unsigned)(total < sms ? total : sms);
    k_conv_dgrad<BN><<<grid, DG_THREADS, smem, st>>>(ta0, ta1, tw, to, g, scale_range, scale_bits, scale_mul);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// dgrad, "flat strip" form (unit-stride problems: every parity class is one).
//
// A strip is ONE TMA box {BK channels, Wp, RH, TN} of the gradient codes whose origin lies (dw_min, dh_min) off the strip's
// first output pixel: it lands in shared memory as rows of BK channels in flat order f' = (tn*RH + y')*Wp + x'.  With the
// OUTPUT pixels numbered on the same pitch, f = (tn*RH + y)*Wp + x, the operand row of output f under tap t is f + off[t],
// off[t] = (dh[t]-dh_min)*Wp + (dw[t]-dw_min): every tap reads the SAME box through a UMMA descriptor that simply starts
// off[t] rows later (descriptors may start inside a swizzle atom: tools/umma_shift_probe.py).  The operand is loaded once
// instead of once per tap; zero padding is the TMA's out-of-bounds fill.  Output rows with x >= Wc or y >= Hc are
// garbage: they are staged like the others and clipped by the TMA store.  A strip is <= 256 flat rows = two TMEM
// accumulators that share every weight stage; two accumulator sets alternate between MMA and epilogue.
// ------------------------------------------------------------------------------------------------
constexpr int DF_MAX_WSTAGES = 8;
constexpr int DF_THREADS = DG_THREADS + 32;      // + a second MMA-issuing warp (warp 10) for the second accumulator of a strip

struct DflatGeom {
    int B, Hc, Wc, Cn, Ck;
    int Wp, RH, RHo, TN;        // box = {BK, Wp, RH, TN}; output rows per image slot RHo (row mode: TH, image mode: RH)
    int TH;                     // output rows advanced per strip (row mode) / Hc (image mode)
    int R_out, J;               // flat output rows per strip (<= 256), accumulators (1 or 2)
    int AR;                     // operand rows allocated per A slot
    int tiles_h, tiles_b;
    int BK, kblocks, ntaps, wstages, b_planes;
    int dh_min, dw_min;
    int off[DG_MAX_TAPS], wtap[DG_MAX_TAPS];
    int nstage;                 // staging slab buffers (1 or 2)
    int mma2;                   // second MMA-issuing warp (QB_DGRAD_MMA2=0 disables)
    int accumulate;             // dx += result (TMA reduce-add stores): the skip-connection gradient is already in dx
    FastDiv d_tiles_m, d_tiles_h;
};

// SBS: the hi and lo weight planes sit side by side on the MMA's N axis (one N = 2*BN instruction per K step instead of two
// N = BN ones: the pixel operand is read from shared memory once, which a 64-column MMA needs to keep the tensor pipe fed);
// the epilogue adds the two column halves.
template <int BN, bool SBS>
__global__ void __launch_bounds__(DF_THREADS, 1)
k_conv_dgrad_flat(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w,
                  const __grid_constant__ CUtensorMap tmap_o, DflatGeom g, const float* __restrict__ scale_range, int scale_bits,
                  float scale_mul) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int rb = g.BK * 2;
    const uint32_t a_slot = ((uint32_t)g.AR * rb + 1023u) & ~1023u;
    const uint32_t w_plane = (uint32_t)BN * rb, w_slot = (uint32_t)g.b_planes * w_plane;
    const uint32_t slab_bytes = ((uint32_t)g.R_out * 128u + 1023u) & ~1023u;
    uint8_t* s_a = smem;                                        // 2 slots
    uint8_t* s_w = s_a + 2 * a_slot;
    uint8_t* staging = s_w + (size_t)g.wstages * w_slot;
    uint8_t* tail = staging + (size_t)g.nstage * slab_bytes;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(tail);       // [2]
    uint64_t* a_empty = a_full + 2;
    uint64_t* w_full = a_empty + 2;                             // [DF_MAX_WSTAGES]
    uint64_t* w_empty = w_full + DF_MAX_WSTAGES;
    uint64_t* tfull_bar = w_empty + DF_MAX_WSTAGES;             // [2]
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_m = g.tiles_h * g.tiles_b;
    const int tiles_n = (g.Cn + BN - 1) / BN;
    const int total = tiles_m * tiles_n;
    constexpr uint32_t ACC = SBS ? (2 * BN < 32 ? 32 : 2 * BN) : (BN < 32 ? 32 : BN);
    constexpr uint32_t JMAX = ACC >= 256 ? 1 : 2;               // accumulators per set (a strip is <= 128 * JMAX rows)
    constexpr uint32_t TMEM_COLS = 2 * JMAX * ACC;              // 2 sets x JMAX accumulators
    // one thread issuing every MMA is what bounds this kernel (see k_conv_wgrad): with two accumulators per strip a second
    // warp issues the second one - independent accumulators, no ordering between the warps
    const uint32_t nmma = (g.J == 2 && g.mma2) ? 2u : 1u;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_o);
        for (int s = 0; s < 2; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], nmma); }
        for (int s = 0; s < g.wstages; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], nmma); }
        for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], nmma); mbar_init(&tempty_bar[i], 8); }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    // operand rows behind the box are read (for garbage outputs only) but never written: no NaN patterns there
    for (uint32_t i = threadIdx.x; i < 2 * a_slot / 16; i += DF_THREADS) reinterpret_cast<uint4*>(s_a)[i] = make_uint4(0u, 0u, 0u, 0u);
    fence_proxy_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    auto decode = [&](int t, int& tn, int& b0, int& h0) {
        tn = (int)fdiv((uint32_t)t, g.d_tiles_m);
        const uint32_t tm = (uint32_t)(t - tn * tiles_m);
        const uint32_t q = fdiv(tm, g.d_tiles_h);
        h0 = (int)(tm - q * (uint32_t)g.tiles_h) * g.TH;
        b0 = (int)q * g.TN;
    };

    if (warp == 0) {
        if (elect_one()) {
            const uint32_t a_box = (uint32_t)(g.Wp * g.RH * g.TN) * rb;
            uint32_t ia = 0;
            RingPos rw;
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                int tn, b0, h0;
                decode(t, tn, b0, h0);
                const int n0 = tn * BN;
                for (int kb = 0; kb < g.kblocks; ++kb, ++ia) {
                    const int sa = ia & 1;
                    mbar_wait_trap(&a_empty[sa], ((ia >> 1) & 1) ^ 1);
                    mbar_arrive_expect_tx(&a_full[sa], a_box);
                    tma_load_4d(s_a + (size_t)sa * a_slot, &tmap_a, &a_full[sa], kb * g.BK, g.dw_min, h0 + g.dh_min, b0);
                    for (int tp = 0; tp < g.ntaps; ++tp, rw.next(g.wstages)) {
                        const int sw = rw.s;
                        mbar_wait_trap(&w_empty[sw], rw.ph ^ 1);
                        mbar_arrive_expect_tx(&w_full[sw], w_slot);
                        for (int p = 0; p < g.b_planes; ++p)
                            tma_load_3d(s_w + (size_t)sw * w_slot + (size_t)p * w_plane, &tmap_w, &w_full[sw], kb * g.BK, g.wtap[tp],
                                        p * g.Cn + n0);
                    }
                }
            }
        }
    } else if (warp == 1 || (warp == 10 && nmma == 2)) {
        const int mw = warp == 1 ? 0 : 1;
        constexpr uint32_t idesc = idesc_bf16(128, SBS ? 2 * BN : BN, 0, 0);
        const int ksteps = g.BK / 16;
        const uint64_t w_desc0 = desc_kmajor(smem_u32(s_w), rb);               // weight stage sw: + sw * w_slot
        uint32_t ia = 0;
        RingPos rw;
        int it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const int set = it & 1;
            mbar_wait_trap(&tempty_bar[set], ((it >> 1) & 1) ^ 1);
            tc_fence_after_sync();
            const uint32_t d0 = tmem_base + (uint32_t)set * JMAX * ACC;
            for (int kb = 0; kb < g.kblocks; ++kb, ++ia) {
                const int sa = ia & 1;
                mbar_wait_trap(&a_full[sa], (ia >> 1) & 1);
                const uint64_t a_desc0 = desc_kmajor(smem_u32(s_a + (size_t)sa * a_slot), rb);
                for (int tp = 0; tp < g.ntaps; ++tp, rw.next(g.wstages)) {
                    const int sw = rw.s;
                    mbar_wait_trap(&w_full[sw], rw.ph);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint64_t w_desc = desc_add(w_desc0, (uint32_t)sw * w_slot);
                        const uint32_t first = (kb | tp) == 0;
                        for (int j = 0; j < g.J; ++j) {
                            if ((uint32_t)j % nmma != (uint32_t)mw) continue;
                            const uint64_t a_desc = desc_add(a_desc0, (uint32_t)(128 * j + g.off[tp]) * (uint32_t)rb);
                            const uint32_t dj = d0 + (uint32_t)j * ACC;
                            if (SBS) {
                                uint32_t acc = first ? 0u : 1u;
                                for (int k = 0; k < ksteps; ++k) {
                                    umma_f16(dj, desc_add(a_desc, k * 32), desc_add(w_desc, k * 32), idesc, acc);
                                    acc = 1u;
                                }
                            } else if (ksteps == 4 && g.b_planes == 2) {
                                umma_f16(dj, a_desc, w_desc, idesc, first ? 0u : 1u);
                                umma_f16(dj, a_desc, desc_add(w_desc, w_plane), idesc, 1u);
#pragma unroll
                                for (int k = 1; k < 4; ++k) {
                                    umma_f16(dj, desc_add(a_desc, k * 32), desc_add(w_desc, k * 32), idesc, 1u);
                                    umma_f16(dj, desc_add(a_desc, k * 32), desc_add(w_desc, w_plane + k * 32), idesc, 1u);
                                }
                            } else {
                                uint32_t acc = first ? 0u : 1u;
                                for (int k = 0; k < ksteps; ++k)
                                    for (int p = 0; p < g.b_planes; ++p) {
                                        umma_f16(dj, desc_add(a_desc, k * 32), desc_add(w_desc, (uint32_t)p * w_plane + k * 32), idesc, acc);
                                        acc = 1u;
                                    }
                            }
                        }
                        umma_commit(&w_empty[sw]);
                        if (tp == g.ntaps - 1) umma_commit(&a_empty[sa]);
                        if (tp == g.ntaps - 1 && kb == g.kblocks - 1) umma_commit(&tfull_bar[set]);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp >= 2 && warp <= 9) {
        // ===== epilogue: warps 2-5 drain accumulator 0, warps 6-9 accumulator 1 (one warp per TMEM lane quarter) =====
        const int e_all = threadIdx.x - 64;
        const int sub = warp & 3, j = (warp - 2) >> 2;
        float scale = scale_mul;
        if (scale_range) scale *= make_qparams_d((double)__ldg(scale_range), (double)__ldg(scale_range + 1), scale_bits).scale;
        const int f = j * 128 + sub * 32 + lane;                 // flat output row of this thread
        const bool live = j < g.J && f < g.R_out;
        const uint32_t row_off = (uint32_t)f * 128u, sw = (uint32_t)(f & 7);
        const uint32_t stage_u32 = smem_u32(staging);
        int it = 0;
        uint32_t slab = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            int tn, b0, h0;
            decode(t, tn, b0, h0);
            const int n0 = tn * BN;
            const int ncols = (g.Cn - n0 < BN) ? g.Cn - n0 : BN;
            const int nslabs = (ncols + 31) >> 5;
            const int set = it & 1;
            mbar_wait_trap(&tfull_bar[set], (it >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + ((uint32_t)set * JMAX + (uint32_t)j) * ACC + ((uint32_t)(sub * 32) << 16);
#pragma unroll 1
            for (int cs = 0; cs < nslabs; ++cs, ++slab) {
                uint32_t r[32];
                if (j < g.J) {
                    tmem_ld_32x32(t_addr + (uint32_t)(cs * 32), r);
                    if (SBS) {
                        uint32_t r2[32];
                        tmem_ld_32x32(t_addr + (uint32_t)(BN + cs * 32), r2);
                        tmem_ld_wait();
#pragma unroll
                        for (int q = 0; q < 32; ++q) r[q] = __float_as_uint(__uint_as_float(r[q]) + __uint_as_float(r2[q]));
                    } else {
                        tmem_ld_wait();
                    }
                }
                if (cs == nslabs - 1) {
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[set]);
                }
                const uint32_t sbuf = g.nstage == 2 ? (slab & 1u) : 0u;
                // the store that last used this staging buffer has finished reading it
                if (e_all == 0) { if (g.nstage == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>(); }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (live) {
                    const uint32_t dst = stage_u32 + sbuf * slab_bytes + row_off;
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const float v0 = __uint_as_float(r[4 * q]) * scale, v1 = __uint_as_float(r[4 * q + 1]) * scale,
                                    v2 = __uint_as_float(r[4 * q + 2]) * scale, v3 = __uint_as_float(r[4 * q + 3]) * scale;
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)q ^ sw) << 4)), "f"(v0),
                                     "f"(v1), "f"(v2), "f"(v3) : "memory");
                    }
                }
                fence_proxy_async();
                asm volatile("bar.sync 2, 256;" ::: "memory");
                if (e_all == 0) {
                    if (g.accumulate) tma_reduce_add_4d(&tmap_o, staging + (size_t)sbuf * slab_bytes, n0 + cs * 32, 0, h0, b0);
                    else tma_store_4d(&tmap_o, staging + (size_t)sbuf * slab_bytes, n0 + cs * 32, 0, h0, b0);
                    tma_store_commit();
                }
            }
        }
        if (e_all == 0) tma_store_wait<0>();
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <int BN, bool SBS>
static int dgrad_flat_launch(const CUtensorMap& ta, const CUtensorMap& tw, const CUtensorMap& to, const DflatGeom& g, size_t smem,
                             const float* scale_range, int scale_bits, float scale_mul, cudaStream_t st) {
    QB_CUDA(cudaFuncSetAttribute(k_conv_dgrad_flat<BN, SBS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t total = (int64_t)g.tiles_h * g.tiles_b * ((g.Cn + BN - 1) / BN);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const unsigned grid = (unsigned)(total < sms ? total : sms);
    k_conv_dgrad_flat<BN, SBS><<<grid, DF_THREADS, smem, st>>>(ta, tw, to, g, scale_range, scale_bits, scale_mul);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// Strip geometry of one parity class; returns the dynamic shared memory it needs, 0 if the flat form does not apply.
static size_t dgrad_flat_geom(DflatGeom& g, int B, int Hc, int Wc, int Cn, int Ck, int BK, int BN, const int* dh, const int* dw,
                              const int* wtap, int ntaps, int max_rows = 256) {
    memset(&g, 0, sizeof(g));
    int dh_min = dh[0], dh_max = dh[0], dw_min = dw[0], dw_max = dw[0];
    for (int i = 1; i < ntaps; ++i) {
        dh_min = dh[i] < dh_min ? dh[i] : dh_min; dh_max = dh[i] > dh_max ? dh[i] : dh_max;
        dw_min = dw[i] < dw_min ? dw[i] : dw_min; dw_max = dw[i] > dw_max ? dw[i] : dw_max;
    }
    const int halo_h = dh_max - dh_min, halo_w = dw_max - dw_min;
    g.B = B; g.Hc = Hc; g.Wc = Wc; g.Cn = Cn; g.Ck = Ck; g.BK = BK; g.ntaps = ntaps; g.b_planes = 2;
    g.dh_min = dh_min; g.dw_min = dw_min;
    // Shared padding: on the flat pitch the right-hand zero columns of row y ARE the left-hand zero columns of row y+1 (and
    // the bottom zero rows of an image the top zero rows of the next one), so the pitch only needs the larger of the two
    // halos: Wp = Wc + max(dw_max, -dw_min) instead of Wc + (dw_max - dw_min).  49 of 64 rows are real for a 7x7 image
    // under a 3x3 filter instead of 49 of 81.
    auto pos = [](int v) { return v > 0 ? v : 0; };
    const int ext_w = pos(dw_max) > pos(-dw_min) ? pos(dw_max) : pos(-dw_min);
    const int ext_h = pos(dh_max) > pos(-dh_min) ? pos(dh_max) : pos(-dh_min);
    g.Wp = Wc + ext_w;
    if (g.Wp > max_rows) return 0;
    const int Hp = Hc + ext_h;
    if (Hp * g.Wp <= max_rows) {                      // image mode: whole padded images
        g.RH = g.RHo = Hp;
        g.TN = max_rows / (Hp * g.Wp);
        if (g.TN > B) g.TN = B;
        g.TH = Hc;
        g.tiles_h = 1;
    } else {                                     // row mode
        g.TN = 1;
        g.TH = max_rows / g.Wp;
        if (g.TH > Hc) g.TH = Hc;
        g.RHo = g.TH;
        g.RH = g.TH + halo_h;
        if (g.RH > 256) return 0;
        g.tiles_h = (Hc + g.TH - 1) / g.TH;
    }
    g.tiles_b = (B + g.TN - 1) / g.TN;
    g.R_out = g.TN * g.RHo * g.Wp;
    g.J = (g.R_out + 127) / 128;
    const int box_rows = g.Wp * g.RH * g.TN;
    int need = 128 * g.J + halo_h * g.Wp + halo_w;
    if (need < box_rows) need = box_rows;
    g.AR = (need + 7) & ~7;
    for (int i = 0; i < ntaps; ++i) {
        g.off[i] = (dh[i] - dh_min) * g.Wp + (dw[i] - dw_min);
        g.wtap[i] = wtap[i];
    }
    g.kblocks = (Ck + BK - 1) / BK;
    static const bool mma2_off = getenv("QB_DGRAD_MMA2") && getenv("QB_DGRAD_MMA2")[0] == '0';
    g.mma2 = mma2_off ? 0 : 1;
    g.d_tiles_m = make_fastdiv((uint32_t)(g.tiles_h * g.tiles_b));
    g.d_tiles_h = make_fastdiv((uint32_t)g.tiles_h);
    const int rb = BK * 2;
    const size_t a_slot = ((size_t)g.AR * rb + 1023) & ~(size_t)1023, w_slot = (size_t)2 * BN * rb;
    const size_t slab = ((size_t)g.R_out * 128 + 1023) & ~(size_t)1023;
    const size_t limit = 226 * 1024, fixed = 1024 + 512 + 2 * a_slot;
    if (fixed + slab + 2 * w_slot > limit) return 0;
    // weight ring first (the stream that must stay in flight), then a second staging buffer if it still fits
    int ws = (int)((limit - fixed - slab) / w_slot);
    g.nstage = 1;
    if (ws >= 5 && fixed + 2 * slab + 4 * w_slot <= limit) {
        g.nstage = 2;
        ws = (int)((limit - fixed - 2 * slab) / w_slot);
    }
    g.wstages = ws > DF_MAX_WSTAGES ? DF_MAX_WSTAGES : ws;
    return fixed + (size_t)g.nstage * slab + (size_t)g.wstages * w_slot;
}

/* dX of y = conv2d(x, w_hat) (groups 1, dilation 1, square stride / padding):
 *   a0 (and a1): bf16 [B,Ho,Wo,Cout] channels-last planes of the output gradient (one plane: integer codes q - zp, or two
 *   planes: hi / lo split);  wt: qb_wt_pack's planes;  dx: fp32 [B,H,W,Cin] channels-last.
 *   dx = scale * SUM, scale = scale_mul * (scale_range ? s(scale_range, scale_bits) : 1)  (s = the zero-point quantizer's
 *   step, quantize.py:78-80).  Linear layers: H = W = kh = kw = 1.                                                       */
static int conv_dgrad_impl(const uint16_t* a0, const uint16_t* a1, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W,
                           int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* scale_range,
                           int scale_bits, float scale_mul, int accumulate, void* stream);

extern "C" int qb_conv_dgrad(const uint16_t* a0, const uint16_t* a1, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W,
                             int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* scale_range,
                             int scale_bits, float scale_mul, void* stream) {
    return conv_dgrad_impl(a0, a1, wt, dx, B, H, W, Cin, Cout, kh, kw, stride, pad, scale_range, scale_bits, scale_mul, 0, stream);
}

/* As qb_conv_dgrad, but dx += result: dx already holds the OTHER gradient that reaches the layer's input - the skip connection
 * of a residual block (models/resnet_quantized.py:62-68: `out += residual`; autograd's separate accumulation pass of 12 B/elem
 * disappears into TMA reduce-add stores).  Single-plane (code) gradients, every pixel parity class of the stride reached by the
 * filter and taken by the flat-strip kernel (checked before anything is added); anything else returns QB_ERR_UNSUPPORTED with dx
 * untouched. */
extern "C" int qb_conv_dgrad_acc(const uint16_t* a0, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W, int64_t Cin,
                                 int64_t Cout, int kh, int kw, int stride, int pad, const float* scale_range, int scale_bits,
                                 float scale_mul, void* stream) {
    return conv_dgrad_impl(a0, nullptr, wt, dx, B, H, W, Cin, Cout, kh, kw, stride, pad, scale_range, scale_bits, scale_mul, 1, stream);
}

static int conv_dgrad_impl(const uint16_t* a0, const uint16_t* a1, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W,
                           int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* scale_range,
                           int scale_bits, float scale_mul, int accumulate, void* stream) {
    QB_CHECK_ARG(a0 && wt && dx, "qb_conv_dgrad: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0, "qb_conv_dgrad: bad shape");
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (Ho < 1 || Wo < 1 || Cin % 4 != 0 || Cout % 8 != 0 || stride > 4 ||
        ((reinterpret_cast<uintptr_t>(a0) | reinterpret_cast<uintptr_t>(a1) | reinterpret_cast<uintptr_t>(wt) |
          reinterpret_cast<uintptr_t>(dx)) & 15) != 0) {
        qb_set_error("qb_conv_dgrad: unsupported geometry (need Cin%%4==0, Cout%%8==0, stride<=4, 16-byte aligned tensors): "
                     "Cin=%lld Cout=%lld k=%dx%d stride=%d", (long long)Cin, (long long)Cout, kh, kw, stride);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(B * H * W < 2147483647LL && B * Ho * Wo < 2147483647LL, "qb_conv_dgrad: pixel count exceeds int32");
    cudaStream_t st = qb_stream(stream);
    const int BK = Cout > 32 ? 64 : (Cout > 16 ? 32 : 16);
    const int BN = Cin >= 128 ? 128 : (Cin > 32 ? 64 : (Cin > 16 ? 32 : 16));
    CUtensorMap ta0, ta1, tw;
    {
        const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)B};
        const uint64_t str[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
        // box filled in per class (tile shape)
        (void)dims; (void)str;
    }
    {   // weights: dims {Cout (K), taps, planes * Cin (rows)}
        const uint64_t dims[3] = {(uint64_t)Cout, (uint64_t)(kh * kw), (uint64_t)(2 * Cin)};
        const uint64_t str[2] = {(uint64_t)Cout * 2, (uint64_t)Cout * 2 * kh * kw};
        const uint32_t box[3] = {(uint32_t)BK, 1u, (uint32_t)BN};
        int rc = qb_make_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, wt, dims, str, box, nullptr, BK * 2);
        if (rc) return rc;
    }
    // parity classes of dX: pixel h = stride*i + ph receives tap r iff (ph + pad - r) % stride == 0, from ho = i + (ph + pad - r)/stride
    bool any_empty = false;
    for (int ph = 0; ph < stride && ph < H; ++ph)
        for (int pw = 0; pw < stride && pw < W; ++pw) {
            int cnt = 0;
            for (int r = 0; r < kh; ++r)
                for (int s = 0; s < kw; ++s)
                    if ((ph + pad - r) % stride == 0 && (pw + pad - s) % stride == 0) ++cnt;
            if (cnt == 0) any_empty = true;
        }
    if (accumulate && (any_empty || a1)) {
        qb_set_error("qb_conv_dgrad_acc: needs a single-plane gradient and a filter that reaches every pixel parity");
        return QB_ERR_UNSUPPORTED;
    }
    if (any_empty) QB_CUDA(cudaMemsetAsync(dx, 0, (size_t)B * H * W * Cin * 4, st));
    // accumulate: a dry pass first - every parity class must be taken by the flat-strip kernel BEFORE anything is added to dx
    for (int pass = accumulate ? 0 : 1; pass < 2; ++pass) {
    const bool dry = pass == 0;
    for (int ph = 0; ph < stride && ph < H; ++ph) {
        for (int pw = 0; pw < stride && pw < W; ++pw) {
            DgradGeom g;
            memset(&g, 0, sizeof(g));
            g.ntaps = 0;
            for (int r = 0; r < kh; ++r)
                for (int s = 0; s < kw; ++s) {
                    // C's % on negative numbers: compare against a non-negative residue
                    const int ah = ph + pad - r, aw = pw + pad - s;
                    if (((ah % stride) + stride) % stride != 0 || ((aw % stride) + stride) % stride != 0) continue;
                    if (g.ntaps == DG_MAX_TAPS) {
                        qb_set_error("qb_conv_dgrad: more than %d taps per parity class (k=%dx%d stride=%d)", DG_MAX_TAPS, kh, kw, stride);
                        return QB_ERR_UNSUPPORTED;
                    }
                    // floor division (ah may be negative)
                    g.dh[g.ntaps] = (ah >= 0 ? ah : ah - stride + 1) / stride;
                    g.dw[g.ntaps] = (aw >= 0 ? aw : aw - stride + 1) / stride;
                    g.wtap[g.ntaps] = r * kw + s;
                    ++g.ntaps;
                }
            if (g.ntaps == 0) continue;
            const int64_t Hc = (H - ph + stride - 1) / stride, Wc = (W - pw + stride - 1) / stride;
            static const bool flat_off = getenv("QB_DGRAD_FLAT") && getenv("QB_DGRAD_FLAT")[0] == '0';
            if (!a1 && !flat_off) {
                DflatGeom fg;
                static const bool sbs128 = getenv("QB_DGRAD_SBS128") && getenv("QB_DGRAD_SBS128")[0] == '1';
                const size_t fsmem = dgrad_flat_geom(fg, (int)B, (int)Hc, (int)Wc, (int)Cin, (int)Cout, BK, BN, g.dh, g.dw, g.wtap, g.ntaps,
                                                     (BN == 128 && sbs128) ? 128 : 256);
                if (fsmem && dry) continue;
                if (fsmem) {
                    fg.accumulate = accumulate;
                    CUtensorMap fa, fo;
                    const uint64_t adims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)B};
                    const uint64_t astr[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
                    const uint32_t abox[4] = {(uint32_t)BK, (uint32_t)fg.Wp, (uint32_t)fg.RH, (uint32_t)fg.TN};
                    int rc = qb_make_tmap(&fa, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a0, adims, astr, abox, nullptr, BK * 2);
                    if (rc) return rc;
                    float* base = dx + ((int64_t)ph * W + pw) * Cin;
                    const uint64_t odims[4] = {(uint64_t)Cin, (uint64_t)Wc, (uint64_t)Hc, (uint64_t)B};
                    const uint64_t ostr[3] = {(uint64_t)stride * Cin * 4, (uint64_t)stride * W * Cin * 4, (uint64_t)H * W * Cin * 4};
                    const uint32_t obox[4] = {32u, (uint32_t)fg.Wp, (uint32_t)fg.RHo, (uint32_t)fg.TN};
                    rc = qb_make_tmap(&fo, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, odims, ostr, obox, nullptr, 128);
                    if (rc) return rc;
                    static const bool sbs_off = getenv("QB_DGRAD_SBS") && getenv("QB_DGRAD_SBS")[0] == '0';
                    if (BN == 128 && sbs128) rc = dgrad_flat_launch<128, true>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    else if (BN == 128) rc = dgrad_flat_launch<128, false>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    else if (BN == 64 && !sbs_off) rc = dgrad_flat_launch<64, true>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    else if (BN == 64) rc = dgrad_flat_launch<64, false>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    else if (BN == 32) rc = dgrad_flat_launch<32, true>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    else rc = dgrad_flat_launch<16, true>(fa, tw, fo, fg, fsmem, scale_range, scale_bits, scale_mul, st);
                    if (rc) return rc;
                    continue;
                }
            }
            if (accumulate) {
                qb_set_error("qb_conv_dgrad_acc: geometry not taken by the flat-strip kernel");
                return QB_ERR_UNSUPPORTED;
            }
            g.B = (int)B; g.Hc = (int)Hc; g.Wc = (int)Wc; g.Cn = (int)Cin; g.Ck = (int)Cout;
            g.BK = BK;
            g.a_planes = a1 ? 2 : 1;
            g.b_planes = 2;
            g.TW = (int)(Wc < 128 ? Wc : 128);
            g.TH = (int)(128 / g.TW < Hc ? 128 / g.TW : Hc);
            g.TN = (g.TH == Hc && g.TW == Wc) ? (int)(128 / (Hc * Wc) < B ? 128 / (Hc * Wc) : B) : 1;
            if (g.TN < 1) g.TN = 1;
            g.tiles_w = (int)((Wc + g.TW - 1) / g.TW);
            g.tiles_h = (int)((Hc + g.TH - 1) / g.TH);
            g.tiles_b = (int)((B + g.TN - 1) / g.TN);
            g.d_tiles_m = make_fastdiv((uint32_t)(g.tiles_w * g.tiles_h * g.tiles_b));
            g.d_tiles_w = make_fastdiv((uint32_t)g.tiles_w);
            g.d_tiles_h = make_fastdiv((uint32_t)g.tiles_h);
            {
                const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)B};
                const uint64_t str[3] = {(uint64_t)Cout * 2, (uint64_t)Wo * Cout * 2, (uint64_t)Ho * Wo * Cout * 2};
                const uint32_t box[4] = {(uint32_t)BK, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
                int rc = qb_make_tmap(&ta0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a0, dims, str, box, nullptr, BK * 2);
                if (rc) return rc;
                rc = qb_make_tmap(&ta1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, a1 ? a1 : a0, dims, str, box, nullptr, BK * 2);
                if (rc) return rc;
            }
            CUtensorMap to;
            {   // strided view of dX holding this class: dims {Cin, Wc, Hc, B}
                float* base = dx + ((int64_t)ph * W + pw) * Cin;
                const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)Wc, (uint64_t)Hc, (uint64_t)B};
                const uint64_t str[3] = {(uint64_t)stride * Cin * 4, (uint64_t)stride * W * Cin * 4, (uint64_t)H * W * Cin * 4};
                const uint32_t box[4] = {32u, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
                int rc = qb_make_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, dims, str, box, nullptr, 128);
                if (rc) return rc;
            }
            int rc;
            if (BN == 128) rc = dgrad_launch<128>(ta0, ta1, tw, to, g, scale_range, scale_bits, scale_mul, st);
            else if (BN == 64) rc = dgrad_launch<64>(ta0, ta1, tw, to, g, scale_range, scale_bits, scale_mul, st);
            else if (BN == 32) rc = dgrad_launch<32>(ta0, ta1, tw, to, g, scale_range, scale_bits, scale_mul, st);
            else rc = dgrad_launch<16>(ta0, ta1, tw, to, g, scale_range, scale_bits, scale_mul, st);
            if (rc) return rc;
        }
    }
    }   // pass
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// wgrad
// ------------------------------------------------------------------------------------------------
constexpr int WG_THREADS = 224;      // TMA warp, MMA warp, 4 drain warps, second MMA warp
// The tensor pipe idles half of the time in this kernel although its operands are there (profiles/ncu_conv_l2_r02e.txt): ONE
// thread issuing the MMAs (descriptor arithmetic, uniform-register moves, barrier waits) is the bottleneck.  With two or more
// accumulators per CTA a second warp issues the odd ones: they are independent, so no ordering between the warps is needed.
constexpr int WG_MAX_STAGES = 8;
constexpr int WG_MAX_ACC = 16;

struct WgradGeom {
    int B, Ho, Wo, Cin, Cout, kh, kw, sh, sw, ph, pw;     // (Cin, kh, kw, strides, pads) may describe a VIRTUAL operand (stem)
    int TW, TH, TN, tiles_w, tiles_h, tiles_b, P;     // pixel tile (= K block), P = TW*TH*TN, P % 16 == 0
    int RC, RN;                 // channels per TMA box ("run") of the x / g operand: 64 / 32 / 16
    int rpm;                    // runs per 128-row M tile = 128 / RC
    int total_runs, mtiles;     // taps * Cin / RC;  ceil(total_runs / rpm)
    int J;                      // M tiles (accumulators) per CTA
    int mgroups, nblocks, slices, BN;
    int g_planes, a_stages, b_stages, total_tiles, mrows;
    // "flat row" mode (unit stride, Cin % 128 == 0): a CTA owns ONE filter row r and one block of 128 input channels; its kw
    // accumulators (the taps (r, 0..kw-1)) all read the SAME x strip - pixels on a padded pitch Wp = Wo + kw - 1, one TMA box
    // per 64-channel run - through MN-major descriptors that start s rows (pixels) later, so x is fetched once per filter row
    // instead of once per tap.  The gradient tile sits on the same pitch (its out-of-range columns are the TMA's zero fill,
    // so products that wrap into the next image row vanish).  P is the K extent (padded to 16), Pbox the pixels a box holds.
    int flat, Wp, Pbox, arows;
    // flat == 2 (unit stride, Cin == 64, all taps in one CTA): the strip also carries the kh-1 halo rows, tap (r, s) starts
    // r*Wp + s rows into it, and ONE M = 128 accumulator holds TWO taps: its second 64-row column block is the same strip read
    // foff-difference rows later - the descriptor's leading-dimension byte offset is simply that distance (blocks overlap).
    int mma2;                   // a second warp issues the odd accumulators (QB_WGRAD_MMA2=0 disables)
    int uissue;                 // sbs with the issue loop on the uniform datapath (experiment: QB_WGRAD_UISSUE=1)
    int sbs;                    // hi | lo gradient planes side by side on N: ONE N = 256 MMA per K step (the single issuing
                                // thread, not the tensor pipe, bounds this kernel), column halves added in the drain
    int xrows;                  // rows of the x box (TH + kh - 1) in mode 2
    int foff[WG_MAX_ACC], flbo[WG_MAX_ACC];     // per accumulator: first strip row, block distance in rows
    FastDiv d_tiles_w, d_tiles_h, d_cin_runs, d_kw;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
k_conv_wgrad(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_g0,
             const __grid_constant__ CUtensorMap tmap_g1, float* __restrict__ partial, WgradGeom g) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t a_slot = g.flat == 2 ? (uint32_t)g.arows * 128u             // one 64-channel strip
                            : g.flat ? 2u * (uint32_t)g.arows * 128u          // two 64-channel runs x [arows][64] bf16
                                   : 128u * (uint32_t)g.P * 2u;               // rpm runs x [P][RC] bf16
    const uint32_t b_plane = (uint32_t)g.BN * (uint32_t)g.P * 2u;
    const uint32_t b_slot = (uint32_t)g.g_planes * b_plane;
    uint8_t* s_a = smem;
    uint8_t* s_b = s_a + (size_t)g.a_stages * a_slot;
    uint8_t* tail = s_b + (size_t)g.b_stages * b_slot;
    uint64_t* a_full = reinterpret_cast<uint64_t*>(tail);
    uint64_t* a_empty = a_full + WG_MAX_STAGES;
    uint64_t* b_full = a_empty + WG_MAX_STAGES;
    uint64_t* b_empty = b_full + WG_MAX_STAGES;
    uint64_t* done_bar = b_empty + WG_MAX_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // unit -> (slice, M group, N block)
    const int unit = blockIdx.x;
    const int nb = unit % g.nblocks;
    const int mg = (unit / g.nblocks) % g.mgroups;
    const int slice = unit / (g.nblocks * g.mgroups);
    const int n0 = nb * g.BN;
    const int mt0 = mg * g.J;
    const int jn = g.flat ? g.J : ((g.mtiles - mt0 < g.J) ? g.mtiles - mt0 : g.J);   // accumulators of this CTA
    const int cgroups = g.Cin >> 7;                                           // flat mode 1: mg = r * cgroups + cg
    const int fr = g.flat == 1 ? mg / cgroups : 0, fcg = g.flat == 1 ? mg - fr * cgroups : 0;
    const int t_begin = (int)(((int64_t)g.total_tiles * slice) / g.slices);
    const int t_end = (int)(((int64_t)g.total_tiles * (slice + 1)) / g.slices);
    const uint32_t acc_stride = g.sbs ? 2u * (uint32_t)g.BN : (uint32_t)(g.BN < 32 ? 32 : g.BN);
    const int nmma = (g.mma2 && !g.flat && !g.uissue && jn >= 2) ? 2 : 1;    // MMA-issuing warps of this CTA

    // a tail M tile has fewer runs than rpm: its unused operand rows must not hold NaN patterns
    // (flat: neither the x rows behind a box - read by the shifted taps - nor the gradient rows between Pbox and P)
    for (uint32_t i = threadIdx.x; i < ((uint32_t)g.a_stages * a_slot + (g.flat ? (uint32_t)g.b_stages * b_slot : 0u)) / 16; i += WG_THREADS)
        reinterpret_cast<uint4*>(s_a)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_g0);
        for (int s = 0; s < g.a_stages; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
        for (int s = 0; s < g.b_stages; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], (uint32_t)nmma); }
        mbar_init(done_bar, (uint32_t)nmma);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<512>(tmem_slot);
    fence_proxy_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;
    const int rbA = g.RC * 2, rbB = g.RN * 2;
    const uint32_t runA = (g.flat ? (uint32_t)g.arows : (uint32_t)g.P) * rbA, runB = (uint32_t)g.P * rbB;   // run pitch = LBO
    const int nruns_b = g.BN / g.RN;

    auto decode = [&](int t, int& b0, int& ho0, int& wo0) {
        uint32_t q = fdiv((uint32_t)t, g.d_tiles_w);
        wo0 = (int)((uint32_t)t - q * (uint32_t)g.tiles_w) * g.TW;
        const uint32_t q2 = fdiv(q, g.d_tiles_h);
        ho0 = (int)(q - q2 * (uint32_t)g.tiles_h) * g.TH;
        b0 = (int)q2 * g.TN;
    };

    if (warp == 0) {
        if (elect_one()) {
            RingPos ra, rb_;
            for (int t = t_begin; t < t_end; ++t) {
                int b0, ho0, wo0;
                decode(t, b0, ho0, wo0);
                {   // gradient tile (B operand), shared by every accumulator of this CTA
                    const int s = rb_.s;
                    const uint32_t ph = rb_.ph;
                    mbar_wait_trap(&b_empty[s], ph ^ 1);
                    mbar_arrive_expect_tx(&b_full[s], g.flat ? (uint32_t)(g.g_planes * g.BN * g.Pbox * 2) : b_slot);
                    uint8_t* sb = s_b + (size_t)s * b_slot;
                    for (int p = 0; p < g.g_planes; ++p)
                        for (int r = 0; r < nruns_b; ++r)
                            tma_load_4d(sb + (size_t)p * b_plane + (size_t)r * runB, p ? &tmap_g1 : &tmap_g0, &b_full[s],
                                        n0 + r * g.RN, wo0, ho0, b0);
                    rb_.next(g.b_stages);
                }
                if (g.flat) {                                  // one x strip per pixel tile: two 64-channel boxes on the padded pitch
                    const int s = ra.s;
                    mbar_wait_trap(&a_empty[s], ra.ph ^ 1);
                    uint8_t* sa = s_a + (size_t)s * a_slot;
                    if (g.flat == 2) {
                        mbar_arrive_expect_tx(&a_full[s], (uint32_t)(g.xrows * g.Wp) * rbA);
                        tma_load_4d(sa, &tmap_x, &a_full[s], 0, -g.pw, ho0 - g.ph, b0);
                    } else {
                        mbar_arrive_expect_tx(&a_full[s], 2u * (uint32_t)g.Pbox * rbA);
                        for (int rr = 0; rr < 2; ++rr)
                            tma_load_4d(sa + (size_t)rr * runA, &tmap_x, &a_full[s], fcg * 128 + rr * 64, -g.pw, ho0 - g.ph + fr, b0);
                    }
                    ra.next(g.a_stages);
                    continue;
                }
                for (int j = 0; j < jn; ++j, ra.next(g.a_stages)) {
                    const int s = ra.s;
                    const uint32_t ph = ra.ph;
                    const int run0 = (mt0 + j) * g.rpm;
                    const int nr = (g.total_runs - run0 < g.rpm) ? g.total_runs - run0 : g.rpm;
                    mbar_wait_trap(&a_empty[s], ph ^ 1);
                    mbar_arrive_expect_tx(&a_full[s], (uint32_t)nr * runA);
                    uint8_t* sa = s_a + (size_t)s * a_slot;
                    for (int rr = 0; rr < nr; ++rr) {
                        const uint32_t run = (uint32_t)(run0 + rr);
                        const uint32_t tap = fdiv(run, g.d_cin_runs);                       // run / (Cin / RC)
                        const int c0 = (int)(run - tap * (uint32_t)(g.Cin / g.RC)) * g.RC;
                        const uint32_t r = fdiv(tap, g.d_kw);
                        const int s_ = (int)(tap - r * (uint32_t)g.kw);
                        tma_load_4d(sa + (size_t)rr * runA, &tmap_x, &a_full[s], c0, wo0 * g.sw - g.pw + s_,
                                    ho0 * g.sh - g.ph + (int)r, b0);
                    }
                }
            }
        }
    } else if (warp == 1 || (warp == 6 && nmma == 2)) {
        const int mw = warp == 1 ? 0 : 1;                       // this warp issues the accumulators j with j % nmma == mw
        const uint32_t idesc = idesc_bf16(128, (uint32_t)g.BN, 1, 1);
        const int ksteps = g.P / 16;
        const uint32_t lead = elect_one() ? 1u : 0u;           // the one lane that issues (uissue mode)
        int j_last = jn - 1;                                    // the last accumulator this warp owns
        while (j_last % nmma != mw) --j_last;
        RingPos ra, rb_;
        const uint64_t a_ring0 = desc_mnmajor(smem_u32(s_a), rbA, runA), b_ring0 = desc_mnmajor(smem_u32(s_b), rbB, runB);
        for (int t = t_begin; t < t_end; ++t, rb_.next(g.b_stages)) {
            const int sbi = rb_.s;
            mbar_wait_trap(&b_full[sbi], rb_.ph);
            const uint64_t b_desc = desc_add(b_ring0, (uint32_t)sbi * b_slot);
            if (g.uissue) {                                     // N = 256 planes side by side, descriptors on the uniform datapath
                const uint32_t idesc2 = idesc_bf16(128, 256u, 1, 1);
                const uint32_t stepA = 16u * (uint32_t)rbA, stepB = 16u * (uint32_t)rbB;
                for (int j = 0; j < jn; ++j, ra.next(g.a_stages)) {
                    const int s = ra.s;
                    mbar_wait_trap(&a_full[s], ra.ph);
                    tc_fence_after_sync();
                    uint64_t a_desc = desc_mnmajor(smem_u32(s_a + (size_t)s * a_slot), rbA, runA);
                    uint64_t b_d = b_desc;
                    const uint32_t d_tmem = tmem_base + (uint32_t)j * acc_stride;
                    uint32_t acc = t != t_begin;
                    for (int ks = 0; ks < ksteps; ++ks) {
                        umma_f16_if(lead, d_tmem, a_desc, b_d, idesc2, acc);
                        a_desc = desc_add(a_desc, stepA);
                        b_d = desc_add(b_d, stepB);
                        acc = 1u;
                    }
                    umma_commit_if(lead, &a_empty[s]);
                    if (j == jn - 1) umma_commit_if(lead, &b_empty[sbi]);
                    if (j == jn - 1 && t == t_end - 1) umma_commit_if(lead, done_bar);
                }
                continue;
            }
            if (g.flat) {
                const int s = ra.s;
                mbar_wait_trap(&a_full[s], ra.ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint64_t a_desc0 = desc_mnmajor(smem_u32(s_a + (size_t)s * a_slot), rbA, runA);
                    const uint32_t stepA = 16u * (uint32_t)rbA, stepB = 16u * (uint32_t)rbB;
                    const uint32_t acc0 = t != t_begin;
                    for (int j = 0; j < jn; ++j) {               // mode 1: tap (fr, j) = the strip read j pixels later
                        const uint64_t a_desc = g.flat == 2
                            ? desc_mnmajor(smem_u32(s_a + (size_t)s * a_slot) + (uint32_t)g.foff[j] * (uint32_t)rbA, rbA,
                                           (uint32_t)g.flbo[j] * (uint32_t)rbA)
                            : desc_add(a_desc0, (uint32_t)j * (uint32_t)rbA);
                        const uint32_t d_tmem = tmem_base + (uint32_t)j * acc_stride;
                        uint32_t acc = acc0;
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, ks * stepB), idesc, acc);
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, b_plane + ks * stepB), idesc, 1u);
                            acc = 1u;
                        }
                    }
                    umma_commit(&a_empty[s]);
                    umma_commit(&b_empty[sbi]);
                    if (t == t_end - 1) umma_commit(done_bar);
                }
                __syncwarp();
                ra.next(g.a_stages);
                continue;
            }
            for (int j = 0; j < jn; ++j, ra.next(g.a_stages)) {
                if (j % nmma != mw) continue;                   // (the slot position still advances)
                const int s = ra.s;
                mbar_wait_trap(&a_full[s], ra.ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint64_t a_desc = desc_add(a_ring0, (uint32_t)s * a_slot);
                    const uint32_t d_tmem = tmem_base + (uint32_t)j * acc_stride;
                    const uint32_t stepA = 16u * (uint32_t)rbA, stepB = 16u * (uint32_t)rbB;
                    uint32_t acc = t != t_begin;
                    if (g.sbs) {                                 // the planes are consecutive 64-channel runs: N = 2*BN spans both
                        const uint32_t idesc2 = idesc_bf16(128, 2u * (uint32_t)g.BN, 1, 1);
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, ks * stepB), idesc2, acc);
                            acc = 1u;
                        }
                    } else if (g.g_planes == 2) {
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, ks * stepB), idesc, acc);
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, b_plane + ks * stepB), idesc, 1u);
                            acc = 1u;
                        }
                    } else {
                        for (int ks = 0; ks < ksteps; ++ks) {
                            umma_f16(d_tmem, desc_add(a_desc, ks * stepA), desc_add(b_desc, ks * stepB), idesc, acc);
                            acc = 1u;
                        }
                    }
                    umma_commit(&a_empty[s]);
                    if (j == j_last) umma_commit(&b_empty[sbi]);
                    if (j == j_last && t == t_end - 1) umma_commit(done_bar);
                }
                __syncwarp();
            }
        }
    } else if (warp >= 2 && warp <= 5 && t_end > t_begin) {
        // ===== drain: partial[slice][row][n], rows = (tap, c) of this CTA's M tiles =====
        const int sub = warp & 3;
        mbar_wait_trap(done_bar, 0);
        tc_fence_after_sync();
        for (int j = 0; j < jn; ++j) {
            const int row = (g.flat == 1 ? ((fr * g.kw + j) * g.Cin + fcg * 128) : (mt0 + j) * 128) + sub * 32 + lane;
            for (int c0 = 0; c0 < g.BN; c0 += 32) {
                uint32_t r[32];
                tmem_ld_32x32(tmem_base + (uint32_t)j * acc_stride + (uint32_t)c0 + ((uint32_t)(sub * 32) << 16), r);
                if (g.sbs) {
                    uint32_t r2[32];
                    tmem_ld_32x32(tmem_base + (uint32_t)j * acc_stride + (uint32_t)g.BN + (uint32_t)c0 + ((uint32_t)(sub * 32) << 16), r2);
                    tmem_ld_wait();
#pragma unroll
                    for (int u = 0; u < 32; ++u) r[u] = __float_as_uint(__uint_as_float(r[u]) + __uint_as_float(r2[u]));
                } else {
                    tmem_ld_wait();
                }
                if (row < g.mrows) {
                    float* dst = partial + ((size_t)slice * g.mrows + row) * g.Cout + n0 + c0;
#pragma unroll
                    for (int u = 0; u < 32; u += 4)
                        if (n0 + c0 + u < g.Cout)
                            *reinterpret_cast<uint4*>(dst + u) = make_uint4(r[u], r[u + 1], r[u + 2], r[u + 3]);
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// Column sums of g: acc[tap][n] = SUM over the output pixels whose tap (r,s) reads inside the image, acc[taps][n] = SUM over all
// pixels (= the bias gradient).  Partials per block in a fixed order (deterministic).  Exits at once when neither is needed.
constexpr int CS_BLOCKS = 296;
__global__ void __launch_bounds__(256)
k_wgrad_colsums(const float* __restrict__ gf, float* __restrict__ cpart, int B, int Ho, int Wo, int Cout, int H, int W, int kh, int kw,
                int stride, int pad, const float* __restrict__ a_range, int a_bits, int want_db) {
    const int taps = kh * kw;
    if (!want_db) {
        const QParams pa = make_qparams_d((double)__ldg(a_range), (double)__ldg(a_range + 1), a_bits);
        if (pa.lo == 0.0f) return;
    }
    const int64_t rows = (int64_t)B * Ho;                                       // (b, ho) rows split over the blocks
    const int64_t r_begin = rows * blockIdx.x / gridDim.x, r_end = rows * (blockIdx.x + 1) / gridDim.x;
    for (int n = threadIdx.x; n < Cout; n += 256) {
        float acc[10];
#pragma unroll
        for (int k = 0; k < 10; ++k) acc[k] = 0.f;
        for (int64_t rw = r_begin; rw < r_end; ++rw) {
            const int ho = (int)(rw % Ho);
            uint32_t vh = 0;
            for (int r = 0; r < kh; ++r) {
                const int hi = ho * stride - pad + r;
                if (hi >= 0 && hi < H) vh |= 1u << r;
            }
            const float* src = gf + (rw * Wo) * Cout + n;
            for (int wo = 0; wo < Wo; ++wo) {
                const float v = __ldg(src + (int64_t)wo * Cout);
                acc[9] += v;
#pragma unroll
                for (int tp = 0; tp < 9; ++tp) {
                    if (tp < taps) {
                        const int r = tp / kw, s = tp - r * kw;
                        const int wi = wo * stride - pad + s;
                        if (((vh >> r) & 1u) && wi >= 0 && wi < W) acc[tp] += v;
                    }
                }
            }
        }
        for (int tp = 0; tp < taps; ++tp) cpart[((size_t)blockIdx.x * (taps + 1) + tp) * Cout + n] = acc[tp];
        cpart[((size_t)blockIdx.x * (taps + 1) + taps) * Cout + n] = acc[9];
    }
}

// dW[n][c][tap] = sa * SUM_slices partial[slice][tap*Cin + c][n] + la * SUM_blocks cpart[blk][tap][n];  db[n] = SUM_blocks cpart[blk][taps][n]
__global__ void __launch_bounds__(256)
k_wgrad_reduce(const float* __restrict__ partial, int slices, const float* __restrict__ cpart, int cblocks, float* __restrict__ dw,
               float* __restrict__ db, int Cout, int Cin, int taps, const float* __restrict__ a_range, int a_bits) {
    const int mrows = taps * Cin;
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const QParams pa = make_qparams_d((double)__ldg(a_range), (double)__ldg(a_range + 1), a_bits);
    if (idx < (int64_t)mrows * Cout) {
        const int n = (int)(idx % Cout), row = (int)(idx / Cout);
        const int tap = row / Cin, c = row - tap * Cin;
        double sq = 0.0;
        const float* src = partial + (size_t)row * Cout + n;
        const size_t pitch = (size_t)mrows * Cout;
        int s = 0;
        for (; s + 8 <= slices; s += 8) {                       // eight independent loads in flight; the sum keeps its slice order
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldcs(src + (size_t)(s + u) * pitch);
#pragma unroll
            for (int u = 0; u < 8; ++u) sq += (double)v[u];
        }
        for (; s < slices; ++s) sq += (double)__ldcs(src + (size_t)s * pitch);
        double v = (double)pa.scale * sq;
        if (pa.lo != 0.0f) {
            double sm = 0.0;
            for (int k = 0; k < cblocks; ++k) sm += (double)cpart[((size_t)k * (taps + 1) + tap) * Cout + n];
            v += (double)pa.lo * sm;
        }
        dw[((size_t)n * Cin + c) * taps + tap] = (float)v;
    }
    if (db && idx < Cout) {
        double sm = 0.0;
        for (int k = 0; k < cblocks; ++k) sm += (double)cpart[((size_t)k * (taps + 1) + taps) * Cout + idx];
        db[idx] = (float)sm;
    }
}

static int wgrad_geom2(WgradGeom& g, int64_t B, int64_t Ho, int64_t Wo, int64_t Cin, int64_t Cout, int kh, int kw, int sh, int sw, int ph,
                       int pw, int g_planes);
static void wgrad_try_flat(WgradGeom& g);
static int wgrad_geom(WgradGeom& g, int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad,
                      int g_planes) {
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (Ho < 1 || Wo < 1 || Cin % 16 != 0 || Cout % 8 != 0 || kh * kw > 9 || stride > 2) return 0;
    return wgrad_geom2(g, B, Ho, Wo, Cin, Cout, kh, kw, stride, stride, pad, pad, g_planes);
}
static int wgrad_geom2(WgradGeom& g, int64_t B, int64_t Ho, int64_t Wo, int64_t Cin, int64_t Cout, int kh, int kw, int sh, int sw, int ph,
                       int pw, int g_planes) {
    memset(&g, 0, sizeof(g));
    g.B = (int)B; g.Ho = (int)Ho; g.Wo = (int)Wo; g.Cin = (int)Cin; g.Cout = (int)Cout; g.kh = kh; g.kw = kw;
    g.sh = sh; g.sw = sw; g.ph = ph; g.pw = pw;
    g.g_planes = g_planes;
    g.RC = Cin % 64 == 0 ? 64 : (Cin % 32 == 0 ? 32 : 16);
    g.RN = Cout > 32 ? 64 : (Cout > 16 ? 32 : 16);
    g.BN = Cout >= 128 ? 128 : (int)((Cout + g.RN - 1) / g.RN) * g.RN;
    g.rpm = 128 / g.RC;
    g.total_runs = kh * kw * (int)Cin / g.RC;
    g.mtiles = (g.total_runs + g.rpm - 1) / g.rpm;
    g.mrows = kh * kw * (int)Cin;
    g.nblocks = (int)((Cout + g.BN - 1) / g.BN);
    static const bool sbs_off = getenv("QB_WGRAD_SBS") && getenv("QB_WGRAD_SBS")[0] == '0';
    // (BN = 64: an N = 64 MMA costs the same 75 cycles as an N = 128 one - profiles/umma_rate_probe_r02h.txt - so putting the two
    // planes of a 64-channel tile side by side halves the tensor time of the C_out = 64 layers)
    g.sbs = (!sbs_off && g_planes == 2 && (g.BN == 128 || g.BN == 64) && g.RN == 64) ? 1 : 0;
    // OFF by default: with the second issuing warp this kernel faults intermittently on the layer1 geometry (five accumulators,
    // odd count) inside a training step although every isolated run passes (the same scheme in k_conv_dgrad_flat, where both
    // warps walk the same stages, is stable and stays on); QB_WGRAD_MMA2=1 enables it for experiments.
    static const bool mma2_on = getenv("QB_WGRAD_MMA2") && getenv("QB_WGRAD_MMA2")[0] == '1';
    g.mma2 = mma2_on ? 1 : 0;
    static const bool uissue_on = getenv("QB_WGRAD_UISSUE") && getenv("QB_WGRAD_UISSUE")[0] == '1';
    g.uissue = (g.sbs && g.BN == 128 && uissue_on) ? 1 : 0;
    const int acc_stride = g.sbs ? 2 * g.BN : (g.BN < 32 ? 32 : g.BN);
    int jmax = 512 / acc_stride;
    if (jmax > WG_MAX_ACC) jmax = WG_MAX_ACC;
    g.mgroups = (g.mtiles + jmax - 1) / jmax;
    g.J = (g.mtiles + g.mgroups - 1) / g.mgroups;
    // pixel tile: P = TW*TH*TN pixels, P % 16 == 0, P <= pmax, covering the (Wo, Ho, B) grid with the least padding
    const int64_t b_bytes_per_px = (int64_t)g_planes * g.BN * 2;
    static const int pmax_env = getenv("QB_WGRAD_PMAX") ? atoi(getenv("QB_WGRAD_PMAX")) : 0;     // experiment: 64 / 96 / 128
    // 128 pixels per K block whenever two gradient stages and two x stages of that size fit: twice the MMAs per barrier round
    // trip of the issuing thread (measured 7 % on the wgrad total against 64-pixel blocks, profiles/bwd_sweep_r02n_pmax128.txt)
    const int pmax = pmax_env >= 16 ? pmax_env : (b_bytes_per_px * 128 * 2 + 2 * 128 * 256 <= 218 * 1024 ? 128 : 64);
    double best = -1.0;
    for (int tw = 1; tw <= 128 && tw <= Wo + 7; ++tw) {
        if (tw > Wo && tw != ((Wo + 7) / 8) * 8) continue;
        for (int th = 1; th * tw <= pmax && th <= Ho + 7; ++th) {
            if (th > Ho && th != ((Ho + 7) / 8) * 8) continue;
            for (int tn = 1; tn * th * tw <= pmax && tn <= B; tn *= 2) {
                const int P = tw * th * tn;
                if (P % 16 != 0) continue;
                const double cover = (double)((Wo + tw - 1) / tw) * tw * (double)((Ho + th - 1) / th) * th * (double)((B + tn - 1) / tn) * tn;
                const double eff = (double)Wo * Ho * B / cover + 1e-4 * P / 128.0;
                if (eff > best) { best = eff; g.TW = tw; g.TH = th; g.TN = tn; }
            }
        }
    }
    if (best < 0) return 0;
    g.P = g.TW * g.TH * g.TN;
    g.tiles_w = (int)((Wo + g.TW - 1) / g.TW);
    g.tiles_h = (int)((Ho + g.TH - 1) / g.TH);
    g.tiles_b = (int)((B + g.TN - 1) / g.TN);
    g.total_tiles = g.tiles_w * g.tiles_h * g.tiles_b;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const int units = g.mgroups * g.nblocks;
    g.slices = units >= sms ? 1 : sms / units;
    if (g.slices > g.total_tiles) g.slices = g.total_tiles;
    if (g.slices < 1) g.slices = 1;
    const size_t a_slot = (size_t)128 * g.P * 2, b_slot = (size_t)g_planes * g.BN * g.P * 2;
    // the gradient tile is the stream every accumulator waits for: keep three of them in flight when the x ring still gets
    // at least J + 2 slots (one pixel tile of x boxes plus a head start on the next)
    g.b_stages = 3;
    int as = (int)((220 * 1024 - 1024 - 512 - (size_t)g.b_stages * b_slot) / a_slot);
    if (as < g.J + 2 || as < 4) {
        g.b_stages = 2;
        as = (int)((220 * 1024 - 1024 - 512 - (size_t)g.b_stages * b_slot) / a_slot);
    }
    g.a_stages = as > WG_MAX_STAGES ? WG_MAX_STAGES : as;
    if (g.a_stages < 2) return 0;
    g.d_tiles_w = make_fastdiv((uint32_t)g.tiles_w);
    g.d_tiles_h = make_fastdiv((uint32_t)g.tiles_h);
    g.d_cin_runs = make_fastdiv((uint32_t)(Cin / g.RC));
    g.d_kw = make_fastdiv((uint32_t)kw);
    wgrad_try_flat(g);
    return 1;
}

// Switch a prepared geometry to the flat-row mode when it applies (see WgradGeom); leaves it untouched otherwise.
static void wgrad_try_flat(WgradGeom& g) {
    // Measured on B200 (profiles/bwd_sweep_r02f.txt against bwd_sweep_r02f_noflat.txt): the flat-row form moves 1.7x fewer
    // operand bytes and is nevertheless SLOWER (layer2 164 vs 147 us, layer3 255 vs 164 us) - the kernel is not bound by the
    // operand feed - so it stays an experiment behind QB_WGRAD_FLAT=1.
    static const bool off = !(getenv("QB_WGRAD_FLAT") && getenv("QB_WGRAD_FLAT")[0] == '1');
    static const bool on64 = getenv("QB_WGRAD_FLAT64") && getenv("QB_WGRAD_FLAT64")[0] == '1';
    const int Wp = g.Wo + g.kw - 1;
    const int taps = g.kh * g.kw, J2 = (taps + 1) / 2;
    if (on64 && g.sh == 1 && g.sw == 1 && g.Cin == 64 && g.g_planes == 2 && taps >= 2 && g.nblocks * g.BN == g.Cout &&
        J2 * (g.BN < 32 ? 32 : g.BN) <= 512 && J2 <= WG_MAX_ACC && Wp <= 128) {
        int bth = 0;
        double best = -1.0;
        for (int th = 1; th <= g.Ho && Wp * th <= 128; ++th) {
            const int P = (Wp * th + 15) & ~15;
            const double eff = (double)g.Wo * g.Ho / ((double)((g.Ho + th - 1) / th) * P);
            if (eff > best) { best = eff; bth = th; }
        }
        if (best >= 0.70) {
            WgradGeom f = g;
            f.sbs = 0; f.uissue = 0;
            f.flat = 2; f.Wp = Wp; f.TW = Wp; f.TH = bth; f.TN = 1;
            f.Pbox = Wp * bth;
            f.P = (f.Pbox + 15) & ~15;
            f.xrows = bth + g.kh - 1;
            const int max_off = (g.kh - 1) * Wp + g.kw - 1;
            int need = f.P + max_off + 1;
            if (need < f.xrows * Wp) need = f.xrows * Wp;
            f.arows = (need + 7) & ~7;
            f.tiles_w = 1;
            f.tiles_h = (g.Ho + bth - 1) / bth;
            f.tiles_b = g.B;
            f.total_tiles = f.tiles_h * f.tiles_b;
            f.mgroups = 1;
            f.J = J2;
            f.mtiles = J2;
            for (int j = 0; j < J2; ++j) {
                const int t1 = 2 * j, t2 = (2 * j + 1 < taps) ? 2 * j + 1 : 2 * j;
                const int o1 = (t1 / g.kw) * Wp + t1 % g.kw, o2 = (t2 / g.kw) * Wp + t2 % g.kw;
                f.foff[j] = o1;
                f.flbo[j] = o2 - o1;
            }
            int sms = qb_sm_count();
            if (sms <= 0) sms = 148;
            f.slices = f.nblocks >= sms ? 1 : sms / f.nblocks;
            if (f.slices > f.total_tiles) f.slices = f.total_tiles;
            const size_t a_slot = (size_t)f.arows * 128, b_slot = (size_t)2 * f.BN * f.P * 2;
            f.b_stages = 3;
            int as = (int)((220 * 1024 - 1024 - 512 - (size_t)f.b_stages * b_slot) / a_slot);
            if (as >= 2) {
                f.a_stages = as > 4 ? 4 : as;
                f.d_tiles_w = make_fastdiv(1u);
                f.d_tiles_h = make_fastdiv((uint32_t)f.tiles_h);
                g = f;
                return;
            }
        }
    }
    if (off || g.sh != 1 || g.sw != 1 || g.kw < 2 || g.kw > 4 || g.Cin % 128 != 0 || g.g_planes != 2 || g.RN != 64 || g.BN != 128 ||
        Wp > 128)
        return;
    // pixel tile {Wp, TH, TN}: K extent P = ceil16(Wp*TH*TN) <= 128, fewest padded K steps per useful pixel
    double best = -1.0;
    int bth = 0, btn = 0;
    for (int th = 1; th <= g.Ho && Wp * th <= 128; ++th)
        for (int tn = 1; tn <= g.B && Wp * th * tn <= 128; tn *= 2) {
            if (tn > 1 && th != g.Ho) continue;                          // several images only when one holds a whole image
            const int P = (Wp * th * tn + 15) & ~15;
            const double cost = (double)((g.Ho + th - 1) / th) * ((g.B + tn - 1) / tn) * P;
            const double eff = (double)g.Wo * g.Ho * g.B / cost;
            if (eff > best) { best = eff; bth = th; btn = tn; }
        }
    if (best < 0.70) return;
    WgradGeom f = g;
    f.sbs = 0; f.uissue = 0;
    f.flat = 1; f.Wp = Wp; f.TW = Wp; f.TH = bth; f.TN = btn;
    f.Pbox = Wp * bth * btn;
    f.P = (f.Pbox + 15) & ~15;
    f.arows = f.P + 8;
    f.tiles_w = 1;
    f.tiles_h = (g.Ho + bth - 1) / bth;
    f.tiles_b = (g.B + btn - 1) / btn;
    f.total_tiles = f.tiles_h * f.tiles_b;
    f.mgroups = g.kh * (g.Cin / 128);
    f.J = g.kw;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const int units = f.mgroups * f.nblocks;
    f.slices = units >= sms ? 1 : sms / units;
    if (f.slices > f.total_tiles) f.slices = f.total_tiles;
    if (f.slices < 1) f.slices = 1;
    const size_t a_slot = (size_t)2 * f.arows * 128, b_slot = (size_t)2 * f.BN * f.P * 2;
    f.b_stages = 2;
    int as = (int)((220 * 1024 - 1024 - 512 - (size_t)f.b_stages * b_slot) / a_slot);
    if (as < 2) return;
    f.a_stages = as > 3 ? 3 : as;
    f.d_tiles_w = make_fastdiv(1u);
    f.d_tiles_h = make_fastdiv((uint32_t)f.tiles_h);
    g = f;
}

// tensor maps of the gradient planes + the launch of k_conv_wgrad for a prepared geometry / x operand map
static int wgrad_core(const WgradGeom& g, const CUtensorMap& tx, const uint16_t* g_hi, const uint16_t* g_lo, float* partial, cudaStream_t st) {
    CUtensorMap tg0, tg1;
    const uint64_t dims[4] = {(uint64_t)g.Cout, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)g.B};
    const uint64_t str[3] = {(uint64_t)g.Cout * 2, (uint64_t)g.Wo * g.Cout * 2, (uint64_t)g.Ho * g.Wo * g.Cout * 2};
    const uint32_t box[4] = {(uint32_t)g.RN, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
    int rc = qb_make_tmap(&tg0, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, g_hi, dims, str, box, nullptr, g.RN * 2);
    if (rc) return rc;
    rc = qb_make_tmap(&tg1, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, g_lo ? g_lo : g_hi, dims, str, box, nullptr, g.RN * 2);
    if (rc) return rc;
    const size_t a_slot = g.flat == 2 ? (size_t)g.arows * 128 : g.flat ? (size_t)2 * g.arows * 128 : (size_t)128 * g.P * 2;
    const size_t smem = 1024 + (size_t)g.a_stages * a_slot + (size_t)g.b_stages * g.g_planes * g.BN * g.P * 2 + 512;
    QB_CUDA(cudaFuncSetAttribute(k_conv_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = (unsigned)(g.slices * g.mgroups * g.nblocks);
    k_conv_wgrad<<<grid, WG_THREADS, smem, st>>>(tx, tg0, tg1, partial, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" size_t qb_conv_wgrad_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad) {
    WgradGeom g;
    if (!wgrad_geom(g, B, H, W, Cin, Cout, kh, kw, stride, pad, 2)) return 0;
    return ((size_t)g.slices * g.mrows * Cout + (size_t)CS_BLOCKS * (kh * kw + 1) * Cout) * 4 + 256;
}

/* dW (and db) of y = conv2d(x_hat, w) + b from
 *   g_hi / g_lo : bf16 hi / lo planes [B,Ho,Wo,Cout] (channels-last) of the output gradient (g_lo may be null: one plane)
 *   g_f32       : the same gradient in fp32 (bias gradient and the zero-padding term when the input range does not start at 0)
 *   x16         : bf16 integer codes q_a [B,H,W,Cin] (channels-last) of the input quantizer with its device range
 * -> dw fp32 [Cout][Cin][kh][kw], db fp32 [Cout] (optional).  Deterministic.                                              */
extern "C" int qb_conv_wgrad(const uint16_t* g_hi, const uint16_t* g_lo, const float* g_f32, const uint16_t* x16, float* dw, float* db,
                             int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad,
                             const float* a_range, int a_bits, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(g_hi && g_f32 && x16 && dw && a_range && ws, "qb_conv_wgrad: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0, "qb_conv_wgrad: bad shape");
    WgradGeom g;
    if (!wgrad_geom(g, B, H, W, Cin, Cout, kh, kw, stride, pad, g_lo ? 2 : 1) ||
        ((reinterpret_cast<uintptr_t>(g_hi) | reinterpret_cast<uintptr_t>(g_lo) | reinterpret_cast<uintptr_t>(x16) |
          reinterpret_cast<uintptr_t>(ws)) & 15) != 0) {
        qb_set_error("qb_conv_wgrad: unsupported geometry (need Cin%%16==0, Cout%%8==0, kh*kw<=9, stride<=2, 16-byte aligned "
                     "tensors): Cin=%lld Cout=%lld k=%dx%d stride=%d", (long long)Cin, (long long)Cout, kh, kw, stride);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(B * H * W < 2147483647LL, "qb_conv_wgrad: pixel count exceeds int32");
    const size_t need = qb_conv_wgrad_ws_bytes(B, H, W, Cin, Cout, kh, kw, stride, pad);
    if (ws_bytes < need) {
        qb_set_error("qb_conv_wgrad: workspace %zu < %zu bytes", ws_bytes, need);
        return QB_ERR_WORKSPACE;
    }
    float* partial = static_cast<float*>(ws);
    float* cpart = partial + (size_t)g.slices * g.mrows * Cout;
    cudaStream_t st = qb_stream(stream);
    CUtensorMap tx;
    {
        const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {(uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)H * W * Cin * 2};
        const uint32_t box[4] = {(uint32_t)g.RC, (uint32_t)(g.TW * stride), (uint32_t)((g.flat == 2 ? g.xrows : g.TH) * stride),
                                 (uint32_t)g.TN};
        const uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
        int rc = qb_make_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, x16, dims, str, box, es, g.RC * 2);
        if (rc) return rc;
    }
    {
        int rc = wgrad_core(g, tx, g_hi, g_lo, partial, st);
        if (rc) return rc;
    }
    k_wgrad_colsums<<<CS_BLOCKS, 256, 0, st>>>(g_f32, cpart, (int)B, g.Ho, g.Wo, (int)Cout, (int)H, (int)W, kh, kw, stride, pad, a_range,
                                              a_bits, db != nullptr);
    QB_LAUNCH_CHECK();
    const int64_t outs = (int64_t)g.mrows * Cout;
    k_wgrad_reduce<<<(unsigned)((outs + 255) / 256), 256, 0, st>>>(partial, g.slices, cpart, CS_BLOCKS, dw, db, (int)Cout, (int)Cin,
                                                                 kh * kw, a_range, a_bits);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// Weight gradient of the FIRST layer (3 input channels; 7x7 / stride 2 / pad 3 of ResNet, 3x3 / stride 2 / pad 1 of MobileNet)
// on the same kernel.  A filter tap is 3 channels = 6 bytes of bf16 codes - no TMA box - but a filter ROW is a contiguous
// window of the NHWC image.  The codes are expanded once into xs[B][H][W + 2*PL][4] bf16 (channel 3 = 1.0 on real pixels, 0 in
// the PL zero columns added left and right), and the kernel is pointed at a VIRTUAL operand through an OVERLAPPING tensor
// map: "pixel" wo of a virtual row is the window of WN pixels x 4 channels (RC = 4*WN = 32 or 16 "channels") that starts at
// stored column 2*wo, i.e. dims {RC, Wo, H, B} with a W stride of two stored pixels = 16 bytes.  The convolution the kernel
// then sees has kw = 1, kh taps, stride (2 in h, 1 in w), pad (pad in h, 0 in w): row (r, s', c') of D is
//     c' < 3 : SUM_px g * q[2ho + r - pad][2wo + s' - PL][c']           -> dW[n][c'][r][s' - (PL - pad)]
//     c' = 3 : SUM_px g * [that input pixel is inside the image]        -> the zero-padding term of la (image inputs have la != 0)
// (h bounds by the TMA's zero fill, w bounds by the physical zero columns).  Replaces the thread-built tf32 kernel of
// conv_stem.cu (1.2 ms, tf32 operand rounding) for these geometries.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_stem_expand16(const uint8_t* __restrict__ codes, uint2* __restrict__ xs, int64_t rows, int W, int PL) {
    // one thread per stored pixel (4 bf16 = 8 bytes)
    const int Ws = W + 2 * PL;
    const int64_t total = rows * Ws;
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
        const int64_t row = i / Ws;
        const int col = (int)(i - row * Ws) - PL;
        uint2 v = make_uint2(0u, 0u);
        if (col >= 0 && col < W) {
            const uint8_t* p = codes + (row * W + col) * 3;
            const uint32_t c0 = __float_as_uint((float)p[0]) >> 16, c1 = __float_as_uint((float)p[1]) >> 16,
                           c2 = __float_as_uint((float)p[2]) >> 16;
            v.x = c0 | (c1 << 16);
            v.y = c2 | (0x3F80u << 16);                     // bf16 1.0
        }
        xs[i] = v;
    }
}

// dW[n][c][r][s] = sa * SUM_slices D[(r, s + PL - pad, c)][n] + la * SUM_slices D[(r, s + PL - pad, 3)][n]
__global__ void __launch_bounds__(256)
k_stem_wgrad16_reduce(const float* __restrict__ partial, int slices, int mrows, float* __restrict__ dw, int Cout, int kh, int kw, int WN,
                      int shift, const float* __restrict__ a_range, int a_bits) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= Cout * 3 * kh * kw) return;
    const int n = idx % Cout;
    const int t = idx / Cout;                    // (r, s, c)
    const int c = t % 3, s = (t / 3) % kw, r = t / (3 * kw);
    const QParams pa = make_qparams_d((double)__ldg(a_range), (double)__ldg(a_range + 1), a_bits);
    const int rq = (r * WN + s + shift) * 4 + c, rm = (r * WN + s + shift) * 4 + 3;
    double sq = 0.0, sm = 0.0;
    for (int k = 0; k < slices; ++k) {
        sq += (double)partial[((size_t)k * mrows + rq) * Cout + n];
        sm += (double)partial[((size_t)k * mrows + rm) * Cout + n];
    }
    dw[((size_t)n * 3 + c) * (kh * kw) + r * kw + s] = (float)((double)pa.scale * sq + (double)pa.lo * sm);
}

static int stem16_geom(WgradGeom& g, int& WN, int& PL, int64_t B, int64_t H, int64_t W, int64_t Cout, int k, int stride, int pad) {
    if (stride != 2 || !((k == 7 && pad == 3) || (k == 3 && pad == 1)) || Cout % 8 != 0 || W % 2 != 0) return 0;
    WN = k == 7 ? 8 : 4;
    PL = k == 7 ? 4 : 2;                          // window of output wo starts at pixel 2*wo - PL (stored column 2*wo)
    const int64_t Ho = (H + 2 * pad - k) / stride + 1, Wo = (W + 2 * pad - k) / stride + 1;
    if (Ho < 1 || Wo < 1) return 0;
    return wgrad_geom2(g, B, Ho, Wo, 4 * WN, Cout, k, 1, 2, 1, pad, 0, 2);
}

extern "C" size_t qb_conv_stem_wgrad16_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t Cout, int k, int stride, int pad) {
    WgradGeom g;
    int WN, PL;
    if (!stem16_geom(g, WN, PL, B, H, W, Cout, k, stride, pad)) return 0;
    const size_t xs = (size_t)B * H * (W + 2 * PL) * 8 + 256;
    return xs + (size_t)g.slices * g.mrows * Cout * 4 + 256;
}

/* dW of the first layer from the bf16 hi / lo planes of its output gradient and the uint8 input codes [B,H,W,3]:
 * dw fp32 [Cout][3][k][k].  0 from qb_conv_stem_wgrad16_ws_bytes = geometry not taken (use qb_conv_stem_wgrad). */
extern "C" int qb_conv_stem_wgrad16(const uint16_t* g_hi, const uint16_t* g_lo, const uint8_t* x_codes, float* dw, int64_t B, int64_t H,
                                    int64_t W, int64_t Cout, int k, int stride, int pad, const float* a_range, int a_bits, void* ws,
                                    size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(g_hi && g_lo && x_codes && dw && a_range && ws, "qb_conv_stem_wgrad16: null pointer");
    WgradGeom g;
    int WN, PL;
    if (!stem16_geom(g, WN, PL, B, H, W, Cout, k, stride, pad) ||
        ((reinterpret_cast<uintptr_t>(g_hi) | reinterpret_cast<uintptr_t>(g_lo) | reinterpret_cast<uintptr_t>(ws)) & 15) != 0) {
        qb_set_error("qb_conv_stem_wgrad16: unsupported geometry (3 channels, 7x7/2/3 or 3x3/2/1, even W, Cout%%8==0)");
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8, "qb_conv_stem_wgrad16: codes are 8-bit");
    const size_t need = qb_conv_stem_wgrad16_ws_bytes(B, H, W, Cout, k, stride, pad);
    if (ws_bytes < need) {
        qb_set_error("qb_conv_stem_wgrad16: workspace %zu < %zu bytes", ws_bytes, need);
        return QB_ERR_WORKSPACE;
    }
    cudaStream_t st = qb_stream(stream);
    const int64_t Ws = W + 2 * PL;
    const size_t xs_bytes = (((size_t)B * H * Ws * 8) + 255) & ~(size_t)255;
    uint16_t* xs = static_cast<uint16_t*>(ws);
    float* partial = reinterpret_cast<float*>(static_cast<char*>(ws) + xs_bytes);
    {
        const int64_t total = B * H * Ws;
        int sms = qb_sm_count();
        if (sms <= 0) sms = 148;
        int64_t want = (total + 255) / 256, cap = (int64_t)sms * 16;
        k_stem_expand16<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(x_codes, reinterpret_cast<uint2*>(xs), B * H, (int)W, PL);
        QB_LAUNCH_CHECK();
    }
    CUtensorMap tx;
    {   // virtual operand: {RC = 4*WN "channels", Wo windows (stride 2 stored pixels = 16 B: overlapping), H rows, B}
        const uint64_t dims[4] = {(uint64_t)(4 * WN), (uint64_t)g.Wo, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {16, (uint64_t)Ws * 8, (uint64_t)H * Ws * 8};
        const uint32_t box[4] = {(uint32_t)g.RC, (uint32_t)g.TW, (uint32_t)(g.TH * 2), (uint32_t)g.TN};
        const uint32_t es[4] = {1, 1, 2, 1};
        int rc = qb_make_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, xs, dims, str, box, es, g.RC * 2);
        if (rc) return rc;
    }
    int rc = wgrad_core(g, tx, g_hi, g_lo, partial, st);
    if (rc) return rc;
    const int outs = (int)Cout * 3 * k * k;
    k_stem_wgrad16_reduce<<<(outs + 255) / 256, 256, 0, st>>>(partial, g.slices, g.mrows, dw, (int)Cout, k, k, WN, PL - pad, a_range, a_bits);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// Hardware probe (tools/umma_shift_probe.py): may a UMMA shared-memory descriptor START at a row that is not a multiple of
// the 8-row swizzle atom?  mn = 0: K-major A (rows = M), D[128 x 64] = A[r0 .. r0+127][0..63] * B[64][64]^T;
// mn = 1: both operands MN-major (rows = K), D[128 x 64] = SUM_{k<64} A[r0+k][0..127] * B[r0+k][0..63].
// use_boff: put ((start >> 7) & 7) into the descriptor's base-offset field.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1)
k_umma_shift_probe(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, float* __restrict__ out,
                   int mn, int r0, int use_boff) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                       // K-major: 144 rows x 128 B; MN-major: 2 boxes of [80][64 ch]
    uint8_t* sB = smem + 32768;               // K-major: 64 rows x 128 B;  MN-major: [80][64 ch]
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 49152);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 2);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<64>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) {
        if (!mn) {
            mbar_arrive_expect_tx(&bar[0], 144 * 128 + 64 * 128);
            tma_load_2d(sA, &tmap_a, &bar[0], 0, 0);
            tma_load_2d(sB, &tmap_b, &bar[0], 0, 0);
        } else {
            mbar_arrive_expect_tx(&bar[0], 3 * 80 * 128);
            tma_load_2d(sA, &tmap_a, &bar[0], 0, 0);
            tma_load_2d(sA + 80 * 128, &tmap_a, &bar[0], 64, 0);
            tma_load_2d(sB, &tmap_b, &bar[0], 0, 0);
        }
        mbar_wait_trap(&bar[0], 0);
        tc_fence_after_sync();
        const uint32_t a = smem_u32(sA) + (uint32_t)r0 * 128u, b = smem_u32(sB) + (mn ? (uint32_t)r0 * 128u : 0u);
        for (int k = 0; k < 4; ++k) {
            uint64_t da, db;
            if (!mn) {
                da = desc_kmajor(a + k * 32, 128);
                db = desc_kmajor(b + k * 32, 128);
            } else if (mn == 2) {         // overlapping column blocks: M rows 64..127 = the SAME 64-channel strip, use_boff rows later
                da = desc_mnmajor(a + k * 16 * 128, 128, (uint32_t)use_boff * 128u);
                db = desc_mnmajor(b + k * 16 * 128, 128, 80 * 128);
            } else {
                da = desc_mnmajor(a + k * 16 * 128, 128, 80 * 128);
                db = desc_mnmajor(b + k * 16 * 128, 128, 80 * 128);
            }
            if (use_boff && mn != 2) {
                da |= (uint64_t)(((a + (mn ? k * 2048 : 0)) >> 7) & 7) << 49;
                if (mn) db |= (uint64_t)(((b + k * 2048) >> 7) & 7) << 49;
            }
            umma_f16(tmem, da, db, idesc_bf16(128, 64, mn != 0, mn != 0), k > 0);
        }
        umma_commit(&bar[1]);
    }
    __syncthreads();
    mbar_wait_trap(&bar[1], 0);
    tc_fence_after_sync();
    for (int c0 = 0; c0 < 64; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem + (uint32_t)c0 + ((uint32_t)(warp * 32) << 16), r);
        tmem_ld_wait();
        for (int u = 0; u < 32; ++u) out[(warp * 32 + lane) * 64 + c0 + u] = __uint_as_float(r[u]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<64>(tmem);
}

// ------------------------------------------------------------------------------------------------
// tensor-pipe rate probe (tools/umma_rate_probe.py): one CTA per SM issues `n` back-to-back tcgen05.mma kind::f16
// (M = 128, N, K = 16) on zero-filled shared memory - no TMA, no waits - and reports the cycles from the first issue to the
// completion of the last.  mode bit 0: A MN-major, bit 1: B MN-major, bit 2: descriptors advance through a 16 KB window (as
// a K loop does) instead of re-reading one address, bit 3: A starts 3 rows into a swizzle atom.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1) k_umma_rate_probe(long long* __restrict__ cycles, int n, int N, int mode) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 128 * 1024);          // A window 64 KB, B window 64 KB
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    for (uint32_t i = threadIdx.x; i < 128 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (threadIdx.x < 32) tmem_alloc<256>(tmem_slot);
    fence_proxy_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) {
        const bool a_mn = mode & 1, b_mn = mode & 2, walk = mode & 4;
        const uint32_t a0 = smem_u32(smem) + ((mode & 8) ? 3u * 128u : 0u), b0 = smem_u32(smem + 64 * 1024);
        const uint32_t idesc = idesc_bf16(128, (uint32_t)N, a_mn, b_mn);
        const uint64_t da0 = a_mn ? desc_mnmajor(a0, 128, 16 * 1024) : desc_kmajor(a0, 128);
        const uint64_t db0 = b_mn ? desc_mnmajor(b0, 128, 16 * 1024) : desc_kmajor(b0, 128);
        const uint32_t stepA = a_mn ? 16u * 128u : 32u, stepB = b_mn ? 16u * 128u : 32u;     // one K = 16 step
        const uint32_t wrapA = a_mn ? 8u : 4u, wrapB = b_mn ? 8u : 4u;
        const long long t0 = clock64();
        for (int i = 0; i < n; ++i) {
            const uint32_t ka = walk ? (uint32_t)i % wrapA : 0u, kb = walk ? (uint32_t)i % wrapB : 0u;
            umma_f16(tmem, desc_add(da0, ka * stepA), desc_add(db0, kb * stepB), idesc, i > 0);
        }
        umma_commit(bar);
        mbar_wait_trap(bar, 0);
        cycles[blockIdx.x] = clock64() - t0;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<256>(tmem);
}

extern "C" int qb_debug_umma_rate(long long* cycles, int blocks, int n, int N, int mode, void* stream) {
    QB_CHECK_ARG(cycles && blocks > 0 && n > 0 && N >= 16 && N <= 256 && N % 16 == 0, "qb_debug_umma_rate: bad argument");
    QB_CUDA(cudaFuncSetAttribute(k_umma_rate_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 132 * 1024));
    k_umma_rate_probe<<<blocks, 128, 132 * 1024, qb_stream(stream)>>>(cycles, n, N, mode);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* a: bf16 [144][64] (mn = 0) or [80][128] (mn = 1); b: bf16 [64][64] (mn = 0) or [80][64] (mn = 1); out fp32 [128][64]. */
extern "C" int qb_debug_umma_shift(const uint16_t* a, const uint16_t* b, float* out, int mn, int r0, int use_boff, void* stream) {
    QB_CHECK_ARG(a && b && out && r0 >= 0 && r0 <= 16, "qb_debug_umma_shift: bad argument");
    CUtensorMap ta, tb;
    if (!mn) {
        const uint64_t da[2] = {64, 144}, sa[1] = {128}, db[2] = {64, 64};
        const uint32_t ba[2] = {64, 144}, bb[2] = {64, 64};
        int rc = qb_make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a, da, sa, ba, nullptr, 128);
        if (rc) return rc;
        rc = qb_make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, b, db, sa, bb, nullptr, 128);
        if (rc) return rc;
    } else {
        const uint64_t da[2] = {128, 80}, sa[1] = {256}, db[2] = {64, 80}, sb[1] = {128};
        const uint32_t bx[2] = {64, 80};
        int rc = qb_make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, a, da, sa, bx, nullptr, 128);
        if (rc) return rc;
        rc = qb_make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, b, db, sb, bx, nullptr, 128);
        if (rc) return rc;
    }
    QB_CUDA(cudaFuncSetAttribute(k_umma_shift_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 51200));
    k_umma_shift_probe<<<1, 128, 51200, qb_stream(stream)>>>(ta, tb, out, mn, r0, use_boff);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
