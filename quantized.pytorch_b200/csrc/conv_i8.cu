// conv_i8.cu - QConv2d's forward contraction as an int8 tcgen05 implicit GEMM over channels-last uint8 codes.
//
// Reference: F.conv2d(x_hat, w_hat, b_hat) on the DEQUANTISED fp32 tensors (models/modules/quantize.py:143-148, :343).
// x_hat = qa*sa + la (per tensor), w_hat[n] = qw*sw[n] + lw[n] (per output channel); zero padding contributes TRUE
// zeros, which in this min-anchored code are not "code 0".  With V(m) the set of in-bounds filter taps of output pixel m
//   out[m,n] = sa*sw[n]*SUM_{V,c} qa qw + sa*lw[n]*SUM_{V,c} qa + la*sw[n]*SUM_{V,c} qw[n] + la*lw[n]*|V|*Cin + bias[n]
// TMA's zero fill makes out-of-bounds taps contribute 0 to the first (integer GEMM) term exactly; the other terms use
//   P[b,h,w]  = SUM_c qa[b,h,w,c]     (per-pixel code sums, qb_rowsum_u8)      -> SUM_{V,c} qa = box filter of P
//   T[n][tap] = SUM_c qw[n,tap,c]     (per-tap weight code sums, qb_fakequant_rows_conv)
// and vanish when la == 0, i.e. for every post-ReLU input.
//
// GEMM view: M = B*Ho*Wo, N = Cout, K = kh*kw*Cin.  One CTA computes a tile of TN x TH x TW output pixels (<= 128 rows
// of the UMMA M=128 tile; unused rows are never stored) x BN output channels.  For filter tap (r,s) and channel block cb
// the A operand is ONE 4-D TMA box {BK channels, TW, TH, TN} of the NHWC code tensor at (cb*BK, wo0*stride-pad+s,
// ho0*stride-pad+r, b0) with element strides (1, stride, stride, 1): no im2col buffer exists anywhere.
#include <stdlib.h>

#include "tc_common.cuh"

constexpr int CV_THREADS = 320;      // TMA warp, MMA warp, 2 x 4 epilogue warps
constexpr int CV_MAX_STAGES = 16;
constexpr int CV_MAX_TAPS = 9;

struct ConvGeom {
    int B, H, W, Cin, Ho, Wo, Cout;
    int kh, kw, stride, pad;
    int TW, TH, TN;             // tile of output pixels (TN*TH*TW <= 128)
    int tiles_w, tiles_h, tiles_b;
    int BK;                     // bytes (= channels) of K per pipeline stage: 32, 64 or 128
    int stages;
    int halo;                   // 1: 3x3 / stride 1 / Cin == BK, Cout == BN: column-shifted halo boxes + resident weights
    FastDiv d_tiles_m, d_tiles_w, d_tiles_h;   // tile-index decoding without integer division
};

// tile t -> (channel tile tn, first image b0, first output row ho0, first output column wo0)
__device__ __forceinline__ void conv_decode_tile(int t, const ConvGeom& g, int tiles_m, int& tn, int& b0, int& ho0, int& wo0) {
    tn = (int)fdiv((uint32_t)t, g.d_tiles_m);
    uint32_t tm = (uint32_t)(t - tn * tiles_m);
    uint32_t q = fdiv(tm, g.d_tiles_w);
    wo0 = (int)(tm - q * (uint32_t)g.tiles_w) * g.TW;
    tm = q;
    q = fdiv(tm, g.d_tiles_h);
    ho0 = (int)(tm - q * (uint32_t)g.tiles_h) * g.TH;
    b0 = (int)q * g.TN;
}

struct ConvEpilogue {
    const float* a_range;       // device {min,max} of the input quantizer (QB_SCALE_DOUBLE)
    int a_bits;
    const float* w_row_min;     // per output channel (QB_SCALE_FLOAT)
    const float* w_row_max;
    int w_bits;
    const int32_t* boxsum;      // S[B*Ho*Wo] = SUM over the in-bounds taps of the per-pixel code sums (k_conv_boxsum)
    const int32_t* tapsum;      // T[Cout][kh*kw]
    const float* bias;          // fake-quantized bias or nullptr
    // optional: QuantMeasure statistics of the OUTPUT (what the following RangeBN's quantize_input needs,
    // quantize.py:223-235), produced while the tile is in registers; same block / workspace layout as qb_measure_stats
    double* stats;              // double[8] or nullptr
    unsigned int* st_counter;
    double* st_psum;
    double* st_psq;
    unsigned int* st_row_max;
    unsigned int* st_row_min;
};

__device__ __forceinline__ uint64_t umma_smem_desc(uint32_t smem_addr, int bk) {
    // K-major rows of `bk` bytes with the bk-byte swizzle: 8-row groups are 8*bk bytes apart
    const uint64_t layout = bk == 128 ? 2 : (bk == 64 ? 4 : 6);
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8 * bk) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= layout << 61;
    return d;
}

constexpr int CV_SLAB_BYTES = 128 * 128;             // one [128 rows][32 fp32] output slab (128-byte swizzle)
constexpr int CV_STAGING_BYTES = 4 * CV_SLAB_BYTES;  // two slabs per epilogue group

__device__ __forceinline__ void epi_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void epi_bar_all() { asm volatile("bar.sync 3, 256;" ::: "memory"); }
__device__ __forceinline__ float4 lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ void lds_2x64(uint32_t addr, uint64_t& a, uint64_t& b) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(a), "=l"(b) : "r"(addr));
}
__device__ __forceinline__ uint64_t lds_64(uint32_t addr) {
    uint64_t a;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(a) : "r"(addr));
    return a;
}

// Persistent kernel: gridDim.x CTAs (one per SM) walk the tile list t = blockIdx.x, += gridDim.x; tiles are ordered with
// the output-channel tile slowest so a CTA's per-channel coefficients rarely change.  Three asynchronous roles:
//   warp 0   TMA producer: keeps the shared-memory ring full ACROSS tile boundaries
//   warp 1   MMA issuer  : accumulates tile j into TMEM buffer j&1 (2 x BN columns allocated)
//   warps 2-9 epilogue   : two groups of four warps (one warp per TMEM lane quarter) take alternate 32-column slabs of
//                          buffer j&1: tcgen05.ld, dequantise, park the [rows][32] slab in a swizzled shared-memory
//                          staging buffer and hand it to a TMA store (coalesced 128-byte rows, ragged tile edges clipped
//                          by the hardware) - overlapped with the MMAs of tile j+1
template <int BN>
__global__ void __launch_bounds__(CV_THREADS, 1)
k_conv_i8(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w,
          const __grid_constant__ CUtensorMap tmap_o, ConvGeom g, ConvEpilogue ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    // halo mode: a stage holds the kw column-shifted boxes {BK, TW, TH+kh-1} of ONE tile (tap (r,s) = box s advanced by
    // r*TW rows = whole 8-row swizzle atoms since TW % 8 == 0); all kh*kw weight slabs stay resident behind the ring
    const uint32_t hbox_bytes = (uint32_t)(g.BK * g.TW * (g.TH + g.kh - 1));
    const uint32_t a_bytes = 128u * g.BK, b_bytes = (uint32_t)BN * g.BK;
    const uint32_t stage_bytes = g.halo ? (uint32_t)g.kw * hbox_bytes : a_bytes + b_bytes;
    uint8_t* wres = smem + (size_t)g.stages * stage_bytes;
    const uint32_t wres_bytes = g.halo ? (uint32_t)(g.kh * g.kw) * b_bytes : 0u;
    float* staging = reinterpret_cast<float*>(wres + wres_bytes);                               // 1024-byte aligned
    uint8_t* tail = reinterpret_cast<uint8_t*>(staging) + CV_STAGING_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
    uint64_t* empty_bar = full_bar + CV_MAX_STAGES;
    uint64_t* tfull_bar = empty_bar + CV_MAX_STAGES;           // [2] accumulator buffer ready
    uint64_t* tempty_bar = tfull_bar + 2;                      // [2] accumulator buffer drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    uint64_t* wfull_bar = reinterpret_cast<uint64_t*>(tail + 296);   // halo mode: resident weights have landed
    // per output channel n: c0 = sa*sw[n], c1 = sa*lw'[n], c2 = bias[n], w = la*(sw*Tall + lw'*K).  c0 / c1 are stored per
    // channel PAIR {c0[2p], c0[2p+1], c1[2p], c1[2p+1]} so that one 128-bit load feeds two packed FMAs
    float* s_pair = reinterpret_cast<float*>(tail + 320);     // [BN/2][4]
    float* s_c2 = s_pair + 2 * BN;                             // [BN]
    float* s_w = s_c2 + BN;                                    // [BN]
    float* s_c3 = s_w + BN;                                    // la*sw[n]      (border pixels of inputs with min != 0)
    float* s_c4 = s_c3 + BN;                                   // la*lw'[n]
    int32_t* s_tap = reinterpret_cast<int32_t*>(s_c4 + BN);   // [BN][taps]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int taps = g.kh * g.kw;
    const int cblocks = (g.Cin + g.BK - 1) / g.BK;                // Cin % BK != 0 (Cin = 16): TMA zero-fills the rest of the row
    const int num_kb = taps * cblocks;
    const int tiles_m = g.tiles_w * g.tiles_h * g.tiles_b;
    const int tiles_n = (g.Cout + BN - 1) / BN;
    const int total_tiles = tiles_m * tiles_n;
    constexpr uint32_t TMEM_COLS = 2 * BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_w);
        tma_prefetch_desc(&tmap_o);
        for (int s = 0; s < g.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 8);                      // one arrival per epilogue warp
        }
        mbar_init(wfull_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        const bool lead = elect_one();
        if (lead && g.halo) {
            mbar_arrive_expect_tx(wfull_bar, wres_bytes);
            for (int tap = 0; tap < taps; ++tap) tma_load_3d(wres + (size_t)tap * b_bytes, &tmap_w, wfull_bar, 0, tap, 0);
            RingPos rp;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, rp.next(g.stages)) {
                int tn, b0, ho0, wo0;                           // tiles_n == 1 in halo mode: tn == 0
                conv_decode_tile(t, g, tiles_m, tn, b0, ho0, wo0);
                const int s = rp.s;
                const uint32_t ph = rp.ph;
                mbar_wait(&empty_bar[s], ph ^ 1);
                mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
                uint8_t* sa = smem + (size_t)s * stage_bytes;
                for (int s_ = 0; s_ < g.kw; ++s_)
                    tma_load_4d(sa + (size_t)s_ * hbox_bytes, &tmap_x, &full_bar[s], 0, wo0 - g.pad + s_, ho0 - g.pad, b0);
            }
        } else if (lead) {
            RingPos rp;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                int tn, b0, ho0, wo0;
                conv_decode_tile(t, g, tiles_m, tn, b0, ho0, wo0);
                const int n0 = tn * BN;
                for (int r = 0; r < g.kh; ++r) {
                    for (int s_ = 0; s_ < g.kw; ++s_) {
                        const int tap = r * g.kw + s_;
                        for (int cb = 0; cb < cblocks; ++cb, rp.next(g.stages)) {
                            const int s = rp.s;
                            const uint32_t ph = rp.ph;
                            mbar_wait(&empty_bar[s], ph ^ 1);
                            mbar_arrive_expect_tx(&full_bar[s], (uint32_t)(g.TN * g.TH * g.TW * g.BK) + b_bytes);
                            uint8_t* sa = smem + (size_t)s * stage_bytes;
                            tma_load_4d(sa, &tmap_x, &full_bar[s], cb * g.BK, wo0 * g.stride - g.pad + s_,
                                        ho0 * g.stride - g.pad + r, b0);
                            tma_load_3d(sa + a_bytes, &tmap_w, &full_bar[s], cb * g.BK, tap, n0);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = umma_idesc(/*S32*/ 2, /*A u8*/ 0, /*B s8*/ 1, 128, BN);
        const int mmas = g.BK / 32;
        const uint64_t ring_desc0 = umma_smem_desc(smem_u32(smem), g.BK);      // stage s: + s * stage_bytes (one add per stage)
        const uint64_t halo_w_desc = umma_smem_desc(smem_u32(wres), g.BK);     // halo mode: resident weight slabs
        uint32_t halo_a[9];                                                     // halo mode: operand offset of tap (r, s) in a stage
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) halo_a[tap] = (uint32_t)(tap % 3) * hbox_bytes + (uint32_t)((tap / 3) * g.TW * g.BK);
        RingPos rp;
        int j = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            const int buf = j & 1;
            mbar_wait(&tempty_bar[buf], ((j >> 1) & 1) ^ 1);       // epilogue has drained this buffer (first use passes)
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
            if (g.halo) {
                if (j == 0) mbar_wait(wfull_bar, 0);
                const int s = rp.s;
                const uint32_t ph = rp.ph;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    // halo mode is 3x3 by construction: the nine taps are unrolled and their operand offsets are the
                    // loop-invariant halo_a[] / tap * b_bytes - the rolled form spent ~75 instructions of uniform-datapath
                    // arithmetic and constant-bank reloads per tap (ncu source view), as long as the MMAs themselves
                    const uint64_t st_desc = desc_add(ring_desc0, (uint32_t)s * stage_bytes);
#pragma unroll
                    for (int tap = 0; tap < 9; ++tap) {
                        const uint64_t a_desc = desc_add(st_desc, halo_a[tap]);
                        const uint64_t b_desc = desc_add(halo_w_desc, (uint32_t)tap * b_bytes);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (k < mmas) umma_i8(d_tmem, desc_add(a_desc, k * 32), desc_add(b_desc, k * 32), idesc, (tap | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);
                    umma_commit(&tfull_bar[buf]);
                }
                __syncwarp();
                rp.next(g.stages);
                continue;
            }
            for (int kb = 0; kb < num_kb; ++kb, rp.next(g.stages)) {
                const int s = rp.s;
                const uint32_t ph = rp.ph;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint64_t a_desc = desc_add(ring_desc0, (uint32_t)s * stage_bytes);
                    const uint64_t b_desc = desc_add(a_desc, a_bytes);
                    if (mmas == 4) {
                        umma_i8(d_tmem, a_desc, b_desc, idesc, kb != 0);
#pragma unroll
                        for (int k = 1; k < 4; ++k) umma_i8(d_tmem, desc_add(a_desc, k * 32), desc_add(b_desc, k * 32), idesc, 1u);
                    } else {
                        for (int k = 0; k < mmas; ++k)
                            umma_i8(d_tmem, desc_add(a_desc, k * 32), desc_add(b_desc, k * 32), idesc, (kb | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);
                    if (kb == num_kb - 1) umma_commit(&tfull_bar[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue: 2 groups x 4 warps; group `grp` owns the slabs cs with cs % 2 == grp =====
        const int e_all = threadIdx.x - 64;                         // 0..255
        const int grp = e_all >> 7, et = e_all & 127;
        const int sub = warp & 3;                                   // TMEM lane quarter this warp may access
        const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
        const double sa = (double)pa.scale, la = (double)pa.lo;
        const bool need_la = pa.lo != 0.0f;
        const int per_img = g.TH * g.TW;
        const int i = sub * 32 + lane;                              // row of the tile = TMEM lane = staging row
        const int nl = i / per_img, rem = i - nl * per_img;
        const int th = rem / g.TW, tw = rem - th * g.TW;
        const bool in_box = nl < g.TN;
        uint8_t* my_staging = reinterpret_cast<uint8_t*>(staging) + grp * 2 * CV_SLAB_BYTES;
        const uint32_t stage_u32 = smem_u32(my_staging);
        const uint32_t pair_u32 = smem_u32(s_pair), c2_u32 = smem_u32(s_c2);
        const uint32_t row_off = (uint32_t)i * 128u, sw = (uint32_t)(i & 7);
        const float kfull = (float)(taps * g.Cin);
        const int bar_id = 1 + grp;
        int cur_tn = -1, j = 0;
        uint32_t slab = 0;                                          // slabs staged by this group so far (buffer = slab & 1)
        // boxsum of this thread's pixel, fetched one tile ahead so its latency hides behind the current tile
        auto fetch_box = [&](int t, bool& valid_out) -> int {
            if (t >= total_tiles) { valid_out = false; return 0; }
            int tn, b0, ho0, wo0;
            conv_decode_tile(t, g, tiles_m, tn, b0, ho0, wo0);
            const int b = b0 + nl, ho = ho0 + th, wo = wo0 + tw;
            valid_out = in_box && (b < g.B) && (ho < g.Ho) && (wo < g.Wo);
            return valid_out ? __ldg(ep.boxsum + ((int64_t)b * g.Ho + ho) * g.Wo + wo) : 0;
        };
        const bool want_stats = ep.stats != nullptr;
        double st_sum = 0.0, st_sq = 0.0;                            // this thread's share of SUM x, SUM x^2 (all its tiles)
        bool valid_next;
        int box_next = fetch_box(blockIdx.x, valid_next);
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            int tn, b0, ho0, wo0;
            conv_decode_tile(t, g, tiles_m, tn, b0, ho0, wo0);
            const int n0 = tn * BN;
            if (tn != cur_tn) {                                      // per-channel coefficients of this channel tile
                epi_bar_all();                                       // everyone is done with the previous ones
                for (int jn = e_all; jn < BN; jn += 256) {
                    const int n = n0 + jn;
                    float4 cf = make_float4(0.f, 0.f, 0.f, 0.f);
                    float c3 = 0.f, c4 = 0.f;
                    if (n < g.Cout) {
                        const QParams pw = make_qparams_f(__ldg(ep.w_row_min + n), __ldg(ep.w_row_max + n), ep.w_bits);
                        const double swd = (double)pw.scale,
                                     lw = (double)pw.lo + (double)(1 << (ep.w_bits - 1)) * (double)pw.scale;
                        int tall = 0;
                        for (int tp = 0; tp < taps; ++tp) {
                            const int v = __ldg(ep.tapsum + (int64_t)n * taps + tp);
                            s_tap[jn * taps + tp] = v;
                            tall += v;
                        }
                        cf.x = (float)(sa * swd);
                        cf.y = (float)(sa * lw);
                        cf.z = ep.bias ? __ldg(ep.bias + n) : 0.f;
                        c3 = (float)(la * swd);
                        c4 = (float)(la * lw);
                        cf.w = fmaf(c3, (float)tall, c4 * kfull);
                    }
                    s_pair[(jn >> 1) * 4 + (jn & 1)] = cf.x;
                    s_pair[(jn >> 1) * 4 + 2 + (jn & 1)] = cf.y;
                    s_c2[jn] = cf.z; s_w[jn] = cf.w; s_c3[jn] = c3; s_c4[jn] = c4;
                }
                epi_bar_all();
                cur_tn = tn;
            }
            const bool valid = valid_next;
            const float rs_a = (float)box_next;
            box_next = fetch_box(t + (int)gridDim.x, valid_next);
            float kvalid = kfull;
            uint32_t mask = 0;
            bool interior = true;
            if (need_la && valid) {
                const int ho = ho0 + th, wo = wo0 + tw;
                int cnt = 0;
                for (int r = 0; r < g.kh; ++r)
                    for (int s_ = 0; s_ < g.kw; ++s_) {
                        const int hi = ho * g.stride - g.pad + r, wi = wo * g.stride - g.pad + s_;
                        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) { mask |= 1u << (r * g.kw + s_); ++cnt; }
                    }
                interior = cnt == taps;
                kvalid = (float)(cnt * g.Cin);
            }
            float t_mn = QB_INF, t_mx = -QB_INF, t_sum = 0.f, t_sq = 0.f;   // statistics of this thread's row of the tile
            const int ncols = (g.Cout - n0 < BN) ? g.Cout - n0 : BN;
            const int nslabs = (ncols + 31) >> 5;
            const int last_cs = ((nslabs - 1 - grp) & ~1) + grp;    // last slab of this group (< grp: the group has none)
            const int buf = j & 1;
            mbar_wait(&tfull_bar[buf], (j >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(sub * 32) << 16);
            if (nslabs <= grp) {                                     // nothing to drain for this group: release right away
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tempty_bar[buf]);
            }
#pragma unroll 1
            for (int cs = grp; cs < nslabs; cs += 2, ++slab) {
                const int c = cs * 32;
                uint32_t r[32];
                tmem_ld_32x32(t_addr + (uint32_t)c, r);
                tmem_ld_wait();
                if (cs == last_cs) {                                 // last read of this accumulator by this warp
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                const uint32_t dst = stage_u32 + (slab & 1u) * (uint32_t)CV_SLAB_BYTES + row_off;
                if (valid) {
                    const uint32_t pa = pair_u32 + (uint32_t)c * 8u, ca = c2_u32 + (uint32_t)c * 4u;
                    const uint64_t rs2 = pack2(rs_a, rs_a);
                    const bool st_fast = want_stats && !need_la && c + 32 <= ncols;
                    uint64_t sum2 = 0ull, sq2 = 0ull;                    // (0.f, 0.f)
                    // the statistics mode is loop invariant: three straight-line copies of the 16-pair loop instead of two
                    // branches per pair (ncu on the 64-channel layers: 13 % of the stall samples were branch resolution)
                    auto dequant_pair = [&](int q, float& va, float& vb) -> uint64_t {
                        uint64_t c0p, c1p;
                        lds_2x64(pa + q * 8, c0p, c1p);
                        const uint64_t c2p = lds_64(ca + q * 4);
                        const uint64_t v2 = fma2(c0p, pack2((float)(int32_t)r[q], (float)(int32_t)r[q + 1]), fma2(c1p, rs2, c2p));
                        unpack2(v2, va, vb);
                        r[q] = __float_as_uint(va);
                        r[q + 1] = __float_as_uint(vb);
                        return v2;
                    };
                    if (st_fast) {
#pragma unroll
                        for (int q = 0; q < 32; q += 2) {
                            float va, vb;
                            const uint64_t v2 = dequant_pair(q, va, vb);
                            sum2 = add2(sum2, v2);
                            sq2 = fma2(v2, v2, sq2);
                            t_mn = min3(t_mn, va, vb);
                            t_mx = max3(t_mx, va, vb);
                        }
                    } else if (want_stats && !need_la) {
#pragma unroll
                        for (int q = 0; q < 32; q += 2) {
                            float va, vb;
                            dequant_pair(q, va, vb);
                            if (c + q < ncols) { t_mn = fminf(t_mn, va); t_mx = fmaxf(t_mx, va); t_sum += va; t_sq = fmaf(va, va, t_sq); }
                            if (c + q + 1 < ncols) { t_mn = fminf(t_mn, vb); t_mx = fmaxf(t_mx, vb); t_sum += vb; t_sq = fmaf(vb, vb, t_sq); }
                        }
                    } else {
#pragma unroll
                        for (int q = 0; q < 32; q += 2) {
                            float va, vb;
                            dequant_pair(q, va, vb);
                        }
                    }
                    if (st_fast) {
                        float a0, a1, b0_, b1_;
                        unpack2(sum2, a0, a1);
                        unpack2(sq2, b0_, b1_);
                        t_sum += a0 + a1;
                        t_sq += b0_ + b1_;
                    }
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)q ^ sw) << 4)),
                                     "r"(r[4 * q]), "r"(r[4 * q + 1]), "r"(r[4 * q + 2]), "r"(r[4 * q + 3]) : "memory");
                    if (need_la) {                                   // inputs whose range does not start at 0 (rare): fix up
#pragma unroll 1                                                     // this thread's own row in place, column by column
                        for (int q = 0; q < 32; ++q) {
                            float add = s_w[c + q];
                            if (!interior) {
                                int tsum = 0;
                                for (int tp = 0; tp < taps; ++tp)
                                    if ((mask >> tp) & 1u) tsum += s_tap[(c + q) * taps + tp];
                                add = fmaf(s_c3[c + q], (float)tsum, s_c4[c + q] * kvalid);
                            }
                            const uint32_t a = dst + ((((uint32_t)q >> 2) ^ sw) << 4) + ((uint32_t)q & 3u) * 4u;
                            float v;
                            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
                            v += add;
                            asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
                            if (want_stats && c + q < ncols) {
                                t_mn = fminf(t_mn, v); t_mx = fmaxf(t_mx, v); t_sum += v; t_sq = fmaf(v, v, t_sq);
                            }
                        }
                    }
                }
                fence_proxy_async();
                if (et == 0) tma_store_wait_read<0>();               // every earlier store of this group has left its slab
                epi_bar(bar_id);
                if (et == 0) {
                    tma_store_4d(&tmap_o, my_staging + (slab & 1u) * CV_SLAB_BYTES, n0 + c, wo0, ho0, b0);
                    tma_store_commit();
                }
            }
            if (want_stats) {
                st_sum += (double)t_sum;
                st_sq += (double)t_sq;
                // per-sample extrema: order-independent integer atomics on monotone codes (exact, like measure.cu);
                // one atomic per warp when all its valid rows belong to one sample
                const int bb = valid ? b0 + nl : -1;
                const unsigned int act = __ballot_sync(0xffffffffu, valid && t_mn <= t_mx);
                if (act) {
                    const int lead = __ffs(act) - 1;
                    const int b_lead = __shfl_sync(0xffffffffu, bb, lead);
                    const bool uniform = __all_sync(0xffffffffu, !(valid && t_mn <= t_mx) || bb == b_lead);
                    if (uniform) {
                        const float wmn = warp_min(t_mn), wmx = warp_max(t_mx);
                        if (lane == lead) {
                            atomicMax(ep.st_row_max + b_lead, enc_ord(wmx));
                            atomicMax(ep.st_row_min + b_lead, ~enc_ord(wmn));
                        }
                    } else if (valid && t_mn <= t_mx) {
                        atomicMax(ep.st_row_max + bb, enc_ord(t_mx));
                        atomicMax(ep.st_row_min + bb, ~enc_ord(t_mn));
                    }
                }
            }
        }
        if (et == 0) tma_store_wait<0>();
        if (want_stats) {
            // CTA partial in a fixed order (warp shuffle tree, then warps 0..7), slot blockIdx.x; the last CTA to arrive
            // combines every slot and the per-sample extrema in a fixed order and leaves the workspace clean
            epi_bar_all();                                                   // both groups' stores have drained: staging is scratch now
            st_sum = warp_sum(st_sum);
            st_sq = warp_sum(st_sq);
            double* s_all = reinterpret_cast<double*>(staging);
            if (lane == 0) { s_all[(warp - 2) * 2] = st_sum; s_all[(warp - 2) * 2 + 1] = st_sq; }
            epi_bar_all();
            __shared__ bool s_last;
            if (e_all == 0) {
                double a = 0.0, c2 = 0.0;
                for (int w = 0; w < 8; ++w) { a += s_all[2 * w]; c2 += s_all[2 * w + 1]; }
                ep.st_psum[blockIdx.x] = a;
                ep.st_psq[blockIdx.x] = c2;
                __threadfence();
                s_last = (atomicAdd(ep.st_counter, 1u) == gridDim.x - 1);
            }
            epi_bar_all();
            if (s_last) {
                __threadfence();
                double smin = 0.0, smax = 0.0, tsum = 0.0, tsq = 0.0;
                float gmn = QB_INF, gmx = -QB_INF;
                for (int b = e_all; b < g.B; b += 256) {
                    const float rmx = dec_ord(__ldcg(ep.st_row_max + b)), rmn = dec_ord(~__ldcg(ep.st_row_min + b));
                    ep.st_row_max[b] = 0u;
                    ep.st_row_min[b] = 0u;
                    smax += (double)rmx; smin += (double)rmn;
                    gmn = fminf(gmn, rmn); gmx = fmaxf(gmx, rmx);
                }
                for (int k = e_all; k < (int)gridDim.x; k += 256) { tsum += __ldcg(ep.st_psum + k); tsq += __ldcg(ep.st_psq + k); }
                smin = warp_sum(smin); smax = warp_sum(smax); tsum = warp_sum(tsum); tsq = warp_sum(tsq);
                gmn = warp_min(gmn); gmx = warp_max(gmx);
                epi_bar_all();
                if (lane == 0) {
                    double* o = s_all + (warp - 2) * 6;
                    o[0] = smin; o[1] = smax; o[2] = tsum; o[3] = tsq; o[4] = (double)gmn; o[5] = (double)gmx;
                }
                epi_bar_all();
                if (e_all == 0) {
                    double rr[4] = {0, 0, 0, 0};
                    for (int w = 0; w < 8; ++w) {
                        for (int k = 0; k < 4; ++k) rr[k] += s_all[w * 6 + k];
                        gmn = fminf(gmn, (float)s_all[w * 6 + 4]);
                        gmx = fmaxf(gmx, (float)s_all[w * 6 + 5]);
                    }
                    ep.stats[0] = rr[0]; ep.stats[1] = rr[1]; ep.stats[2] = rr[2]; ep.stats[3] = rr[3];
                    ep.stats[4] = (double)gmn; ep.stats[5] = (double)gmx;
                    ep.stats[6] = (double)g.B * (double)g.Ho * (double)g.Wo * (double)g.Cout;
                    ep.stats[7] = (double)g.B;
                    *ep.st_counter = 0u;
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

// S[b,ho,wo] = SUM over the in-bounds filter taps of P[b,hi,wi] (P = per-pixel sums of the input codes): the row sums
// of the implicit im2col matrix, zero padding excluded.  One thread per output pixel; P is L2 resident.
__global__ void k_conv_boxsum(const int32_t* __restrict__ P, int32_t* __restrict__ S, int B, int H, int W, int Ho, int Wo,
                              int kh, int kw, int stride, int pad) {
    const int64_t m = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= (int64_t)B * Ho * Wo) return;
    const int wo = (int)(m % Wo), ho = (int)((m / Wo) % Ho), b = (int)(m / ((int64_t)Wo * Ho));
    int acc = 0;
    for (int r = 0; r < kh; ++r) {
        const int hi = ho * stride - pad + r;
        if (hi < 0 || hi >= H) continue;
        for (int s = 0; s < kw; ++s) {
            const int wi = wo * stride - pad + s;
            if (wi >= 0 && wi < W) acc += __ldg(P + ((int64_t)b * H + hi) * W + wi);
        }
    }
    S[m] = acc;
}

// ------------------------------------------------------------------------------------------------
// weight preparation: per-output-channel range + fake-quant (quantize.py:318-334) and, for the tensor cores, the codes
// re-ordered to [Cout][kh*kw][Cin] (K-major, channel innermost like the NHWC activations) plus T[n][tap].
// The codes are CENTRED: qc = q - 2^(bits-1) as signed int8, w_hat = qc*sw + (lw + 2^(bits-1)*sw).  Weights are nearly
// symmetric, so the centred offset lw' is ~20x smaller than lw and SUM qa*qc is itself small: the fp32 epilogue then
// has no large cancelling terms (with raw codes the K=4608 layers lost two digits).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_fq_rows_conv(const float* __restrict__ w, float* wq, uint8_t* codes_khwc, int32_t* tapsum, int Cin, int taps, int bits,
               float* row_min, float* row_max, int compute_range) {
    __shared__ float sm[64];
    __shared__ int s_tap[64];
    const int64_t n = blockIdx.x;
    const int cols = Cin * taps;
    const float* src = w + n * cols;                       // [Cin][taps]
    float lo, hi;
    if (compute_range) {
        float mn = QB_INF, mx = -QB_INF;
        for (int j = threadIdx.x; j < cols; j += 256) {
            const float v = src[j];
            mn = fminf(mn, v);
            mx = fmaxf(mx, v);
        }
        mn = warp_min(mn);
        mx = warp_max(mx);
        if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = mn; sm[32 + (threadIdx.x >> 5)] = mx; }
        __syncthreads();
        mn = warp_min((threadIdx.x & 31) < 8 ? sm[threadIdx.x & 31] : QB_INF);
        mx = warp_max((threadIdx.x & 31) < 8 ? sm[32 + (threadIdx.x & 31)] : -QB_INF);
        lo = mn;
        hi = mx;
        if (threadIdx.x == 0) { row_min[n] = lo; row_max[n] = hi; }
    } else {
        lo = row_min[n];
        hi = row_max[n];
    }
    if (threadIdx.x < 64) s_tap[threadIdx.x] = 0;
    __syncthreads();
    const QParams p = make_qparams_f(lo, hi, bits);
    for (int j = threadIdx.x; j < cols; j += 256) {
        const float q = q_affine_code(src[j], p);
        if (wq) wq[n * cols + j] = q_affine_deq(q, p);
        const int c = j / taps, tp = j - c * taps;
        const int qc = (int)q - (1 << (bits - 1));           // centred code in [-2^(b-1), 2^(b-1)): signed int8 operand
        if (codes_khwc) codes_khwc[n * cols + (int64_t)tp * Cin + c] = (uint8_t)(int8_t)qc;
        if (tapsum) atomicAdd(&s_tap[tp], qc);               // integer: order independent
    }
    __syncthreads();
    if (tapsum && threadIdx.x < taps) tapsum[n * taps + threadIdx.x] = s_tap[threadIdx.x];
}

extern "C" int qb_fakequant_rows_conv(const float* w, float* wq, uint8_t* codes_khwc, int32_t* tapsum, int64_t Cout,
                                      int64_t Cin, int taps, int bits, float* row_min, float* row_max, int compute_range,
                                      void* stream) {
    QB_CHECK_ARG(w && row_min && row_max && (wq || codes_khwc), "qb_fakequant_rows_conv: null pointer");
    QB_CHECK_ARG(Cout > 0 && Cin > 0 && taps > 0 && taps <= 64, "qb_fakequant_rows_conv: bad shape (taps<=64)");
    QB_CHECK_ARG(bits >= 1 && bits <= 8, "qb_fakequant_rows_conv: uint8 codes need bits<=8");
    k_fq_rows_conv<<<(unsigned)Cout, 256, 0, qb_stream(stream)>>>(w, wq, codes_khwc, tapsum, (int)Cin, taps, bits, row_min,
                                                                 row_max, compute_range);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// host launcher
// ------------------------------------------------------------------------------------------------
static size_t conv_tail_bytes(int BN, int taps) { return 320 + (size_t)BN * 24 + (size_t)BN * taps * 4; }

template <int BN>
static int launch_conv(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& to, const ConvGeom& g,
                       const ConvEpilogue& ep, cudaStream_t st) {
    const size_t stage = g.halo ? (size_t)g.kw * g.BK * g.TW * (g.TH + g.kh - 1) : (size_t)(128 + BN) * g.BK;
    const size_t smem = (size_t)g.stages * stage + CV_STAGING_BYTES + 1024 + conv_tail_bytes(BN, g.kh * g.kw) +
                        (g.halo ? (size_t)g.kh * g.kw * BN * g.BK : 0);
    QB_CUDA(cudaFuncSetAttribute(k_conv_i8<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   // per device: not cached
    const int64_t total = (int64_t)g.tiles_w * g.tiles_h * g.tiles_b * ((g.Cout + BN - 1) / BN);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const unsigned grid = (unsigned)(total < sms ? total : sms);       // persistent: one CTA per SM
    k_conv_i8<BN><<<grid, CV_THREADS, smem, st>>>(tx, tw, to, g, ep);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_i8_fwd(const uint8_t* x_codes, const uint8_t* w_codes_khwc, float* out, int64_t B, int64_t H,
                              int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad,
                              const float* a_range, int a_bits, const float* w_row_min, const float* w_row_max, int w_bits,
                              const int32_t* pixsum, int32_t* boxsum_ws, const int32_t* tapsum, const float* bias,
                              double* out_stats, void* stats_ws, size_t stats_ws_bytes, void* stream) {
    QB_CHECK_ARG(x_codes && w_codes_khwc && out && a_range && w_row_min && w_row_max && pixsum && boxsum_ws && tapsum,
                 "qb_conv_i8_fwd: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0,
                 "qb_conv_i8_fwd: bad shape");
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (Cin % 16 != 0 || kh * kw > CV_MAX_TAPS || stride > 2 || Ho < 1 || Wo < 1 || Cout % 16 != 0 ||
        ((reinterpret_cast<uintptr_t>(x_codes) | reinterpret_cast<uintptr_t>(w_codes_khwc) | reinterpret_cast<uintptr_t>(out)) & 15) != 0) {
        qb_set_error("qb_conv_i8_fwd: unsupported geometry (need Cin%%16==0, Cout%%16==0, kh*kw<=9, stride<=2, 16-byte "
                     "aligned tensors): Cin=%lld Cout=%lld k=%dx%d stride=%d", (long long)Cin, (long long)Cout, kh, kw, stride);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(B * Ho * Wo < 2147483647LL && B * H * W < 2147483647LL, "qb_conv_i8_fwd: pixel count exceeds int32");
    ConvGeom g;
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.Cin = (int)Cin; g.Ho = (int)Ho; g.Wo = (int)Wo; g.Cout = (int)Cout;
    g.kh = kh; g.kw = kw; g.stride = stride; g.pad = pad;
    g.BK = (Cin % 128 == 0) ? 128 : (Cin % 64 == 0 ? 64 : 32);
    // output-pixel tile: whole rows when they fit, several rows / images per tile when the image is small
    g.TW = (int)(Wo < 128 ? Wo : 128);
    g.TH = (int)(128 / g.TW < Ho ? 128 / g.TW : Ho);
    g.TN = (g.TH == Ho && g.TW == Wo) ? (int)(128 / (Ho * Wo) < B ? 128 / (Ho * Wo) : B) : 1;
    if (g.TN < 1) g.TN = 1;
    g.tiles_w = (int)((Wo + g.TW - 1) / g.TW);
    g.tiles_h = (int)((Ho + g.TH - 1) / g.TH);
    g.tiles_b = (int)((B + g.TN - 1) / g.TN);
    const int BN = Cout >= 256 ? 256 : (Cout >= 128 ? 128 : 64);
    // halo mode (see the kernel): 3x3, stride 1, pad 1, all input channels in one K block, all output channels in one
    // N tile, and an 8 x 16 pixel tile so that a filter-row shift is a whole number of swizzle atoms
    static const bool halo_off = getenv("QB_CONV_HALO") && getenv("QB_CONV_HALO")[0] == '0';
    g.halo = (!halo_off && kh == 3 && kw == 3 && stride == 1 && pad == 1 && Cin == g.BK && Cout == BN && BN == 64 && Wo >= 8 &&
              Ho >= 16) ? 1 : 0;
    if (g.halo) {
        g.TW = 8; g.TH = 16; g.TN = 1;
        g.tiles_w = (int)((Wo + g.TW - 1) / g.TW);
        g.tiles_h = (int)((Ho + g.TH - 1) / g.TH);
        g.tiles_b = (int)B;
    }
    g.d_tiles_m = make_fastdiv((uint32_t)(g.tiles_w * g.tiles_h * g.tiles_b));
    g.d_tiles_w = make_fastdiv((uint32_t)g.tiles_w);
    g.d_tiles_h = make_fastdiv((uint32_t)g.tiles_h);
    {
        const size_t stage = g.halo ? (size_t)kw * g.BK * g.TW * (g.TH + kh - 1) : (size_t)(128 + BN) * g.BK;
        const size_t fixed = CV_STAGING_BYTES + 1024 + conv_tail_bytes(BN, kh * kw) + (g.halo ? (size_t)kh * kw * BN * g.BK : 0);
        int s = (int)((225 * 1024 - fixed) / stage);
        g.stages = s > CV_MAX_STAGES ? CV_MAX_STAGES : (s < 2 ? 2 : s);
    }
    cudaStream_t st = qb_stream(stream);
    {
        const int64_t M = B * Ho * Wo;
        k_conv_boxsum<<<(unsigned)((M + 255) / 256), 256, 0, st>>>(pixsum, boxsum_ws, (int)B, (int)H, (int)W, (int)Ho, (int)Wo,
                                                                  kh, kw, stride, pad);
        QB_LAUNCH_CHECK();
    }
    CUtensorMap tx, tw, to;
    {   // activations: dims innermost-first {Cin, W, H, B}
        const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {(uint64_t)Cin, (uint64_t)W * Cin, (uint64_t)H * W * Cin};
        const uint32_t box[4] = {(uint32_t)g.BK, (uint32_t)(g.TW * stride), (uint32_t)((g.TH + (g.halo ? kh - 1 : 0)) * stride),
                                 (uint32_t)g.TN};
        const uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
        int rc = qb_make_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, x_codes, dims, str, box, es, g.BK);
        if (rc) return rc;
    }
    {   // weights: [Cout][tap][Cin] (a box row = BK channels of ONE tap; zero fill beyond Cin)
        const uint64_t dims[3] = {(uint64_t)Cin, (uint64_t)(kh * kw), (uint64_t)Cout};
        const uint64_t str[2] = {(uint64_t)Cin, (uint64_t)kh * kw * Cin};
        const uint32_t box[3] = {(uint32_t)g.BK, 1u, (uint32_t)BN};
        int rc = qb_make_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w_codes_khwc, dims, str, box, nullptr, g.BK);
        if (rc) return rc;
    }
    {   // output fp32 NHWC: dims {Cout, Wo, Ho, B}; one staged slab = 32 channels x the tile's pixels
        const uint64_t dims[4] = {(uint64_t)Cout, (uint64_t)Wo, (uint64_t)Ho, (uint64_t)B};
        const uint64_t str[3] = {(uint64_t)Cout * 4, (uint64_t)Wo * Cout * 4, (uint64_t)Ho * Wo * Cout * 4};
        const uint32_t box[4] = {32u, (uint32_t)g.TW, (uint32_t)g.TH, (uint32_t)g.TN};
        int rc = qb_make_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, dims, str, box, nullptr, 128);
        if (rc) return rc;
    }
    ConvEpilogue ep{a_range, a_bits, w_row_min, w_row_max, w_bits, boxsum_ws, tapsum, bias,
                    nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (out_stats) {
        if (!stats_ws || stats_ws_bytes < qb_measure_ws_bytes(B)) {
            qb_set_error("qb_conv_i8_fwd: statistics workspace %zu < %zu bytes", stats_ws_bytes, qb_measure_ws_bytes(B));
            return QB_ERR_WORKSPACE;
        }
        char* base = static_cast<char*>(stats_ws);
        ep.stats = out_stats;
        ep.st_counter = reinterpret_cast<unsigned int*>(base);
        ep.st_psum = reinterpret_cast<double*>(base + 64);
        ep.st_psq = ep.st_psum + QB_MS_MAX_BLOCKS;
        ep.st_row_max = reinterpret_cast<unsigned int*>(base + QB_MS_FIXED_BYTES);
        ep.st_row_min = ep.st_row_max + B;
    }
    if (BN == 256) return launch_conv<256>(tx, tw, to, g, ep, st);
    if (BN == 128) return launch_conv<128>(tx, tw, to, g, ep, st);
    return launch_conv<64>(tx, tw, to, g, ep, st);
}
