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
#include "tc_common.cuh"

constexpr int CV_THREADS = 192;
constexpr int CV_MAX_TAPS = 9;

struct ConvGeom {
    int B, H, W, Cin, Ho, Wo, Cout;
    int kh, kw, stride, pad;
    int TW, TH, TN;             // tile of output pixels (TN*TH*TW <= 128)
    int tiles_w, tiles_h, tiles_b;
    int BK;                     // bytes (= channels) of K per pipeline stage: 32, 64 or 128
    int stages;
};

struct ConvEpilogue {
    const float* a_range;       // device {min,max} of the input quantizer (QB_SCALE_DOUBLE)
    int a_bits;
    const float* w_row_min;     // per output channel (QB_SCALE_FLOAT)
    const float* w_row_max;
    int w_bits;
    const int32_t* pixsum;      // P[B*H*W]
    const int32_t* tapsum;      // T[Cout][kh*kw]
    const float* bias;          // fake-quantized bias or nullptr
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

// Persistent kernel: gridDim.x CTAs (one per SM) walk the tile list t = blockIdx.x, += gridDim.x; tiles are ordered with
// the output-channel tile slowest so a CTA's per-channel coefficients rarely change.  Three asynchronous roles:
//   warp 0   TMA producer: keeps the shared-memory ring full ACROSS tile boundaries
//   warp 1   MMA issuer  : accumulates tile j into TMEM buffer j&1 (2 x BN columns allocated)
//   warps 2-5 epilogue   : drain buffer j&1 (tcgen05.ld), dequantise, store - overlapped with the MMAs of tile j+1
template <int BN>
__global__ void __launch_bounds__(CV_THREADS, 1)
k_conv_i8(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w, float* __restrict__ out,
          ConvGeom g, ConvEpilogue ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t a_bytes = 128u * g.BK, b_bytes = (uint32_t)BN * g.BK, stage_bytes = a_bytes + b_bytes;
    uint8_t* tail = smem + (size_t)g.stages * stage_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
    uint64_t* empty_bar = full_bar + 8;
    uint64_t* tfull_bar = empty_bar + 8;                       // [2] accumulator buffer ready
    uint64_t* tempty_bar = tfull_bar + 2;                      // [2] accumulator buffer drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float* s_c1 = reinterpret_cast<float*>(tail + 256);       // sa*sw[n]
    float* s_c2 = s_c1 + BN;                                   // sa*lw'[n]
    float* s_c3 = s_c2 + BN;                                   // la*sw[n]
    float* s_c4 = s_c3 + BN;                                   // la*lw'[n]
    float* s_bias = s_c4 + BN;
    int32_t* s_tap = reinterpret_cast<int32_t*>(s_bias + BN);  // [BN][taps]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int taps = g.kh * g.kw;
    const int cblocks = g.Cin / g.BK;
    const int num_kb = taps * cblocks;
    const int tiles_m = g.tiles_w * g.tiles_h * g.tiles_b;
    const int tiles_n = (g.Cout + BN - 1) / BN;
    const int total_tiles = tiles_m * tiles_n;
    constexpr uint32_t TMEM_COLS = 2 * BN;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_x);
        tma_prefetch_desc(&tmap_w);
        for (int s = 0; s < g.stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);
            mbar_init(&tempty_bar[i], 4);                      // one arrival per epilogue warp
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
            uint32_t it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                const int tn = t / tiles_m;
                int tm = t - tn * tiles_m;
                const int tw_i = tm % g.tiles_w; tm /= g.tiles_w;
                const int th_i = tm % g.tiles_h; tm /= g.tiles_h;
                const int b0 = tm * g.TN, ho0 = th_i * g.TH, wo0 = tw_i * g.TW, n0 = tn * BN;
                for (int tap = 0; tap < taps; ++tap) {
                    const int r = tap / g.kw, s_ = tap - r * g.kw;
                    for (int cb = 0; cb < cblocks; ++cb, ++it) {
                        const int s = it % g.stages;
                        const uint32_t ph = (it / g.stages) & 1;
                        mbar_wait(&empty_bar[s], ph ^ 1);
                        mbar_arrive_expect_tx(&full_bar[s], (uint32_t)(g.TN * g.TH * g.TW * g.BK) + b_bytes);
                        uint8_t* sa = smem + (size_t)s * stage_bytes;
                        tma_load_4d(sa, &tmap_x, &full_bar[s], cb * g.BK, wo0 * g.stride - g.pad + s_,
                                    ho0 * g.stride - g.pad + r, b0);
                        tma_load_2d(sa + a_bytes, &tmap_w, &full_bar[s], tap * g.Cin + cb * g.BK, n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        constexpr uint32_t idesc = umma_idesc(/*S32*/ 2, /*A u8*/ 0, /*B s8*/ 1, 128, BN);
        const int mmas = g.BK / 32;
        uint32_t it = 0;
        int j = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            const int buf = j & 1;
            mbar_wait(&tempty_bar[buf], ((j >> 1) & 1) ^ 1);       // epilogue has drained this buffer (first use passes)
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem_base + (uint32_t)(buf * BN);
            for (int kb = 0; kb < num_kb; ++kb, ++it) {
                const int s = it % g.stages;
                const uint32_t ph = (it / g.stages) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
                    const uint32_t b_addr = a_addr + a_bytes;
                    for (int k = 0; k < mmas; ++k)
                        umma_i8(d_tmem, umma_smem_desc(a_addr + k * 32, g.BK), umma_smem_desc(b_addr + k * 32, g.BK), idesc,
                                (kb | k) != 0);
                    umma_commit(&empty_bar[s]);
                    if (kb == num_kb - 1) umma_commit(&tfull_bar[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ===== epilogue warps (128 threads) =====
        const int et = threadIdx.x - 64;
        const int sub = warp & 3;                                   // TMEM lane quarter this warp may access
        const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
        const double sa = (double)pa.scale, la = (double)pa.lo;
        const bool need_la = pa.lo != 0.0f;
        const int per_img = g.TH * g.TW;
        const int i = sub * 32 + lane;
        const int nl = i / per_img, rem = i - nl * per_img;
        const int th = rem / g.TW, tw = rem - th * g.TW;
        int cur_tn = -1, j = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++j) {
            const int tn = t / tiles_m;
            int tm = t - tn * tiles_m;
            const int tw_i = tm % g.tiles_w; tm /= g.tiles_w;
            const int th_i = tm % g.tiles_h; tm /= g.tiles_h;
            const int b0 = tm * g.TN, ho0 = th_i * g.TH, wo0 = tw_i * g.TW, n0 = tn * BN;
            if (tn != cur_tn) {                                      // per-channel coefficients of this channel tile
                asm volatile("bar.sync 1, 128;" ::: "memory");      // everyone is done with the previous ones
                for (int jn = et; jn < BN; jn += 128) {
                    const int n = n0 + jn;
                    float c1 = 0.f, c2 = 0.f, c3 = 0.f, c4 = 0.f, bs = 0.f;
                    if (n < g.Cout) {
                        const QParams pw = make_qparams_f(__ldg(ep.w_row_min + n), __ldg(ep.w_row_max + n), ep.w_bits);
                        const double sw = (double)pw.scale,
                                     lw = (double)pw.lo + (double)(1 << (ep.w_bits - 1)) * (double)pw.scale;
                        c1 = (float)(sa * sw); c2 = (float)(sa * lw); c3 = (float)(la * sw); c4 = (float)(la * lw);
                        bs = ep.bias ? __ldg(ep.bias + n) : 0.f;
                        for (int tp = 0; tp < taps; ++tp) s_tap[jn * taps + tp] = __ldg(ep.tapsum + (int64_t)n * taps + tp);
                    }
                    s_c1[jn] = c1; s_c2[jn] = c2; s_c3[jn] = c3; s_c4[jn] = c4; s_bias[jn] = bs;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                cur_tn = tn;
            }
            // this thread's output pixel of the tile
            const int b = b0 + nl, ho = ho0 + th, wo = wo0 + tw;
            const bool valid = (nl < g.TN) && (b < g.B) && (ho < g.Ho) && (wo < g.Wo);
            float rs_a = 0.f, kvalid = 0.f;
            uint32_t mask = 0;
            if (valid) {
                int cnt = 0, acc = 0;
                for (int tp = 0; tp < taps; ++tp) {
                    const int r = tp / g.kw, s_ = tp - r * g.kw;
                    const int hi = ho * g.stride - g.pad + r, wi = wo * g.stride - g.pad + s_;
                    if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) {
                        mask |= 1u << tp;
                        ++cnt;
                        acc += __ldg(ep.pixsum + ((int64_t)b * g.H + hi) * g.W + wi);
                    }
                }
                rs_a = (float)acc;
                kvalid = (float)(cnt * g.Cin);
            }
            const bool interior = mask == ((1u << taps) - 1u);
            float* dst = out + (((int64_t)b * g.Ho + ho) * g.Wo + wo) * g.Cout + n0;
            const int buf = j & 1;
            mbar_wait(&tfull_bar[buf], (j >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(sub * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32(t_addr + (uint32_t)c, r);
                tmem_ld_wait();
                if (c + 32 >= BN) {                                  // last read of this buffer: hand it back to the MMA warp
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tempty_bar[buf]);
                }
                if (valid) {
#pragma unroll
                    for (int q = 0; q < 32; ++q) {
                        const int jn = c + q;
                        float v = fmaf(s_c1[jn], (float)(int32_t)r[q], fmaf(s_c2[jn], rs_a, s_bias[jn]));
                        if (need_la) {
                            int tsum = 0;
                            for (int tp = 0; tp < taps; ++tp)
                                if (interior || ((mask >> tp) & 1u)) tsum += s_tap[jn * taps + tp];
                            v += fmaf(s_c3[jn], (float)tsum, s_c4[jn] * kvalid);
                        }
                        r[q] = __float_as_uint(v);
                    }
                    if (n0 + c + 32 <= g.Cout && (g.Cout & 3) == 0) {
#pragma unroll
                        for (int q = 0; q < 32; q += 4)
                            __stcs(reinterpret_cast<float4*>(dst + c + q),
                                   make_float4(__uint_as_float(r[q]), __uint_as_float(r[q + 1]), __uint_as_float(r[q + 2]),
                                               __uint_as_float(r[q + 3])));
                    } else {
                        for (int q = 0; q < 32; ++q)
                            if (n0 + c + q < g.Cout) dst[c + q] = __uint_as_float(r[q]);
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
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
template <int BN>
static int launch_conv(const CUtensorMap& tx, const CUtensorMap& tw, float* out, const ConvGeom& g, const ConvEpilogue& ep,
                       cudaStream_t st) {
    const size_t stage = (size_t)(128 + BN) * g.BK;
    const size_t tail = 256 + (size_t)BN * 20 + (size_t)BN * g.kh * g.kw * 4;
    const size_t smem = (size_t)g.stages * stage + 1024 + tail;
    static size_t configured = 0;
    if (smem > configured) {
        QB_CUDA(cudaFuncSetAttribute(k_conv_i8<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured = smem;
    }
    const int64_t total = (int64_t)g.tiles_w * g.tiles_h * g.tiles_b * ((g.Cout + BN - 1) / BN);
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const unsigned grid = (unsigned)(total < sms ? total : sms);       // persistent: one CTA per SM
    k_conv_i8<BN><<<grid, CV_THREADS, smem, st>>>(tx, tw, out, g, ep);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_i8_fwd(const uint8_t* x_codes, const uint8_t* w_codes_khwc, float* out, int64_t B, int64_t H,
                              int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad,
                              const float* a_range, int a_bits, const float* w_row_min, const float* w_row_max, int w_bits,
                              const int32_t* pixsum, const int32_t* tapsum, const float* bias, void* stream) {
    QB_CHECK_ARG(x_codes && w_codes_khwc && out && a_range && w_row_min && w_row_max && pixsum && tapsum,
                 "qb_conv_i8_fwd: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && Cin > 0 && Cout > 0 && kh > 0 && kw > 0 && stride > 0 && pad >= 0,
                 "qb_conv_i8_fwd: bad shape");
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (Cin % 32 != 0 || kh * kw > CV_MAX_TAPS || stride > 2 || Ho < 1 || Wo < 1 || Cout % 16 != 0 ||
        ((reinterpret_cast<uintptr_t>(x_codes) | reinterpret_cast<uintptr_t>(w_codes_khwc)) & 15) != 0) {
        qb_set_error("qb_conv_i8_fwd: unsupported geometry (need Cin%%32==0, Cout%%16==0, kh*kw<=9, stride<=2, 16-byte "
                     "aligned codes): Cin=%lld Cout=%lld k=%dx%d stride=%d", (long long)Cin, (long long)Cout, kh, kw, stride);
        return QB_ERR_UNSUPPORTED;
    }
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
    {
        const size_t stage = (size_t)(128 + BN) * g.BK;
        const size_t tail = 256 + (size_t)BN * 20 + (size_t)BN * kh * kw * 4 + 1024;
        int s = (int)((220 * 1024 - tail) / stage);
        g.stages = s > 8 ? 8 : (s < 2 ? 2 : s);
    }
    CUtensorMap tx, tw;
    {   // activations: dims innermost-first {Cin, W, H, B}
        const uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)B};
        const uint64_t str[3] = {(uint64_t)Cin, (uint64_t)W * Cin, (uint64_t)H * W * Cin};
        const uint32_t box[4] = {(uint32_t)g.BK, (uint32_t)(g.TW * stride), (uint32_t)(g.TH * stride), (uint32_t)g.TN};
        const uint32_t es[4] = {1, (uint32_t)stride, (uint32_t)stride, 1};
        int rc = qb_make_tmap(&tx, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, x_codes, dims, str, box, es, g.BK);
        if (rc) return rc;
    }
    {   // weights: [Cout][K = kh*kw*Cin]
        const uint64_t K = (uint64_t)kh * kw * Cin;
        const uint64_t dims[2] = {K, (uint64_t)Cout}, str[1] = {K};
        const uint32_t box[2] = {(uint32_t)g.BK, (uint32_t)BN};
        int rc = qb_make_tmap(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, w_codes_khwc, dims, str, box, nullptr, g.BK);
        if (rc) return rc;
    }
    ConvEpilogue ep{a_range, a_bits, w_row_min, w_row_max, w_bits, pixsum, tapsum, bias};
    cudaStream_t st = qb_stream(stream);
    if (BN == 256) return launch_conv<256>(tx, tw, out, g, ep, st);
    if (BN == 128) return launch_conv<128>(tx, tw, out, g, ep, st);
    return launch_conv<64>(tx, tw, out, g, ep, st);
}
