// conv_dw.cu - QConv2d forward for DEPTHWISE 3x3 convolutions (groups == C_in == C_out: the 13 depthwise layers of
// models/mobilenet_quantized.py:21-60) on the integer codes.  Reference call: F.conv2d(qinput, qweight, qbias, stride,
// padding, dilation, groups) on the fake-quantized tensors (models/modules/quantize.py:143-148, 343).
//
// A depthwise filter contracts 9 taps per output: 18 operations per 5 bytes of HBM traffic (1 B of input codes, 4 B of
// fp32 output), so this is a streaming kernel bounded by HBM, not GEMM-shaped work - no tensor cores.  With
//   x_hat = qa*sa + la  (0 in the padding),   w_hat[n][t] = qc[n][t]*sw[n] + lw'[n]   (centred codes, see conv_i8.cu)
//   out[p,n] = sa*sw[n]*SUM_V qa*qc + sa*lw'[n]*SUM_V qa + la*(sw[n]*SUM_V qc + lw'[n]*|V|) + bias[n]
// where V = the in-bounds taps of output pixel p: exact zero padding for any input range.
//
// Mapping: one thread = TWO horizontally adjacent output pixels x 4 consecutive channels (one 32-bit word of NHWC codes
// per tap, a 3 x (3 + stride) patch of words shared by the two pixels, one float4 of output per pixel).  SUM qa*qc: one IDP.4A per (tap, channel) against a weight word holding the channel's centred code in
// that channel's byte lane and zeros elsewhere (extract + multiply + accumulate in one instruction); SUM qa: packed
// 16-bit lanes (<= 9*255 per lane).  The 36 weight words and the interior coefficients live in registers (the border-only ones
// in shared memory) and are rebuilt only when a thread's channel group changes (never, when 256 % (C/4) == 0); the
// (sample, row, column) of a thread's item advances incrementally - no division in the loop.  A CTA walks a contiguous range
// of outputs, consecutive threads -> consecutive channel groups -> consecutive pixels: loads and stores of a warp are
// contiguous 128 B / 512 B runs, the 9x input re-use is served by L1/L2.  Optionally the QuantMeasure statistics block
// of the OUTPUT (the range reduction of the RangeBN that follows, quantize.py:223-235) is produced on the way, in
// qb_measure_stats' workspace protocol.
#include "common.cuh"

constexpr int DW_THREADS = 256;

__device__ __forceinline__ int dp4a_us(uint32_t a_u8x4, uint32_t b_s8x4, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
    return d;
}

struct DwGeom {
    int B, H, W, C, Ho, Wo, Wp, pad, C4;
    uint32_t items;         // B*Ho*Wp*C4 (Wp = pixel pairs per output row)
    uint32_t per_cta;       // multiple of DW_THREADS
    int dcg, dq;            // DW_THREADS % C4, DW_THREADS / C4
};

struct DwEpi {
    const float* a_range;
    int a_bits;
    const float* w_row_min;
    const float* w_row_max;
    int w_bits;
    const float* bias;
    double* stats;
    unsigned int* st_counter;
    double* st_psum;
    double* st_psq;
    unsigned int* st_row_max;
    unsigned int* st_row_min;
};

template <bool STATS, int STRIDE>
__global__ void __launch_bounds__(DW_THREADS, 2)
k_conv_dw3x3(const uint8_t* __restrict__ x, const int8_t* __restrict__ w, float* __restrict__ out, DwGeom g, DwEpi ep) {
    constexpr int NC = 3 + STRIDE;                                   // patch columns shared by the two pixels
    __shared__ float s_border[3][4][DW_THREADS];                     // c3 = la*sw, c4 = la*lw', bias: border pixels only
    const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
    const double sa = (double)pa.scale, la = (double)pa.lo;
    uint32_t wsel[9][4];
    float c0[4], c1[4], cw[4];
    int cur_cg = -1;
    float mn = QB_INF, mx = -QB_INF;                                 // statistics of the produced outputs
    double dsum = 0.0, dsq = 0.0;
    int cur_b = -1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tid = threadIdx.x;

    const uint32_t begin = blockIdx.x * g.per_cta;
    const uint32_t end = (g.items - begin > g.per_cta) ? begin + g.per_cta : g.items;
    // this thread's first item, decomposed once; afterwards advanced incrementally
    int cg, wp, ho, b;
    {
        const uint32_t item = begin + tid;
        const uint32_t q = item / (uint32_t)g.C4;
        cg = (int)(item - q * g.C4);
        const uint32_t row = q / (uint32_t)g.Wp;
        wp = (int)(q - row * g.Wp);
        b = (int)(row / (uint32_t)g.Ho);
        ho = (int)(row - (uint32_t)b * g.Ho);
    }
    for (uint32_t base = begin; base < end; base += DW_THREADS) {
        const uint32_t item = base + tid;
        const bool valid = item < end;
        float4 o0 = make_float4(0.f, 0.f, 0.f, 0.f), o1 = o0;
        bool two = false;
        if (valid) {
            if (cg != cur_cg) {                                       // (re)build this thread's weights and coefficients
                cur_cg = cg;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    const int n = cg * 4 + ch;
                    const QParams pw = make_qparams_f(__ldg(ep.w_row_min + n), __ldg(ep.w_row_max + n), ep.w_bits);
                    const double swd = (double)pw.scale,
                                 lw = (double)pw.lo + (double)(1 << (ep.w_bits - 1)) * (double)pw.scale;
                    int tall = 0;
#pragma unroll
                    for (int t = 0; t < 9; ++t) {
                        const int q = (int)__ldg(w + (int64_t)n * 9 + t);
                        tall += q;
                        wsel[t][ch] = ((uint32_t)q & 0xffu) << (8 * ch);
                    }
                    const float c3 = (float)(la * swd), c4 = (float)(la * lw), cb = ep.bias ? __ldg(ep.bias + n) : 0.f;
                    c0[ch] = (float)(sa * swd);
                    c1[ch] = (float)(sa * lw);
                    cw[ch] = fmaf(c3, (float)tall, c4 * 9.0f) + cb;
                    s_border[0][ch][tid] = c3;
                    s_border[1][ch][tid] = c4;
                    s_border[2][ch][tid] = cb;
                }
            }
            const int wo = wp * 2;
            two = wo + 1 < g.Wo;
            const int hi0 = ho * STRIDE - g.pad, wi0 = wo * STRIDE - g.pad;
            const uint8_t* xb = x + (((int64_t)b * g.H + hi0) * g.W + wi0) * g.C + cg * 4;
            float* ob = out + ((((int64_t)b * g.Ho + ho) * g.Wo + wo) * g.C + cg * 4);
            const bool interior = two && hi0 >= 0 && wi0 >= 0 && hi0 + 2 < g.H && wi0 + NC - 1 < g.W;
            if (interior) {
                uint32_t a[3][NC];
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int c = 0; c < NC; ++c)
                        a[r][c] = __ldg(reinterpret_cast<const uint32_t*>(xb + ((int64_t)r * g.W + c) * g.C));
                uint32_t cl[NC], chh[NC];                             // per-column code sums, packed 16-bit lanes
#pragma unroll
                for (int c = 0; c < NC; ++c) {
                    cl[c] = (a[0][c] & 0x00ff00ffu) + (a[1][c] & 0x00ff00ffu) + (a[2][c] & 0x00ff00ffu);
                    chh[c] = ((a[0][c] >> 8) & 0x00ff00ffu) + ((a[1][c] >> 8) & 0x00ff00ffu) + ((a[2][c] >> 8) & 0x00ff00ffu);
                }
                float res[2][4];
#pragma unroll
                for (int p = 0; p < 2; ++p) {
                    int acc[4] = {0, 0, 0, 0};
#pragma unroll
                    for (int t = 0; t < 9; ++t)
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch) acc[ch] = dp4a_us(a[t / 3][p * STRIDE + t % 3], wsel[t][ch], acc[ch]);
                    const uint32_t s_lo = cl[p * STRIDE] + cl[p * STRIDE + 1] + cl[p * STRIDE + 2];
                    const uint32_t s_hi = chh[p * STRIDE] + chh[p * STRIDE + 1] + chh[p * STRIDE + 2];
                    const float S[4] = {(float)(s_lo & 0xffffu), (float)(s_hi & 0xffffu), (float)(s_lo >> 16), (float)(s_hi >> 16)};
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) res[p][ch] = fmaf(c0[ch], (float)acc[ch], fmaf(c1[ch], S[ch], cw[ch]));
                }
                o0 = make_float4(res[0][0], res[0][1], res[0][2], res[0][3]);
                o1 = make_float4(res[1][0], res[1][1], res[1][2], res[1][3]);
            } else {
#pragma unroll 1
                for (int p = 0; p < (two ? 2 : 1); ++p) {
                    int acc[4] = {0, 0, 0, 0}, wn[4] = {0, 0, 0, 0}, nin = 0;
                    uint32_t s_lo = 0, s_hi = 0;
#pragma unroll
                    for (int t = 0; t < 9; ++t) {
                        const int hi = hi0 + t / 3, wi = wi0 + p * STRIDE + t % 3;
                        if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) {
                            const uint32_t av = __ldg(reinterpret_cast<const uint32_t*>(
                                xb + ((int64_t)(t / 3) * g.W + (p * STRIDE + t % 3)) * g.C));
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch) {
                                acc[ch] = dp4a_us(av, wsel[t][ch], acc[ch]);
                                wn[ch] = dp4a_us(0x01010101u, wsel[t][ch], wn[ch]);     // the centred code itself
                            }
                            s_lo += av & 0x00ff00ffu;
                            s_hi += (av >> 8) & 0x00ff00ffu;
                            ++nin;
                        }
                    }
                    const float S[4] = {(float)(s_lo & 0xffffu), (float)(s_hi & 0xffffu), (float)(s_lo >> 16), (float)(s_hi >> 16)};
                    const float fn = (float)nin;
                    float res[4];
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        const float c3 = s_border[0][ch][tid], c4 = s_border[1][ch][tid], cb = s_border[2][ch][tid];
                        res[ch] = fmaf(c0[ch], (float)acc[ch], fmaf(c1[ch], S[ch], fmaf(c3, (float)wn[ch], c4 * fn) + cb));
                    }
                    if (p == 0) o0 = make_float4(res[0], res[1], res[2], res[3]);
                    else o1 = make_float4(res[0], res[1], res[2], res[3]);
                }
            }
            __stcs(reinterpret_cast<float4*>(ob), o0);
            if (two) __stcs(reinterpret_cast<float4*>(ob + g.C), o1);
        }
        if (STATS) {
            // a warp's 32 items are consecutive, so the sample index changes rarely; when it does anywhere in the warp,
            // every lane hands the extrema of its previous sample to the per-sample slots
            if (__any_sync(0xffffffffu, valid && b != cur_b)) {
                if (cur_b >= 0 && mn <= mx) {
                    atomicMax(&ep.st_row_max[cur_b], enc_ord(mx));
                    atomicMax(&ep.st_row_min[cur_b], ~enc_ord(mn));
                }
                mn = QB_INF; mx = -QB_INF;
            }
            if (valid) {
                cur_b = b;
                mn = fminf(fminf(mn, fminf(o0.x, o0.y)), fminf(o0.z, o0.w));
                mx = fmaxf(fmaxf(mx, fmaxf(o0.x, o0.y)), fmaxf(o0.z, o0.w));
                float fs = (o0.x + o0.y) + (o0.z + o0.w), fq = (o0.x * o0.x + o0.y * o0.y) + (o0.z * o0.z + o0.w * o0.w);
                if (two) {
                    mn = fminf(fminf(mn, fminf(o1.x, o1.y)), fminf(o1.z, o1.w));
                    mx = fmaxf(fmaxf(mx, fmaxf(o1.x, o1.y)), fmaxf(o1.z, o1.w));
                    fs += (o1.x + o1.y) + (o1.z + o1.w);
                    fq += (o1.x * o1.x + o1.y * o1.y) + (o1.z * o1.z + o1.w * o1.w);
                }
                dsum += (double)fs;
                dsq += (double)fq;
            }
        }
        // advance (cg, wp, ho, b) by DW_THREADS items
        cg += g.dcg;
        int dq = g.dq;
        if (cg >= g.C4) { cg -= g.C4; ++dq; }
        wp += dq;
        if (wp >= g.Wp) {
            if (wp < 3 * g.Wp) {
                do { wp -= g.Wp; if (++ho == g.Ho) { ho = 0; ++b; } } while (wp >= g.Wp);
            } else {
                const uint32_t rows = (uint32_t)wp / (uint32_t)g.Wp;
                wp -= (int)(rows * g.Wp);
                const uint32_t hh = (uint32_t)ho + rows;
                const uint32_t db = hh / (uint32_t)g.Ho;
                ho = (int)(hh - db * g.Ho);
                b += (int)db;
            }
        }
    }
    if (!STATS) return;
    // ---- statistics: per-sample extrema, per-CTA partial sums, last CTA combines (qb_measure_stats' protocol) ----
    {
        const int b0 = __shfl_sync(0xffffffffu, cur_b, 0);
        if (__all_sync(0xffffffffu, cur_b == b0)) {
            const float a = warp_min(mn), c = warp_max(mx);
            if (lane == 0 && b0 >= 0 && a <= c) {
                atomicMax(&ep.st_row_max[b0], enc_ord(c));
                atomicMax(&ep.st_row_min[b0], ~enc_ord(a));
            }
        } else if (cur_b >= 0 && mn <= mx) {
            atomicMax(&ep.st_row_max[cur_b], enc_ord(mx));
            atomicMax(&ep.st_row_min[cur_b], ~enc_ord(mn));
        }
    }
    __shared__ double sd[2][DW_THREADS / 32];
    __shared__ bool last;
    dsum = warp_sum(dsum);
    dsq = warp_sum(dsq);
    if (lane == 0) { sd[0][warp] = dsum; sd[1][warp] = dsq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, c = 0.0;
        for (int k = 0; k < DW_THREADS / 32; ++k) { a += sd[0][k]; c += sd[1][k]; }
        ep.st_psum[blockIdx.x] = a;
        ep.st_psq[blockIdx.x] = c;
        __threadfence();
        last = (atomicAdd(ep.st_counter, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    double smin = 0.0, smax = 0.0, tsum = 0.0, tsq = 0.0;
    float gmn = QB_INF, gmx = -QB_INF;
    for (int b = threadIdx.x; b < g.B; b += DW_THREADS) {
        const float rmx = dec_ord(__ldcg(&ep.st_row_max[b]));
        const float rmn = dec_ord(~__ldcg(&ep.st_row_min[b]));
        ep.st_row_max[b] = 0u;
        ep.st_row_min[b] = 0u;
        smax += (double)rmx;
        smin += (double)rmn;
        gmn = fminf(gmn, rmn);
        gmx = fmaxf(gmx, rmx);
    }
    for (int i = threadIdx.x; i < (int)gridDim.x; i += DW_THREADS) {
        tsum += __ldcg(&ep.st_psum[i]);
        tsq += __ldcg(&ep.st_psq[i]);
    }
    smin = warp_sum(smin); smax = warp_sum(smax); tsum = warp_sum(tsum); tsq = warp_sum(tsq);
    gmn = warp_min(gmn); gmx = warp_max(gmx);
    __shared__ double fin[4][DW_THREADS / 32];
    __shared__ float fmm[2][DW_THREADS / 32];
    if (lane == 0) {
        fin[0][warp] = smin; fin[1][warp] = smax; fin[2][warp] = tsum; fin[3][warp] = tsq;
        fmm[0][warp] = gmn; fmm[1][warp] = gmx;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double r[4] = {0, 0, 0, 0};
        for (int k = 0; k < DW_THREADS / 32; ++k) {
            for (int j = 0; j < 4; ++j) r[j] += fin[j][k];
            gmn = fminf(gmn, fmm[0][k]);
            gmx = fmaxf(gmx, fmm[1][k]);
        }
        ep.stats[0] = r[0]; ep.stats[1] = r[1]; ep.stats[2] = r[2]; ep.stats[3] = r[3];
        ep.stats[4] = (double)gmn; ep.stats[5] = (double)gmx;
        ep.stats[6] = (double)g.B * (double)g.Ho * (double)g.Wo * (double)g.C; ep.stats[7] = (double)g.B;
        *ep.st_counter = 0u;
    }
}

extern "C" int qb_conv_dw_supported(int64_t C, int kh, int kw, int stride, int pad) {
    return (C > 0 && C % 4 == 0 && kh == 3 && kw == 3 && (stride == 1 || stride == 2) && pad >= 0 && pad <= 2) ? 1 : 0;
}

template <int STRIDE>
static void launch_dw(bool stats, unsigned grid, cudaStream_t st, const uint8_t* x, const int8_t* w, float* out, const DwGeom& g,
                      const DwEpi& ep) {
    if (stats) k_conv_dw3x3<true, STRIDE><<<grid, DW_THREADS, 0, st>>>(x, w, out, g, ep);
    else k_conv_dw3x3<false, STRIDE><<<grid, DW_THREADS, 0, st>>>(x, w, out, g, ep);
}

extern "C" int qb_conv_dw_fwd(const uint8_t* x_codes, const int8_t* w_codes, float* out, int64_t B, int64_t H, int64_t W,
                              int64_t C, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                              const float* w_row_min, const float* w_row_max, int w_bits, const float* bias,
                              double* out_stats, void* stats_ws, size_t stats_ws_bytes, void* stream) {
    QB_CHECK_ARG(x_codes && w_codes && out && a_range && w_row_min && w_row_max, "qb_conv_dw_fwd: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0, "qb_conv_dw_fwd: bad shape");
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8 && w_bits >= 1 && w_bits <= 8, "qb_conv_dw_fwd: uint8 codes need bits<=8");
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (!qb_conv_dw_supported(C, kh, kw, stride, pad) || Ho < 1 || Wo < 1 ||
        ((reinterpret_cast<uintptr_t>(x_codes) & 3) | (reinterpret_cast<uintptr_t>(out) & 15)) != 0) {
        qb_set_error("qb_conv_dw_fwd: unsupported geometry (need 3x3, C%%4==0, stride 1 or 2, pad<=2, aligned tensors): C=%lld "
                     "k=%dx%d stride=%d pad=%d", (long long)C, kh, kw, stride, pad);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(B * Ho * Wo * C < 2147483647LL * 4 && B * H * W * C < 2147483647LL * 4, "qb_conv_dw_fwd: tensor too large");
    DwGeom g;
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.C = (int)C; g.Ho = (int)Ho; g.Wo = (int)Wo; g.pad = pad;
    g.Wp = (int)((Wo + 1) / 2);
    g.C4 = (int)(C / 4);
    g.items = (uint32_t)(B * Ho * g.Wp * g.C4);
    g.dcg = DW_THREADS % g.C4;
    g.dq = DW_THREADS / g.C4;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    // 2 resident CTAs per SM x 6 waves, each CTA a contiguous range (a multiple of the CTA size, so a thread keeps its
    // channel group whenever 256 % (C/4) == 0)
    const int64_t ctas = (int64_t)sms * 12;
    int64_t per = ((int64_t)g.items + ctas - 1) / ctas;
    per = (per + DW_THREADS - 1) / DW_THREADS * DW_THREADS;
    g.per_cta = (uint32_t)per;
    const int64_t grid = ((int64_t)g.items + per - 1) / per;
    DwEpi ep{a_range, a_bits, w_row_min, w_row_max, w_bits, bias, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t st = qb_stream(stream);
    if (out_stats) {
        if (!stats_ws || stats_ws_bytes < qb_measure_ws_bytes(B) || grid > QB_MS_MAX_BLOCKS) {
            qb_set_error("qb_conv_dw_fwd: statistics workspace %zu < %zu bytes", stats_ws_bytes, qb_measure_ws_bytes(B));
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
    if (stride == 1) launch_dw<1>(out_stats != nullptr, (unsigned)grid, st, x_codes, w_codes, out, g, ep);
    else launch_dw<2>(out_stats != nullptr, (unsigned)grid, st, x_codes, w_codes, out, g, ep);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// Depthwise BACKWARD (conv2d_biprec, quantize.py:142-148, for groups == C): fp32 streaming kernels over NHWC tensors.
//   input gradient  dX[b,hi,wi,c] = SUM_t gq[b,ho,wo,c] * w_hat[c,t]     (gq = the num_bits_grad-quantized gradient)
//   weight gradient dW[c,t] = SUM_{b,ho,wo} x_hat[b,ho*s-p+kh,wo*s-p+kw,c] * g[b,ho,wo,c],   db[c] = SUM g
// Both read each tensor once from HBM (9x window re-use through L1) - 8 B/elem for dgrad, 8 B/elem + the partial blocks
// for wgrad - where cuDNN's grouped kernels and a separate bias reduction took 5.2 ms per MobileNet step.  Thread = one
// pixel x 4 channels with a thread-constant channel group (weights / accumulators in registers), CTA = contiguous item
// range advanced without divisions.  The weight gradient is reduced deterministically: per-thread registers -> fixed-order
// sum over the CTA's replicas of a channel group in shared memory -> per-CTA partial block -> second kernel sums the
// blocks in CTA order.
// ------------------------------------------------------------------------------------------------
constexpr int DW_STRIP = 4;             // stride-1 backward kernels: adjacent pixels per thread sharing the window loads

struct DwbGeom {
    int B, H, W, C, Ho, Wo, stride, pad, C4, Wq;
    uint32_t items, per_cta;
    int dcg, dq;
};

struct DwPos {
    int cg, w, h, b;
    __device__ __forceinline__ void init(uint32_t item, int C4, int Wd, int Hd) {
        const uint32_t q = item / (uint32_t)C4;
        cg = (int)(item - q * C4);
        const uint32_t row = q / (uint32_t)Wd;
        w = (int)(q - row * Wd);
        b = (int)(row / (uint32_t)Hd);
        h = (int)(row - (uint32_t)b * Hd);
    }
    __device__ __forceinline__ void advance(int dcg, int dq, int C4, int Wd, int Hd) {
        cg += dcg;
        if (cg >= C4) { cg -= C4; ++dq; }
        w += dq;
        if (w >= Wd) {
            const uint32_t rows = (uint32_t)w / (uint32_t)Wd;
            w -= (int)(rows * Wd);
            const uint32_t hh = (uint32_t)h + rows;
            const uint32_t db = hh / (uint32_t)Hd;
            h = (int)(hh - db * Hd);
            b += (int)db;
        }
    }
};

template <int STRIDE>
__global__ void __launch_bounds__(DW_THREADS, 2)
k_conv_dw_dgrad(const float* __restrict__ gq, const float* __restrict__ w, float* __restrict__ dx, DwbGeom g) {
    float wr[9][4];
    int cur_cg = -1;
    const uint32_t begin = blockIdx.x * g.per_cta;
    const uint32_t end = (g.items - begin > g.per_cta) ? begin + g.per_cta : g.items;
    DwPos p;
    p.init(begin + threadIdx.x, g.C4, g.W, g.H);
    for (uint32_t base = begin; base < end; base += DW_THREADS) {
        if (base + threadIdx.x < end) {
            if (p.cg != cur_cg) {
                cur_cg = p.cg;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch)
#pragma unroll
                    for (int t = 0; t < 9; ++t) wr[t][ch] = __ldg(w + (int64_t)(p.cg * 4 + ch) * 9 + t);
            }
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            const float* gb = gq + (int64_t)p.b * g.Ho * g.Wo * g.C + p.cg * 4;
            const int h0 = p.h + g.pad - 2, w0 = p.w + g.pad - 2;         // stride 1: first contributing output row / column
            if (STRIDE == 1 && h0 >= 0 && w0 >= 0 && h0 + 2 < g.Ho && w0 + 2 < g.Wo) {
                float4 v[9];                                              // interior: nine unconditional loads in flight
                const float* g0 = gb + ((int64_t)h0 * g.Wo + w0) * g.C;
#pragma unroll
                for (int t = 0; t < 9; ++t)
                    v[t] = __ldg(reinterpret_cast<const float4*>(g0 + ((int64_t)(t / 3) * g.Wo + t % 3) * g.C));
#pragma unroll
                for (int t = 0; t < 9; ++t) {                             // output (h0 + r, w0 + c) meets tap (2 - r, 2 - c)
                    const int k = (2 - t / 3) * 3 + (2 - t % 3);
                    acc.x = fmaf(v[t].x, wr[k][0], acc.x);
                    acc.y = fmaf(v[t].y, wr[k][1], acc.y);
                    acc.z = fmaf(v[t].z, wr[k][2], acc.z);
                    acc.w = fmaf(v[t].w, wr[k][3], acc.w);
                }
            } else
#pragma unroll
            for (int kh = 0; kh < 3; ++kh) {
                const int th = p.h + g.pad - kh;
                if (th < 0 || (STRIDE == 2 && (th & 1))) continue;
                const int ho = th / STRIDE;
                if (ho >= g.Ho) continue;
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    const int tw = p.w + g.pad - kw;
                    if (tw < 0 || (STRIDE == 2 && (tw & 1))) continue;
                    const int wo = tw / STRIDE;
                    if (wo >= g.Wo) continue;
                    const float4 v = __ldg(reinterpret_cast<const float4*>(gb + ((int64_t)ho * g.Wo + wo) * g.C));
                    acc.x = fmaf(v.x, wr[kh * 3 + kw][0], acc.x);
                    acc.y = fmaf(v.y, wr[kh * 3 + kw][1], acc.y);
                    acc.z = fmaf(v.z, wr[kh * 3 + kw][2], acc.z);
                    acc.w = fmaf(v.w, wr[kh * 3 + kw][3], acc.w);
                }
            }
            __stcs(reinterpret_cast<float4*>(dx + (((int64_t)p.b * g.H + p.h) * g.W + p.w) * g.C + p.cg * 4), acc);
        }
        p.advance(g.dcg, g.dq, g.C4, g.W, g.H);
    }
}

// partial[cta][C][10]: 9 taps + bias gradient
template <int STRIDE>
__global__ void __launch_bounds__(DW_THREADS, 2)
k_conv_dw_wgrad(const float* __restrict__ gr, const float* __restrict__ xq, float* __restrict__ partial, DwbGeom g) {
    extern __shared__ float s_red[];                                  // [10*4][DW_THREADS]
    float acc[10][4];
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) acc[t][ch] = 0.f;
    const uint32_t begin = blockIdx.x * g.per_cta;
    const uint32_t end = (g.items - begin > g.per_cta) ? begin + g.per_cta : g.items;
    DwPos p;
    p.init(begin + threadIdx.x, g.C4, g.Wo, g.Ho);
    const int cg = p.cg;                                               // constant: DW_THREADS % C4 == 0 (host-checked)
    for (uint32_t base = begin; base < end; base += DW_THREADS) {
        if (base + threadIdx.x < end) {
            const float4 gv = __ldcs(reinterpret_cast<const float4*>(
                gr + (((int64_t)p.b * g.Ho + p.h) * g.Wo + p.w) * g.C + cg * 4));
            acc[9][0] += gv.x; acc[9][1] += gv.y; acc[9][2] += gv.z; acc[9][3] += gv.w;
            const int hi0 = p.h * STRIDE - g.pad, wi0 = p.w * STRIDE - g.pad;
            const float* xb = xq + (((int64_t)p.b * g.H + hi0) * g.W + wi0) * g.C + cg * 4;
            if (hi0 >= 0 && wi0 >= 0 && hi0 + 2 < g.H && wi0 + 2 < g.W) {
                float4 xv[9];                                             // interior: nine unconditional loads in flight
#pragma unroll
                for (int t = 0; t < 9; ++t)
                    xv[t] = __ldg(reinterpret_cast<const float4*>(xb + ((int64_t)(t / 3) * g.W + t % 3) * g.C));
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    acc[t][0] = fmaf(xv[t].x, gv.x, acc[t][0]);
                    acc[t][1] = fmaf(xv[t].y, gv.y, acc[t][1]);
                    acc[t][2] = fmaf(xv[t].z, gv.z, acc[t][2]);
                    acc[t][3] = fmaf(xv[t].w, gv.w, acc[t][3]);
                }
            } else
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const int hi = hi0 + t / 3, wi = wi0 + t % 3;
                if (hi >= 0 && hi < g.H && wi >= 0 && wi < g.W) {
                    const float4 xv = __ldg(reinterpret_cast<const float4*>(xb + ((int64_t)(t / 3) * g.W + t % 3) * g.C));
                    acc[t][0] = fmaf(xv.x, gv.x, acc[t][0]);
                    acc[t][1] = fmaf(xv.y, gv.y, acc[t][1]);
                    acc[t][2] = fmaf(xv.z, gv.z, acc[t][2]);
                    acc[t][3] = fmaf(xv.w, gv.w, acc[t][3]);
                }
            }
        }
        p.advance(0, g.dq, g.C4, g.Wo, g.Ho);
    }
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) s_red[(t * 4 + ch) * DW_THREADS + threadIdx.x] = acc[t][ch];
    __syncthreads();
    // thread j < C4*40 sums value v = j / C4 of channel group j % C4 over the CTA's replicas (threads cg, cg+C4, ...)
    float* dst = partial + (int64_t)blockIdx.x * g.C * 10;
    for (int j = threadIdx.x; j < g.C4 * 40; j += DW_THREADS) {
        const int v = j / g.C4, c4 = j - v * g.C4;
        float s = 0.f;
        for (int r = c4; r < DW_THREADS; r += g.C4) s += s_red[v * DW_THREADS + r];
        dst[(c4 * 4 + (v & 3)) * 10 + (v >> 2)] = s;                   // [c][tap 0..8, bias]
    }
}

// stride 1: a thread owns DW_STRIP horizontally adjacent pixels x 4 channels and walks the three window rows, so a row of
// DW_STRIP + 2 float4 loads (zero-filled outside the tensor: no separate border path) serves 3 * DW_STRIP products:
// 4.5 L1 loads per pixel instead of 9
__global__ void __launch_bounds__(DW_THREADS, 2)
k_conv_dw_dgrad_strip(const float* __restrict__ gq, const float* __restrict__ w, float* __restrict__ dx, DwbGeom g) {
    constexpr int P = DW_STRIP;
    float wr[9][4];
    int cur_cg = -1;
    const uint32_t begin = blockIdx.x * g.per_cta;
    const uint32_t end = (g.items - begin > g.per_cta) ? begin + g.per_cta : g.items;
    DwPos p;
    p.init(begin + threadIdx.x, g.C4, g.Wq, g.H);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t base = begin; base < end; base += DW_THREADS) {
        if (base + threadIdx.x < end) {
            if (p.cg != cur_cg) {
                cur_cg = p.cg;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch)
#pragma unroll
                    for (int t = 0; t < 9; ++t) wr[t][ch] = __ldg(w + (int64_t)(p.cg * 4 + ch) * 9 + t);
            }
            const int wbase = p.w * P;
            float4 acc[P];
#pragma unroll
            for (int i = 0; i < P; ++i) acc[i] = zero;
            const float* gb = gq + (int64_t)p.b * g.Ho * g.Wo * g.C + p.cg * 4;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int ho = p.h + g.pad - 2 + r;
                if (ho < 0 || ho >= g.Ho) continue;
                float4 v[P + 2];
#pragma unroll
                for (int c = 0; c < P + 2; ++c) {
                    const int wo = wbase + g.pad - 2 + c;
                    v[c] = (wo >= 0 && wo < g.Wo) ? __ldg(reinterpret_cast<const float4*>(gb + ((int64_t)ho * g.Wo + wo) * g.C)) : zero;
                }
#pragma unroll
                for (int i = 0; i < P; ++i)
#pragma unroll
                    for (int c3 = 0; c3 < 3; ++c3) {                   // output column wbase+i+pad-2+c3 meets tap (2-r, 2-c3)
                        const int k = (2 - r) * 3 + (2 - c3);
                        acc[i].x = fmaf(v[i + c3].x, wr[k][0], acc[i].x);
                        acc[i].y = fmaf(v[i + c3].y, wr[k][1], acc[i].y);
                        acc[i].z = fmaf(v[i + c3].z, wr[k][2], acc[i].z);
                        acc[i].w = fmaf(v[i + c3].w, wr[k][3], acc[i].w);
                    }
            }
            float* ob = dx + (((int64_t)p.b * g.H + p.h) * g.W + wbase) * g.C + p.cg * 4;
#pragma unroll
            for (int i = 0; i < P; ++i)
                if (wbase + i < g.W) __stcs(reinterpret_cast<float4*>(ob + (int64_t)i * g.C), acc[i]);
        }
        p.advance(g.dcg, g.dq, g.C4, g.Wq, g.H);
    }
}

__global__ void __launch_bounds__(DW_THREADS, 2)
k_conv_dw_wgrad_strip(const float* __restrict__ gr, const float* __restrict__ xq, float* __restrict__ partial, DwbGeom g) {
    constexpr int P = DW_STRIP;
    extern __shared__ float s_red[];                                  // [10*4][DW_THREADS]
    float acc[10][4];
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) acc[t][ch] = 0.f;
    const uint32_t begin = blockIdx.x * g.per_cta;
    const uint32_t end = (g.items - begin > g.per_cta) ? begin + g.per_cta : g.items;
    DwPos p;
    p.init(begin + threadIdx.x, g.C4, g.Wq, g.Ho);
    const int cg = p.cg;                                               // constant: DW_THREADS % C4 == 0 (host-checked)
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t base = begin; base < end; base += DW_THREADS) {
        if (base + threadIdx.x < end) {
            const int wbase = p.w * P;
            float4 gv[P];
            const float* gb = gr + (((int64_t)p.b * g.Ho + p.h) * g.Wo + wbase) * g.C + cg * 4;
#pragma unroll
            for (int i = 0; i < P; ++i) {
                gv[i] = (wbase + i < g.Wo) ? __ldcs(reinterpret_cast<const float4*>(gb + (int64_t)i * g.C)) : zero;
                acc[9][0] += gv[i].x; acc[9][1] += gv[i].y; acc[9][2] += gv[i].z; acc[9][3] += gv[i].w;
            }
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int hi = p.h - g.pad + r;
                if (hi < 0 || hi >= g.H) continue;
                const float* xb = xq + (((int64_t)p.b * g.H + hi) * g.W) * g.C + cg * 4;
                float4 xv[P + 2];
#pragma unroll
                for (int c = 0; c < P + 2; ++c) {
                    const int wi = wbase - g.pad + c;
                    xv[c] = (wi >= 0 && wi < g.W) ? __ldg(reinterpret_cast<const float4*>(xb + (int64_t)wi * g.C)) : zero;
                }
#pragma unroll
                for (int i = 0; i < P; ++i)
#pragma unroll
                    for (int kw = 0; kw < 3; ++kw) {
                        acc[r * 3 + kw][0] = fmaf(xv[i + kw].x, gv[i].x, acc[r * 3 + kw][0]);
                        acc[r * 3 + kw][1] = fmaf(xv[i + kw].y, gv[i].y, acc[r * 3 + kw][1]);
                        acc[r * 3 + kw][2] = fmaf(xv[i + kw].z, gv[i].z, acc[r * 3 + kw][2]);
                        acc[r * 3 + kw][3] = fmaf(xv[i + kw].w, gv[i].w, acc[r * 3 + kw][3]);
                    }
            }
        }
        p.advance(0, g.dq, g.C4, g.Wq, g.Ho);
    }
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) s_red[(t * 4 + ch) * DW_THREADS + threadIdx.x] = acc[t][ch];
    __syncthreads();
    float* dst = partial + (int64_t)blockIdx.x * g.C * 10;
    for (int j = threadIdx.x; j < g.C4 * 40; j += DW_THREADS) {
        const int v = j / g.C4, c4 = j - v * g.C4;
        float s = 0.f;
        for (int r = c4; r < DW_THREADS; r += g.C4) s += s_red[v * DW_THREADS + r];
        dst[(c4 * 4 + (v & 3)) * 10 + (v >> 2)] = s;
    }
}

__global__ void __launch_bounds__(256)
k_conv_dw_wgrad_reduce(const float* __restrict__ partial, int nparts, int C, float* __restrict__ dw, float* __restrict__ db) {
    const int i = blockIdx.x * 256 + threadIdx.x;                      // over C*10
    if (i >= C * 10) return;
    float s = 0.f;
    for (int k = 0; k < nparts; ++k) s += partial[(int64_t)k * C * 10 + i];
    const int c = i / 10, t = i - c * 10;
    if (t < 9) dw[c * 9 + t] = s;
    else if (db) db[c] = s;
}

static int dwb_geom(DwbGeom& g, int64_t B, int64_t H, int64_t W, int64_t C, int stride, int pad, bool over_outputs, int ctas_per_sm) {
    const int64_t Ho = (H + 2 * pad - 3) / stride + 1, Wo = (W + 2 * pad - 3) / stride + 1;
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.C = (int)C; g.Ho = (int)Ho; g.Wo = (int)Wo; g.stride = stride; g.pad = pad;
    g.C4 = (int)(C / 4);
    const int64_t Wd = over_outputs ? Wo : W, Hd = over_outputs ? Ho : H;
    g.Wq = (int)(stride == 1 ? (Wd + DW_STRIP - 1) / DW_STRIP : Wd);      // stride 1: strips of DW_STRIP pixels per item
    const int64_t items = B * Hd * g.Wq * g.C4;
    if (items <= 0 || items >= 4294967295LL || Ho < 1 || Wo < 1) return QB_ERR_UNSUPPORTED;
    g.items = (uint32_t)items;
    g.dcg = DW_THREADS % g.C4;
    g.dq = DW_THREADS / g.C4;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const int64_t ctas = (int64_t)sms * ctas_per_sm;
    int64_t per = (items + ctas - 1) / ctas;
    per = (per + DW_THREADS - 1) / DW_THREADS * DW_THREADS;
    g.per_cta = (uint32_t)per;
    return QB_OK;
}

extern "C" int qb_conv_dw_dgrad(const float* grad_q, const float* w_hat, float* dx, int64_t B, int64_t H, int64_t W, int64_t C,
                                int kh, int kw, int stride, int pad, void* stream) {
    QB_CHECK_ARG(grad_q && w_hat && dx, "qb_conv_dw_dgrad: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0, "qb_conv_dw_dgrad: bad shape");
    DwbGeom g;
    if (!qb_conv_dw_supported(C, kh, kw, stride, pad) || dwb_geom(g, B, H, W, C, stride, pad, false, 12) != QB_OK ||
        ((reinterpret_cast<uintptr_t>(grad_q) | reinterpret_cast<uintptr_t>(dx)) & 15) != 0) {
        qb_set_error("qb_conv_dw_dgrad: unsupported geometry (3x3, C%%4==0, stride 1 or 2, 16-byte aligned tensors)");
        return QB_ERR_UNSUPPORTED;
    }
    const unsigned grid = (unsigned)(((int64_t)g.items + g.per_cta - 1) / g.per_cta);
    if (stride == 1) k_conv_dw_dgrad_strip<<<grid, DW_THREADS, 0, qb_stream(stream)>>>(grad_q, w_hat, dx, g);
    else k_conv_dw_dgrad<2><<<grid, DW_THREADS, 0, qb_stream(stream)>>>(grad_q, w_hat, dx, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

static unsigned dw_wgrad_grid(const DwbGeom& g) { return (unsigned)(((int64_t)g.items + g.per_cta - 1) / g.per_cta); }

extern "C" size_t qb_conv_dw_wgrad_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t C, int stride, int pad) {
    DwbGeom g;
    if (C <= 0 || C % 4 != 0 || dwb_geom(g, B, H, W, C, stride, pad, true, 4) != QB_OK) return 0;
    return (size_t)dw_wgrad_grid(g) * (size_t)C * 10 * sizeof(float);
}

extern "C" int qb_conv_dw_wgrad(const float* grad_out, const float* x_hat, float* dw, float* db, int64_t B, int64_t H, int64_t W,
                                int64_t C, int kh, int kw, int stride, int pad, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(grad_out && x_hat && dw && ws, "qb_conv_dw_wgrad: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && C > 0, "qb_conv_dw_wgrad: bad shape");
    DwbGeom g;
    if (!qb_conv_dw_supported(C, kh, kw, stride, pad) || dwb_geom(g, B, H, W, C, stride, pad, true, 4) != QB_OK ||
        g.C4 > DW_THREADS || DW_THREADS % g.C4 != 0 ||
        ((reinterpret_cast<uintptr_t>(grad_out) | reinterpret_cast<uintptr_t>(x_hat)) & 15) != 0) {
        qb_set_error("qb_conv_dw_wgrad: unsupported geometry (3x3, stride 1 or 2, C/4 a divisor of 256, 16-byte aligned tensors)");
        return QB_ERR_UNSUPPORTED;
    }
    const unsigned grid = dw_wgrad_grid(g);
    if (ws_bytes < (size_t)grid * (size_t)C * 10 * sizeof(float)) {
        qb_set_error("qb_conv_dw_wgrad: workspace %zu < %zu bytes", ws_bytes, (size_t)grid * (size_t)C * 10 * sizeof(float));
        return QB_ERR_WORKSPACE;
    }
    const size_t smem = (size_t)40 * DW_THREADS * sizeof(float);
    // per call, not cached: the attribute belongs to the CURRENT device's copy of the function
    if (stride == 1) QB_CUDA(cudaFuncSetAttribute(k_conv_dw_wgrad_strip, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    else QB_CUDA(cudaFuncSetAttribute(k_conv_dw_wgrad<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaStream_t st = qb_stream(stream);
    float* partial = static_cast<float*>(ws);
    if (stride == 1) k_conv_dw_wgrad_strip<<<grid, DW_THREADS, smem, st>>>(grad_out, x_hat, partial, g);
    else k_conv_dw_wgrad<2><<<grid, DW_THREADS, smem, st>>>(grad_out, x_hat, partial, g);
    QB_LAUNCH_CHECK();
    k_conv_dw_wgrad_reduce<<<(unsigned)((C * 10 + 255) / 256), 256, 0, st>>>(partial, (int)grid, (int)C, dw, db);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
