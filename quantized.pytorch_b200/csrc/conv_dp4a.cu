// conv_dp4a.cu - QConv2d forward for the FIRST layer of a network (3 input channels: ResNet 7x7/2, MobileNet 3x3/2), the
// case the tcgen05 implicit GEMM cannot take (C_in % 32 != 0): a direct convolution on the integer codes with DP4A.
//
// Same algebra as conv_i8.cu (reference: F.conv2d on dequantised tensors, models/modules/quantize.py:143-148, :343):
//   out[m,n] = sa*sw[n]*SUM_V qa*qc + sa*lw'[n]*SUM_V qa + la*sw[n]*SUM_V qc[n] + la*lw'[n]*|V|*Cin + bias[n]
// with V the in-bounds filter taps of output pixel m.  An image input is not post-ReLU, so la != 0 and every term is
// live; border pixels get SUM_V qc[n] from a 2-D prefix table of the per-tap weight code sums (4 loads per channel).
//
// With K = kh*kw*3 = 147 the tensor pipe is irrelevant; the work is 8.6 G lane-DP4As and a 0.8 GB output stream:
//   * a filter ROW (kw*Cin <= 24 bytes) is contiguous both in the NHWC codes and in the [Cout][tap][Cin] weights, so it is
//     <= 6 packed words; a thread keeps the whole KH x WPR-word patch of PPT output pixels in registers
//   * weights sit in shared memory as [Cout][KH][32 B]; all threads of a warp read the same address (broadcast), one
//     LDS.128 + one LDS.64 per 6*PPT DP4As.  A broadcast LDS.128 still delivers 512 B to the register file (4 LSU cycles),
//     so with one pixel per thread the kernel was LSU bound (1.46 ms); weights as constant-bank operands (LDCU per pair
//     of DP4As) were slower still (2.34 ms).  PPT pixels per thread divide the weight traffic by PPT.
//   * k_dp4a_prep packs weights (zero padded rows), epilogue coefficients and prefix tables ONCE per call into a global
//     workspace; the persistent CTAs stage it with 16-byte loads
//   * input patch rows are copied with aligned 16-byte loads (byte path only for chunks that straddle an image edge)
//   * every thread writes its results as 16-byte stores (4 channels); the halves of a 32-byte sector are written by
//     consecutive instructions of the same thread and merge in L2
// cuDNN's kernel for this layer (sm80_xmma ... indexed_wo_smem, TF32) took 1.25-1.32 ms at 256x3x224x224, plus its
// NCHW<->NHWC conversions in the training step.
#include "common.cuh"

constexpr int CD_THREADS = 128;
constexpr int CD_TW = 32, CD_ROWS = CD_THREADS / CD_TW;     // a CTA tile is (CD_ROWS * PPT) x 32 output pixels
constexpr int CD_MAX_COUT = 128, CD_PT = 64;                // prefix table: (kh+1) x (kw+1) <= 64 entries

struct Dp4aGeom {
    int B, H, W, Cin, Ho, Wo, Cout, kh, kw, stride, pad;
    int tiles_h, tiles_w, rows_in, pitch_in;       // input patch: rows_in x pitch_in bytes (pitch % 16 == 0)
    FastDiv d_tw, d_th, d_chunks;
};

struct Dp4aEpi {
    const float* a_range; int a_bits;
    const float* w_row_min; const float* w_row_max; int w_bits;
    const int32_t* tapsum;     // [Cout][kh*kw] sums over Cin of the centred weight codes
    const float* bias;
};

// workspace layout (bytes): [ w: Cout*KH*32 | cf: float4[Cout] | c34: float2[Cout] | P: int32[Cout][CD_PT] ]
struct Dp4aWs {
    uint32_t* w; float4* cf; float2* c34; int32_t* P;
};
__host__ __device__ inline size_t dp4a_ws_bytes(int Cout, int KH) { return (size_t)Cout * ((size_t)KH * 32 + 16 + 8 + CD_PT * 4); }
__host__ __device__ inline Dp4aWs dp4a_ws(void* base, int Cout, int KH) {
    Dp4aWs s;
    unsigned char* b = static_cast<unsigned char*>(base);
    s.w = reinterpret_cast<uint32_t*>(b);
    s.cf = reinterpret_cast<float4*>(b + (size_t)Cout * KH * 32);
    s.c34 = reinterpret_cast<float2*>(b + (size_t)Cout * (KH * 32 + 16));
    s.P = reinterpret_cast<int32_t*>(b + (size_t)Cout * (KH * 32 + 24));
    return s;
}

__device__ __forceinline__ int dp4a_us(uint32_t a_u8x4, uint32_t b_s8x4, int c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a_u8x4), "r"(b_s8x4), "r"(c));
    return d;
}

__global__ void __launch_bounds__(256) k_dp4a_prep(void* ws, const int8_t* __restrict__ w, Dp4aGeom g, Dp4aEpi ep) {
    const Dp4aWs d = dp4a_ws(ws, g.Cout, g.kh);
    const int tid = threadIdx.x, KH = g.kh;
    const int row_bytes = g.kw * g.Cin, taps = KH * g.kw, PW = g.kw + 1;
    for (int i = tid; i < g.Cout * KH * 8; i += 256) {
        const int nk = i >> 3, j = i & 7;                     // nk = n*KH + kh
        uint32_t v = 0;
        for (int bt = 0; bt < 4; ++bt) {
            const int col = j * 4 + bt;
            if (col < row_bytes) v |= (uint32_t)(uint8_t)w[(size_t)nk * row_bytes + col] << (8 * bt);
        }
        d.w[i] = v;
    }
    const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
    const double sa = (double)pa.scale, la = (double)pa.lo;
    for (int n = tid; n < g.Cout; n += 256) {
        int32_t* P = d.P + (size_t)n * CD_PT;                 // P[a][b] = sum of taps with kh < a, kw < b
        for (int b = 0; b < PW; ++b) P[b] = 0;
        int tall = 0;
        for (int a = 1; a <= KH; ++a) {
            P[a * PW] = 0;
            int run = 0;
            for (int b = 1; b <= g.kw; ++b) {
                const int v = __ldg(ep.tapsum + (size_t)n * taps + (a - 1) * g.kw + (b - 1));
                run += v;
                tall += v;
                P[a * PW + b] = P[(a - 1) * PW + b] + run;
            }
        }
        const QParams pw = make_qparams_f(__ldg(ep.w_row_min + n), __ldg(ep.w_row_max + n), ep.w_bits);
        const double swd = (double)pw.scale, lw = (double)pw.lo + (double)(1 << (ep.w_bits - 1)) * (double)pw.scale;
        const float bias = ep.bias ? __ldg(ep.bias + n) : 0.f;
        const float c3 = (float)(la * swd), c4 = (float)(la * lw);
        d.cf[n] = make_float4((float)(sa * swd), (float)(sa * lw), fmaf(c3, (float)tall, c4 * (float)(taps * g.Cin)) + bias, bias);
        d.c34[n] = make_float2(c3, c4);
    }
}

template <int KH, int WPR, int PPT>      // filter rows, 32-bit words per filter row, output pixels per thread
__global__ void __launch_bounds__(CD_THREADS, PPT >= 4 ? 2 : 3)
k_conv_dp4a(const uint8_t* __restrict__ x, float* __restrict__ out, const void* __restrict__ ws, Dp4aGeom g) {
    extern __shared__ __align__(16) unsigned char smem[];
    const size_t ws_bytes = dp4a_ws_bytes(g.Cout, KH);
    const Dp4aWs sw = dp4a_ws(smem, g.Cout, KH);                                           // same layout as the workspace
    uint32_t* s_in = reinterpret_cast<uint32_t*>(smem + ((ws_bytes + 15) & ~(size_t)15));   // rows_in x pitch_in bytes
    const int tid = threadIdx.x;
    for (int i = tid; i < (int)(ws_bytes / 16); i += CD_THREADS)
        reinterpret_cast<uint4*>(smem)[i] = __ldg(reinterpret_cast<const uint4*>(ws) + i);
    const int row_bytes = g.kw * g.Cin;                       // <= 4 * WPR
    const int PW = g.kw + 1;
    const int th = tid / CD_TW, tw = tid - th * CD_TW;
    const int row_len = g.W * g.Cin, chunks = g.pitch_in >> 4;
    const int tiles_per_img = g.tiles_h * g.tiles_w, total = g.B * tiles_per_img;
    const uint32_t last_mask = (row_bytes & 3) ? ((1u << (8 * (row_bytes & 3))) - 1u) : 0xffffffffu;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
        uint32_t q = fdiv((uint32_t)t, g.d_tw);
        const int twi = t - (int)q * g.tiles_w;
        const uint32_t b = fdiv(q, g.d_th);
        const int thi = (int)q - (int)b * g.tiles_h;
        const int ho0 = thi * (CD_ROWS * PPT), wo0 = twi * CD_TW;
        const int h_start = ho0 * g.stride - g.pad, w_start = wo0 * g.stride - g.pad;
        __syncthreads();                                      // previous tile's patch is consumed (and the staged operands are visible)
        // ---- input patch -> shared memory: row r holds the image bytes from the 16-byte aligned address at or below its
        //      first needed byte; zero outside the image ----
        for (int i = tid; i < g.rows_in * chunks; i += CD_THREADS) {
            const int r = (int)fdiv((uint32_t)i, g.d_chunks), c = i - r * chunks;
            const int h = h_start + r;
            const long long row_addr = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h) * row_len;
            const long long first = row_addr + (long long)w_start * g.Cin;
            const long long a0 = (first & ~15LL) + 16LL * c;                       // global address of this chunk
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (h >= 0 && h < g.H) {
                if (a0 >= row_addr && a0 + 16 <= row_addr + row_len) {
                    v = __ldg(reinterpret_cast<const uint4*>(a0));
                } else if (a0 + 16 > row_addr && a0 < row_addr + row_len) {        // straddles an image edge
                    uint32_t wd[4] = {0u, 0u, 0u, 0u};
                    for (int bt = 0; bt < 16; ++bt) {
                        const long long a = a0 + bt;
                        if (a >= row_addr && a < row_addr + row_len)
                            wd[bt >> 2] |= (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(a)) << (8 * (bt & 3));
                    }
                    v = make_uint4(wd[0], wd[1], wd[2], wd[3]);
                }
            }
            *reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(s_in) + (size_t)r * g.pitch_in + 16 * c) = v;
        }
        __syncthreads();
        // ---- this thread's patches: PPT pixels x KH rows of WPR words, re-aligned from the byte offset of the pixel ----
        uint32_t p[PPT][KH][WPR];
        float fs_a[PPT];
#pragma unroll
        for (int i = 0; i < PPT; ++i) {
            int s_a = 0;
#pragma unroll
            for (int kh = 0; kh < KH; ++kh) {
                const int r = (th + CD_ROWS * i) * g.stride + kh;
                const long long first = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h_start + r) * row_len +
                                        (long long)w_start * g.Cin;
                const int o = tw * g.stride * g.Cin + (int)(first & 15LL);
                const int sh = (o & 3) * 8;
                const uint32_t* rowp = s_in + ((size_t)r * g.pitch_in + (o & ~3)) / 4;
                uint32_t prev = rowp[0];
#pragma unroll
                for (int j = 0; j < WPR; ++j) {
                    const uint32_t nxt = rowp[j + 1];
                    uint32_t v = __funnelshift_r(prev, nxt, sh);
                    if (j == WPR - 1) v &= last_mask;         // bytes beyond the filter row belong to the next pixel
                    p[i][kh][j] = v;
                    s_a = dp4a_us(v, 0x01010101u, s_a);
                    prev = nxt;
                }
            }
            fs_a[i] = (float)s_a;
        }
        // in-bounds tap rectangle of each pixel
        const int wo = wo0 + tw;
        const int w0 = wo * g.stride - g.pad;
        const int kc = w0 < 0 ? -w0 : 0, kd = (w0 + g.kw > g.W) ? g.W - w0 : g.kw;         // kw in [kc, kd)
        int i00[PPT], i01[PPT], i10[PPT], i11[PPT];
        float kvalid[PPT];
        bool interior = true, valid[PPT];
        float* optr[PPT];
#pragma unroll
        for (int i = 0; i < PPT; ++i) {
            const int ho = ho0 + th + CD_ROWS * i;
            const int h0 = ho * g.stride - g.pad;
            int ka = h0 < 0 ? -h0 : 0, kb = (h0 + KH > g.H) ? g.H - h0 : KH;               // kh in [ka, kb)
            valid[i] = ho < g.Ho && wo < g.Wo;
            if (!valid[i]) { ka = 0; kb = KH; }
            interior = interior && ka == 0 && kb == KH;
            kvalid[i] = (float)((kb - ka) * (kd - kc) * g.Cin);
            i00[i] = ka * PW + kc; i01[i] = ka * PW + kd; i10[i] = kb * PW + kc; i11[i] = kb * PW + kd;
            optr[i] = out + (((size_t)b * g.Ho + ho) * g.Wo + wo) * g.Cout;
        }
        interior = interior && ((kc == 0 && kd == g.kw) || wo >= g.Wo);
        for (int n4 = 0; n4 < g.Cout; n4 += 4) {
            float res[PPT][4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int n = n4 + u;
                const uint4* wr = reinterpret_cast<const uint4*>(sw.w + (size_t)n * KH * 8);
                int acc[PPT][2];
#pragma unroll
                for (int i = 0; i < PPT; ++i) acc[i][0] = acc[i][1] = 0;
#pragma unroll
                for (int kh = 0; kh < KH; ++kh) {
                    const uint4 wa = wr[kh * 2];
                    uint2 wb = make_uint2(0u, 0u);
                    if (WPR > 4) wb = *reinterpret_cast<const uint2*>(&wr[kh * 2 + 1]);
                    const uint32_t wv[6] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y};
                    // word-major, pixel-minor: consecutive DP4As belong to different accumulator chains (2 * PPT of them)
#pragma unroll
                    for (int j = 0; j < WPR; ++j)
#pragma unroll
                        for (int i = 0; i < PPT; ++i) acc[i][j & 1] = dp4a_us(p[i][kh][j], wv[j], acc[i][j & 1]);
                }
                const float4 cf = sw.cf[n];
#pragma unroll
                for (int i = 0; i < PPT; ++i) {
                    float add = cf.z;
                    if (!interior) {
                        const int32_t* P = sw.P + (size_t)n * CD_PT;
                        const int tsum = P[i11[i]] - P[i01[i]] - P[i10[i]] + P[i00[i]];
                        const float2 c34 = sw.c34[n];
                        add = fmaf(c34.x, (float)tsum, c34.y * kvalid[i]) + cf.w;
                    }
                    res[i][u] = fmaf(cf.x, (float)(acc[i][0] + acc[i][1]), fmaf(cf.y, fs_a[i], add));
                }
            }
#pragma unroll
            for (int i = 0; i < PPT; ++i)
                if (valid[i]) *reinterpret_cast<float4*>(optr[i] + n4) = make_float4(res[i][0], res[i][1], res[i][2], res[i][3]);
        }
    }
}

extern "C" int qb_conv_dp4a_supported(int64_t Cin, int64_t Cout, int kh, int kw, int stride) {
    const int64_t rb = (int64_t)kw * Cin;
    const bool shape = (kh == 7 && rb > 20 && rb <= 24) || (kh == 3 && rb > 8 && rb <= 12);
    return (shape && Cin >= 1 && Cin <= 8 && Cout >= 4 && Cout <= CD_MAX_COUT && Cout % 4 == 0 && stride >= 1 && stride <= 2) ? 1 : 0;
}

extern "C" size_t qb_conv_dp4a_ws_bytes(int64_t Cout, int kh) {
    return (Cout > 0 && Cout <= CD_MAX_COUT && kh > 0 && kh <= 7) ? dp4a_ws_bytes((int)Cout, kh) : 0;
}

template <int KH, int WPR, int PPT>
static int dp4a_launch(const uint8_t* x, float* out, const void* ws, Dp4aGeom& g, cudaStream_t st) {
    g.tiles_h = (g.Ho + CD_ROWS * PPT - 1) / (CD_ROWS * PPT);
    g.tiles_w = (g.Wo + CD_TW - 1) / CD_TW;
    g.rows_in = (CD_ROWS * PPT - 1) * g.stride + KH;
    // 15 bytes of alignment slack + the bytes of 32 pixels' windows + one extra word read by the re-alignment
    g.pitch_in = (15 + ((CD_TW - 1) * g.stride) * g.Cin + (WPR + 1) * 4 + 15) & ~15;
    g.d_tw = make_fastdiv((uint32_t)g.tiles_w);
    g.d_th = make_fastdiv((uint32_t)g.tiles_h);
    g.d_chunks = make_fastdiv((uint32_t)(g.pitch_in >> 4));
    const size_t smem = ((dp4a_ws_bytes(g.Cout, KH) + 15) & ~(size_t)15) + (size_t)g.rows_in * g.pitch_in;
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const int per_sm = PPT >= 4 ? 2 : 3;
    const int64_t total = (int64_t)g.B * g.tiles_h * g.tiles_w;
    const unsigned grid = (unsigned)(total < (int64_t)per_sm * sms ? total : (int64_t)per_sm * sms);
    QB_CUDA(cudaFuncSetAttribute(k_conv_dp4a<KH, WPR, PPT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_dp4a<KH, WPR, PPT><<<grid, CD_THREADS, smem, st>>>(x, out, ws, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_dp4a_fwd(const uint8_t* x_codes, const int8_t* w_codes_khwc, float* out, int64_t B, int64_t H, int64_t W,
                                int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                                const float* w_row_min, const float* w_row_max, int w_bits, const int32_t* tapsum,
                                const float* bias, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x_codes && w_codes_khwc && out && a_range && w_row_min && w_row_max && tapsum && ws,
                 "qb_conv_dp4a_fwd: null pointer");
    QB_CHECK_ARG(B > 0 && H > 0 && W > 0 && pad >= 0, "qb_conv_dp4a_fwd: bad shape");
    if (!qb_conv_dp4a_supported(Cin, Cout, kh, kw, stride)) {
        qb_set_error("qb_conv_dp4a_fwd: unsupported geometry Cin=%lld Cout=%lld k=%dx%d stride=%d (7x7 with 21..24-byte rows or 3x3 "
                     "with 9..12-byte rows, Cout %% 4 == 0 and <= 128, stride <= 2)", (long long)Cin, (long long)Cout, kh, kw, stride);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8 && w_bits >= 1 && w_bits <= 8, "qb_conv_dp4a_fwd: codes are 8-bit");
    QB_CHECK_ARG(ws_bytes >= qb_conv_dp4a_ws_bytes(Cout, kh) && (reinterpret_cast<uintptr_t>(ws) & 15) == 0,
                 "qb_conv_dp4a_fwd: workspace too small or not 16-byte aligned");
    QB_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "qb_conv_dp4a_fwd: out must be 16-byte aligned");
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    QB_CHECK_ARG(Ho > 0 && Wo > 0 && B * Ho * Wo < 2147483647LL / 64, "qb_conv_dp4a_fwd: output size");
    Dp4aGeom g;
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.Cin = (int)Cin; g.Ho = (int)Ho; g.Wo = (int)Wo; g.Cout = (int)Cout;
    g.kh = kh; g.kw = kw; g.stride = stride; g.pad = pad;
    Dp4aEpi ep{a_range, a_bits, w_row_min, w_row_max, w_bits, tapsum, bias};
    cudaStream_t st = qb_stream(stream);
    k_dp4a_prep<<<1, 256, 0, st>>>(ws, w_codes_khwc, g, ep);
    QB_LAUNCH_CHECK();
    static const int ppt_env = [] { const char* e = getenv("QB_DP4A_PPT"); return e ? atoi(e) : 0; }();
    if (kh == 7) {
        if (ppt_env == 1) return dp4a_launch<7, 6, 1>(x_codes, out, ws, g, st);
        if (ppt_env == 4) return dp4a_launch<7, 6, 4>(x_codes, out, ws, g, st);
        return dp4a_launch<7, 6, 2>(x_codes, out, ws, g, st);
    }
    if (ppt_env == 1) return dp4a_launch<3, 3, 1>(x_codes, out, ws, g, st);
    if (ppt_env == 2) return dp4a_launch<3, 3, 2>(x_codes, out, ws, g, st);
    return dp4a_launch<3, 3, 4>(x_codes, out, ws, g, st);
}
