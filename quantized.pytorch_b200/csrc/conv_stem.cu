// conv_stem.cu - QConv2d forward for the FIRST layer of a network (3 input channels: ResNet 7x7/2, MobileNet 3x3/2), the
// case the TMA-fed implicit GEMM of conv_i8.cu cannot take (C_in % 32 != 0, so a filter tap is not a 16-byte TMA row).
//
// Same algebra as conv_i8.cu (reference: F.conv2d on dequantised tensors, models/modules/quantize.py:143-148, :343):
//   out[m,n] = sa*sw[n]*SUM_V qa*qc + sa*lw'[n]*SUM_V qa + la*sw[n]*SUM_V qc[n] + la*lw'[n]*|V|*Cin + bias[n]
// with V the in-bounds filter taps of output pixel m.  An image input is not post-ReLU, so la != 0 and every term is live.
//
// Structure: tcgen05 kind::i8 with a THREAD-BUILT im2col operand.
//   * a filter ROW (kw*Cin <= 24 bytes) is contiguous both in the NHWC codes and in the [Cout][tap][Cin] weights, so
//     K is laid out as KH rows of WPR 32-bit words (zero padded in the weights): 168 bytes for 7x7x3, 36 for 3x3x3
//   * CTA = 128 threads = 128 output pixels (4 rows x 32) = the 128 rows of the MMA = the 128 TMEM lanes.  Each thread
//     re-aligns the KH x WPR words of ITS pixel from the shared-memory input patch (funnel shifts), takes the code sum
//     SUM_V qa with DP4A on the way, and writes its A row in the 128-byte swizzled K-major layout; one thread then issues
//     <= 6 tcgen05.mma (M=128, N=Cout, K=32) against the resident, pre-swizzled weight operand
//   * the same thread reads ITS accumulator row back (tcgen05.ld 32x32b), applies the dequantising epilogue and writes
//     16-byte stores.  The zero-padding terms depend only on which filter rows / columns are in bounds: the host lists
//     the <= 8 x 8 distinct cases, k_stem_prep tabulates R[case][n] = la*sw*SUM_V qc + la*lw'*|V|*Cin + bias, and the
//     epilogue is two FMAs per output for interior and border pixels alike
//   * no intra-CTA pipelining: 3-4 CTAs per SM overlap each other's phases
// History (256x3x224x224 -> 64 channels): cuDNN fp32 1.32 ms (+ layout conversions in the step); direct DP4A convolution,
// weights broadcast from shared memory 1.46 ms (LSU bound), 2 pixels per thread 0.93 ms (IDP.4A runs on the half-rate
// fmaheavy pipe: 8.6 G lane-ops is a 0.54 ms floor); this kernel: see profiles/.
#include "tc_common.cuh"

constexpr int CS_THREADS = 128;
constexpr int CS_TW = 32, CS_TH = 4;
constexpr int CS_MAX_COUT = 128, CS_MAX_CLS = 8;
constexpr uint32_t CS_TMEM_COLS = 128;

struct StemCls {                 // distinct in-bounds tap ranges [a, b) along one axis
    int n;
    signed char a[CS_MAX_CLS], b[CS_MAX_CLS];
};

struct StemGeom {
    int B, H, W, Cin, Ho, Wo, Cout, kh, kw, stride, pad;
    int tiles_h, tiles_w, rows_in, pitch_in;       // input patch: rows_in x pitch_in bytes (pitch % 16 == 0)
    int katoms, ksteps;                            // 128-byte K atoms per operand row, 32-byte MMA steps
    FastDiv d_tw, d_th, d_chunks;
    StemCls ch, cw;
};

struct StemEpi {
    const float* a_range; int a_bits;
    const float* w_row_min; const float* w_row_max; int w_bits;
    const int32_t* tapsum;     // [Cout][kh*kw] sums over Cin of the centred weight codes
    const float* bias;
};

// workspace = what every CTA stages: [ B operand, swizzled: katoms x Cout x 128 B | cf: float2[Cout] | R: float[nh*nw][Cout] ]
__host__ __device__ inline size_t stem_ws_bytes(int Cout, int katoms, int ncls) {
    return (size_t)katoms * Cout * 128 + (size_t)Cout * 8 + (size_t)ncls * Cout * 4;
}
__host__ __device__ inline uint32_t stem_swz(uint32_t row, uint32_t kbyte, uint32_t rows_per_atom) {   // byte offset of (row, k)
    const uint32_t atom = kbyte >> 7, kk = kbyte & 127u;
    return atom * rows_per_atom * 128u + row * 128u + ((((kk >> 4) ^ (row & 7u)) << 4) | (kk & 15u));
}

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(256) k_stem_prep(unsigned char* ws, const int8_t* __restrict__ w, StemGeom g, StemEpi ep, int WPR) {
    const int tid = threadIdx.x, KH = g.kh;
    const int row_bytes = g.kw * g.Cin, taps = KH * g.kw;
    const size_t b_bytes = (size_t)g.katoms * g.Cout * 128;
    float2* cf = reinterpret_cast<float2*>(ws + b_bytes);
    float* R = reinterpret_cast<float*>(ws + b_bytes + (size_t)g.Cout * 8);
    for (size_t i = tid; i < b_bytes / 4; i += 256) reinterpret_cast<uint32_t*>(ws)[i] = 0u;
    __syncthreads();
    for (int i = tid; i < g.Cout * KH * WPR; i += 256) {
        const int n = i / (KH * WPR), r = i - n * (KH * WPR);
        const int kh = r / WPR, j = r - kh * WPR;
        uint32_t v = 0;
        for (int bt = 0; bt < 4; ++bt) {
            const int col = j * 4 + bt;
            if (col < row_bytes) v |= (uint32_t)(uint8_t)w[((size_t)n * KH + kh) * row_bytes + col] << (8 * bt);
        }
        *reinterpret_cast<uint32_t*>(ws + stem_swz((uint32_t)n, (uint32_t)r * 4u, (uint32_t)g.Cout)) = v;
    }
    const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
    const double sa = (double)pa.scale, la = (double)pa.lo;
    const int ncls = g.ch.n * g.cw.n;
    for (int i = tid; i < ncls * g.Cout; i += 256) {
        const int cls = i / g.Cout, n = i - cls * g.Cout;
        const int ih = cls / g.cw.n, iw = cls - ih * g.cw.n;
        const QParams pw = make_qparams_f(__ldg(ep.w_row_min + n), __ldg(ep.w_row_max + n), ep.w_bits);
        const double swd = (double)pw.scale, lw = (double)pw.lo + (double)(1 << (ep.w_bits - 1)) * (double)pw.scale;
        int tsum = 0;
        for (int a = g.ch.a[ih]; a < g.ch.b[ih]; ++a)
            for (int b = g.cw.a[iw]; b < g.cw.b[iw]; ++b) tsum += __ldg(ep.tapsum + (size_t)n * taps + a * g.kw + b);
        const float kvalid = (float)((g.ch.b[ih] - g.ch.a[ih]) * (g.cw.b[iw] - g.cw.a[iw]) * g.Cin);
        const float bias = ep.bias ? __ldg(ep.bias + n) : 0.f;
        R[i] = fmaf((float)(la * swd), (float)tsum, (float)(la * lw) * kvalid) + bias;
        if (cls == 0) cf[n] = make_float2((float)(sa * swd), (float)(sa * lw));
    }
}

template <int KH, int WPR>      // filter rows, 32-bit words per filter row
__global__ void __launch_bounds__(CS_THREADS, 3)
k_conv_stem(const uint8_t* __restrict__ x, float* __restrict__ out, const unsigned char* __restrict__ ws, StemGeom g) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int ncls = g.ch.n * g.cw.n;
    const size_t a_bytes = (size_t)g.katoms * 128 * 128;
    const size_t b_bytes = (size_t)g.katoms * g.Cout * 128;                  // multiple of 1024 (Cout % 16 == 0)
    const size_t wsb = stem_ws_bytes(g.Cout, g.katoms, ncls);
    unsigned char* s_A = smem;
    unsigned char* s_B = smem + a_bytes;                                     // staged workspace starts here
    const float2* s_cf = reinterpret_cast<const float2*>(s_B + b_bytes);
    const float* s_R = reinterpret_cast<const float*>(s_B + b_bytes + (size_t)g.Cout * 8);
    unsigned char* s_in = s_B + ((wsb + 15) & ~(size_t)15);
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_in + (((size_t)g.rows_in * g.pitch_in + 15) & ~(size_t)15));
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);

    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (int)(wsb / 16); i += CS_THREADS)
        reinterpret_cast<uint4*>(s_B)[i] = __ldg(reinterpret_cast<const uint4*>(ws) + i);
    for (int i = tid; i < (int)(a_bytes / 16); i += CS_THREADS) reinterpret_cast<uint4*>(s_A)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (tid == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<CS_TMEM_COLS>(tmem_slot);
    fence_proxy_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const uint32_t idesc = umma_idesc(/*S32*/ 2, /*A u8*/ 0, /*B s8*/ 1, 128, (uint32_t)g.Cout);
    const uint32_t a_u32 = smem_u32(s_A), b_u32 = smem_u32(s_B);

    const int row_bytes = g.kw * g.Cin;                       // <= 4 * WPR
    const int th = tid / CS_TW, tw = tid - th * CS_TW;
    const int row_len = g.W * g.Cin, chunks = g.pitch_in >> 4;
    const int total = g.B * g.tiles_h * g.tiles_w;
    const uint32_t last_mask = (row_bytes & 3) ? ((1u << (8 * (row_bytes & 3))) - 1u) : 0xffffffffu;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
        uint32_t q = fdiv((uint32_t)t, g.d_tw);
        const int twi = t - (int)q * g.tiles_w;
        const uint32_t b = fdiv(q, g.d_th);
        const int thi = (int)q - (int)b * g.tiles_h;
        const int ho0 = thi * CS_TH, wo0 = twi * CS_TW;
        const int h_start = ho0 * g.stride - g.pad, w_start = wo0 * g.stride - g.pad;
        // ---- input patch -> shared memory: row r holds the image bytes from the 16-byte aligned address at or below its
        //      first needed byte; zero outside the image ----
        for (int i = tid; i < g.rows_in * chunks; i += CS_THREADS) {
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
            *reinterpret_cast<uint4*>(s_in + (size_t)r * g.pitch_in + 16 * c) = v;
        }
        __syncthreads();
        // ---- this thread's A row: KH x WPR words re-aligned from the byte offset of its pixel; code sum on the way ----
        int s_a = 0;
        {
            uint32_t p[KH * WPR + 3];
#pragma unroll
            for (int kh = 0; kh < KH; ++kh) {
                const int r = th * g.stride + kh;
                const long long first = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h_start + r) * row_len +
                                        (long long)w_start * g.Cin;
                const int o = tw * g.stride * g.Cin + (int)(first & 15LL);
                const int sh = (o & 3) * 8;
                const uint32_t* rowp = reinterpret_cast<const uint32_t*>(s_in + (size_t)r * g.pitch_in + (o & ~3));
                uint32_t prev = rowp[0];
#pragma unroll
                for (int j = 0; j < WPR; ++j) {
                    const uint32_t nxt = rowp[j + 1];
                    uint32_t v = __funnelshift_r(prev, nxt, sh);
                    if (j == WPR - 1) v &= last_mask;         // bytes beyond the filter row belong to the next pixel
                    p[kh * WPR + j] = v;
                    s_a = dp4a_uu(v, 0x01010101u, s_a);
                    prev = nxt;
                }
            }
            p[KH * WPR] = p[KH * WPR + 1] = p[KH * WPR + 2] = 0u;
#pragma unroll
            for (int c = 0; c < (KH * WPR + 3) / 4; ++c)
                *reinterpret_cast<uint4*>(s_A + stem_swz((uint32_t)tid, (uint32_t)c * 16u, 128u)) =
                    make_uint4(p[4 * c], p[4 * c + 1], p[4 * c + 2], p[4 * c + 3]);
        }
        fence_proxy_async();                                  // generic-proxy writes of A -> visible to the tensor core
        tc_fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after_sync();
            for (int k = 0; k < g.ksteps; ++k) {
                const uint32_t off_a = (uint32_t)(k >> 2) * 128u * 128u + (uint32_t)(k & 3) * 32u;
                const uint32_t off_b = (uint32_t)(k >> 2) * (uint32_t)g.Cout * 128u + (uint32_t)(k & 3) * 32u;
                umma_i8(tmem, umma_smem_desc_sw128(a_u32 + off_a), umma_smem_desc_sw128(b_u32 + off_b), idesc, k > 0 ? 1u : 0u);
            }
            umma_commit(bar);
        }
        // in-bounds tap ranges of this pixel -> case index of the padding table
        const int ho = ho0 + th, wo = wo0 + tw;
        const bool valid = ho < g.Ho && wo < g.Wo;
        int cls = 0;
        if (valid) {
            const int h0 = ho * g.stride - g.pad, w0 = wo * g.stride - g.pad;
            const int ka = h0 < 0 ? -h0 : 0, kb = (h0 + KH > g.H) ? g.H - h0 : KH;
            const int kc = w0 < 0 ? -w0 : 0, kd = (w0 + g.kw > g.W) ? g.W - w0 : g.kw;
            int ih = 0, iw = 0;
#pragma unroll
            for (int c = 1; c < CS_MAX_CLS; ++c) {
                if (c < g.ch.n && g.ch.a[c] == ka && g.ch.b[c] == kb) ih = c;
                if (c < g.cw.n && g.cw.a[c] == kc && g.cw.b[c] == kd) iw = c;
            }
            cls = ih * g.cw.n + iw;
        }
        const float* Rrow = s_R + (size_t)cls * g.Cout;
        float* optr = out + (((size_t)b * g.Ho + ho) * g.Wo + wo) * g.Cout;
        const float fs_a = (float)s_a;
        {   // bounded wait: a lost MMA completion must not hang the GPU
            uint32_t spins = 0;
            while (!mbar_try_wait(bar, phase)) {
                if (++spins > (1u << 26)) __trap();
            }
            phase ^= 1u;
        }
        tc_fence_after_sync();
        for (int c0 = 0; c0 < g.Cout; c0 += 32) {
            uint32_t acc[32];
            tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, acc);
            tmem_ld_wait();
            if (valid) {
#pragma unroll
                for (int u = 0; u < 32; u += 4) {
                    const float4 cfa = *reinterpret_cast<const float4*>(s_cf + c0 + u);
                    const float4 cfb = *reinterpret_cast<const float4*>(s_cf + c0 + u + 2);
                    const float4 rr = *reinterpret_cast<const float4*>(Rrow + c0 + u);
                    float4 o;
                    o.x = fmaf(cfa.x, (float)(int)acc[u], fmaf(cfa.y, fs_a, rr.x));
                    o.y = fmaf(cfa.z, (float)(int)acc[u + 1], fmaf(cfa.w, fs_a, rr.y));
                    o.z = fmaf(cfb.x, (float)(int)acc[u + 2], fmaf(cfb.y, fs_a, rr.z));
                    o.w = fmaf(cfb.z, (float)(int)acc[u + 3], fmaf(cfb.w, fs_a, rr.w));
                    if (c0 + u < g.Cout) *reinterpret_cast<float4*>(optr + c0 + u) = o;
                }
            }
        }
        tc_fence_before_sync();                               // accumulator reads done before the next tile's MMA
        __syncthreads();                                      // ... and the patch / A rows are free again
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<CS_TMEM_COLS>(tmem);
}

// distinct in-bounds ranges [a, b) of a k-wide window over n_out positions of an n_in-long axis; class 0 = interior
static bool stem_classes(StemCls& c, int n_in, int n_out, int k, int stride, int pad) {
    c.n = 1;
    c.a[0] = 0; c.b[0] = (signed char)k;
    for (int o = 0; o < n_out; ++o) {
        const int p0 = o * stride - pad;
        const int a = p0 < 0 ? -p0 : 0, b = (p0 + k > n_in) ? n_in - p0 : k;
        bool found = false;
        for (int i = 0; i < c.n; ++i) found = found || (c.a[i] == a && c.b[i] == b);
        if (!found) {
            if (c.n == CS_MAX_CLS) return false;
            c.a[c.n] = (signed char)a; c.b[c.n] = (signed char)b;
            ++c.n;
        }
    }
    for (int i = c.n; i < CS_MAX_CLS; ++i) { c.a[i] = 0; c.b[i] = 0; }
    return true;
}

static bool stem_geom(StemGeom& g, int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad) {
    const int64_t rb = (int64_t)kw * Cin;
    const bool shape = (kh == 7 && rb > 20 && rb <= 24) || (kh == 3 && rb > 8 && rb <= 12);
    if (!(shape && Cin >= 1 && Cin <= 8 && Cout >= 16 && Cout <= CS_MAX_COUT && Cout % 16 == 0 && stride >= 1 && stride <= 2 && pad >= 0 &&
          B > 0 && H > 0 && W > 0 && H < (1 << 20) && W < (1 << 20)))
        return false;
    const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
    if (H + 2 * pad < kh || W + 2 * pad < kw || Ho <= 0 || Wo <= 0 || B * Ho * Wo >= 2147483647LL / 128) return false;
    g.B = (int)B; g.H = (int)H; g.W = (int)W; g.Cin = (int)Cin; g.Ho = (int)Ho; g.Wo = (int)Wo; g.Cout = (int)Cout;
    g.kh = kh; g.kw = kw; g.stride = stride; g.pad = pad;
    const int wpr = kh == 7 ? 6 : 3;
    g.ksteps = (kh * wpr * 4 + 31) / 32;
    g.katoms = (g.ksteps + 3) / 4;
    g.tiles_h = (g.Ho + CS_TH - 1) / CS_TH;
    g.tiles_w = (g.Wo + CS_TW - 1) / CS_TW;
    g.rows_in = (CS_TH - 1) * stride + kh;
    // 15 bytes of alignment slack + the bytes of 32 pixels' windows + one extra word read by the re-alignment
    g.pitch_in = (15 + ((CS_TW - 1) * stride) * g.Cin + (wpr + 1) * 4 + 15) & ~15;
    g.d_tw = make_fastdiv((uint32_t)g.tiles_w);
    g.d_th = make_fastdiv((uint32_t)g.tiles_h);
    g.d_chunks = make_fastdiv((uint32_t)(g.pitch_in >> 4));
    return stem_classes(g.ch, g.H, g.Ho, kh, stride, pad) && stem_classes(g.cw, g.W, g.Wo, kw, stride, pad);
}

extern "C" int qb_conv_stem_supported(int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad) {
    StemGeom g;
    return stem_geom(g, 1, H, W, Cin, Cout, kh, kw, stride, pad) ? 1 : 0;
}

extern "C" size_t qb_conv_stem_ws_bytes(int64_t Cout) {
    return (Cout > 0 && Cout <= CS_MAX_COUT) ? stem_ws_bytes((int)Cout, 2, CS_MAX_CLS * CS_MAX_CLS) : 0;
}

template <int KH, int WPR>
static int stem_launch(const uint8_t* x, float* out, const unsigned char* ws, const StemGeom& g, cudaStream_t st) {
    const size_t wsb = stem_ws_bytes(g.Cout, g.katoms, g.ch.n * g.cw.n);
    const size_t smem = 1024 + (size_t)g.katoms * 128 * 128 + ((wsb + 15) & ~(size_t)15) +
                        (((size_t)g.rows_in * g.pitch_in + 15) & ~(size_t)15) + 16;
    if (smem > 200 * 1024) {
        qb_set_error("qb_conv_stem_fwd: %zu bytes of shared memory needed", smem);
        return QB_ERR_UNSUPPORTED;
    }
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int per_sm = (int)((size_t)220 * 1024 / smem);
    if (per_sm > 4) per_sm = 4;                               // 4 x 128 TMEM columns
    if (per_sm < 1) per_sm = 1;
    const int64_t total = (int64_t)g.B * g.tiles_h * g.tiles_w;
    const unsigned grid = (unsigned)(total < (int64_t)per_sm * sms ? total : (int64_t)per_sm * sms);
    QB_CUDA(cudaFuncSetAttribute(k_conv_stem<KH, WPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_stem<KH, WPR><<<grid, CS_THREADS, smem, st>>>(x, out, ws, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_stem_fwd(const uint8_t* x_codes, const int8_t* w_codes_khwc, float* out, int64_t B, int64_t H, int64_t W,
                                int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                                const float* w_row_min, const float* w_row_max, int w_bits, const int32_t* tapsum,
                                const float* bias, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(x_codes && w_codes_khwc && out && a_range && w_row_min && w_row_max && tapsum && ws,
                 "qb_conv_stem_fwd: null pointer");
    StemGeom g;
    if (!stem_geom(g, B, H, W, Cin, Cout, kh, kw, stride, pad)) {
        qb_set_error("qb_conv_stem_fwd: unsupported geometry Cin=%lld Cout=%lld k=%dx%d stride=%d pad=%d H=%lld W=%lld (7x7 with "
                     "21..24-byte rows or 3x3 with 9..12-byte rows, Cout %% 16 == 0 and <= 128, stride <= 2, <= 8 border cases "
                     "per axis)", (long long)Cin, (long long)Cout, kh, kw, stride, pad, (long long)H, (long long)W);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8 && w_bits >= 1 && w_bits <= 8, "qb_conv_stem_fwd: codes are 8-bit");
    QB_CHECK_ARG(ws_bytes >= qb_conv_stem_ws_bytes(Cout) && (reinterpret_cast<uintptr_t>(ws) & 15) == 0,
                 "qb_conv_stem_fwd: workspace too small or not 16-byte aligned");
    QB_CHECK_ARG((reinterpret_cast<uintptr_t>(out) & 15) == 0, "qb_conv_stem_fwd: out must be 16-byte aligned");
    StemEpi ep{a_range, a_bits, w_row_min, w_row_max, w_bits, tapsum, bias};
    cudaStream_t st = qb_stream(stream);
    k_stem_prep<<<1, 256, 0, st>>>(static_cast<unsigned char*>(ws), w_codes_khwc, g, ep, kh == 7 ? 6 : 3);
    QB_LAUNCH_CHECK();
    if (kh == 7) return stem_launch<7, 6>(x_codes, out, static_cast<const unsigned char*>(ws), g, st);
    return stem_launch<3, 3>(x_codes, out, static_cast<const unsigned char*>(ws), g, st);
}
