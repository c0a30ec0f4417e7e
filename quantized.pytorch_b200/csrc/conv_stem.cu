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
//     <= 6 tcgen05.mma (M=128, N=Cout, K=32) against the resident, pre-swizzled weight operand; the next tile's
//     input patch is fetched into registers meanwhile
//   * the same thread reads ITS accumulator row back (tcgen05.ld 32x32b), applies the dequantising epilogue and stages
//     it for a TMA store (32-channel slabs, alternating between the A rows and a second buffer).  The zero-padding terms depend only on which filter rows / columns are in bounds: the host lists
//     the <= 8 x 8 distinct cases, k_stem_prep tabulates R[case][n] = la*sw*SUM_V qc + la*lw'*|V|*Cin + bias, and the
//     epilogue is two FMAs per output for interior and border pixels alike
//   * otherwise no intra-CTA pipelining: 4 CTAs per SM (45 KB each) overlap each other's phases
// History (256x3x224x224 -> 64 channels): cuDNN fp32 1.32 ms (+ layout conversions in the step); direct DP4A convolution,
// weights broadcast from shared memory 1.46 ms (LSU bound), 2 pixels per thread 0.93 ms (IDP.4A runs on the half-rate
// fmaheavy pipe: 8.6 G lane-ops is a 0.54 ms floor); this kernel: see profiles/.
#include "tc_common.cuh"

constexpr int CS_THREADS = 128;
constexpr int CS_TW = 32, CS_TH = 4;
constexpr int CS_MAX_COUT = 128, CS_MAX_CLS = 8;
constexpr uint32_t CS_TMEM_COLS = 128;
constexpr int CS_SLAB_BYTES = 128 * 128;                 // one staged output slab: 128 pixels x 32 fp32 channels

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

// Operand rows are K-major: bytes [0,128) of every row in a 128-byte-swizzled atom (rows x 128 B), bytes [128,192) - only
// the 7x7 filter needs them - in a 64-byte-swizzled atom behind it (rows x 64 B), so a 192-byte row costs 192 bytes.
__host__ __device__ inline size_t stem_operand_bytes(int rows, int katoms) { return (size_t)rows * (katoms == 2 ? 192 : 128); }
// workspace = what every CTA stages: [ B operand, swizzled | cf: float2[Cout] | R: float[nh*nw][Cout] ]
__host__ __device__ inline size_t stem_ws_bytes(int Cout, int katoms, int ncls) {
    return stem_operand_bytes(Cout, katoms) + (size_t)Cout * 8 + (size_t)ncls * Cout * 4;
}
__host__ __device__ inline uint32_t stem_swz(uint32_t row, uint32_t kbyte, uint32_t rows) {   // byte offset of (row, k)
    if (kbyte < 128u) return row * 128u + ((((kbyte >> 4) ^ (row & 7u)) << 4) | (kbyte & 15u));
    const uint32_t kk = kbyte - 128u;
    return rows * 128u + row * 64u + ((((kk >> 4) ^ ((row >> 1) & 3u)) << 4) | (kk & 15u));
}
// K-major shared-memory descriptor with the 128- or 64-byte swizzle: 8-row groups are 8 rows apart
__device__ __forceinline__ uint64_t stem_desc(uint32_t smem_addr, uint32_t row_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)((8u * row_bytes) >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)(row_bytes == 128u ? 2 : 4) << 61;
    return d;
}

__device__ __forceinline__ int dp4a_uu(uint32_t a, uint32_t b, int c) {
    int d;
    asm("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

__global__ void __launch_bounds__(256) k_stem_prep(unsigned char* ws, const int8_t* __restrict__ w, StemGeom g, StemEpi ep, int WPR) {
    const int tid = threadIdx.x, KH = g.kh;
    const int row_bytes = g.kw * g.Cin, taps = KH * g.kw;
    const size_t b_bytes = stem_operand_bytes(g.Cout, g.katoms);
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
__global__ void __launch_bounds__(CS_THREADS, 4)
k_conv_stem(const __grid_constant__ CUtensorMap tmap_o, const uint8_t* __restrict__ x, const unsigned char* __restrict__ ws,
            StemGeom g) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the address space
    const int ncls = g.ch.n * g.cw.n;
    const size_t a_bytes = stem_operand_bytes(128, g.katoms);               // multiples of 1024 (Cout % 16 == 0)
    const size_t b_bytes = stem_operand_bytes(g.Cout, g.katoms);
    const size_t wsb = stem_ws_bytes(g.Cout, g.katoms, ncls);
    unsigned char* s_A = smem;
    unsigned char* s_O = smem + a_bytes;                                     // second output staging slab (the first is s_A)
    unsigned char* s_B = s_O + CS_SLAB_BYTES;                                // staged workspace starts here
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
    if (warp == 0) {
        if (g.Cout <= 64) tmem_alloc<64>(tmem_slot);
        else tmem_alloc<CS_TMEM_COLS>(tmem_slot);
    }
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
    // chunk i (16 bytes) of the input patch of tile t: row r holds the image bytes from the 16-byte aligned address at or
    // below its first needed byte; zero outside the image
    auto load_chunk = [&](int t, int i) -> uint4 {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (t >= total || i >= g.rows_in * chunks) return v;
        uint32_t q = fdiv((uint32_t)t, g.d_tw);
        const int twi = t - (int)q * g.tiles_w;
        const uint32_t b = fdiv(q, g.d_th);
        const int thi = (int)q - (int)b * g.tiles_h;
        const int r = (int)fdiv((uint32_t)i, g.d_chunks), c = i - r * chunks;
        const int h = thi * CS_TH * g.stride - g.pad + r;
        const long long row_addr = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h) * row_len;
        const long long first = row_addr + (long long)(twi * CS_TW * g.stride - g.pad) * g.Cin;
        const long long a0 = (first & ~15LL) + 16LL * c;                           // global address of this chunk
        if (h >= 0 && h < g.H) {
            if (a0 >= row_addr && a0 + 16 <= row_addr + row_len) {
                v = __ldg(reinterpret_cast<const uint4*>(a0));
            } else if (a0 + 16 > row_addr && a0 < row_addr + row_len) {            // straddles an image edge
                uint32_t wd[4] = {0u, 0u, 0u, 0u};
                for (int bt = 0; bt < 16; ++bt) {
                    const long long a = a0 + bt;
                    if (a >= row_addr && a < row_addr + row_len)
                        wd[bt >> 2] |= (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(a)) << (8 * (bt & 3));
                }
                v = make_uint4(wd[0], wd[1], wd[2], wd[3]);
            }
        }
        return v;
    };
    // the patch of the NEXT tile is fetched into registers while the current tile's MMA and epilogue run
    uint4 pre0 = load_chunk((int)blockIdx.x, tid), pre1 = load_chunk((int)blockIdx.x, tid + CS_THREADS);
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
        uint32_t q = fdiv((uint32_t)t, g.d_tw);
        const int twi = t - (int)q * g.tiles_w;
        const uint32_t b = fdiv(q, g.d_th);
        const int thi = (int)q - (int)b * g.tiles_h;
        const int ho0 = thi * CS_TH, wo0 = twi * CS_TW;
        const int h_start = ho0 * g.stride - g.pad, w_start = wo0 * g.stride - g.pad;
        if (tid < g.rows_in * chunks) *reinterpret_cast<uint4*>(s_in + (size_t)tid * 16) = pre0;          // chunk i sits at byte 16*i
        if (tid + CS_THREADS < g.rows_in * chunks) *reinterpret_cast<uint4*>(s_in + (size_t)(tid + CS_THREADS) * 16) = pre1;
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
        pre0 = load_chunk(t + (int)gridDim.x, tid);
        pre1 = load_chunk(t + (int)gridDim.x, tid + CS_THREADS);
        fence_proxy_async();                                  // generic-proxy writes of A -> visible to the tensor core
        tc_fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after_sync();
            for (int k = 0; k < g.ksteps; ++k) {
                const uint32_t off_a = (uint32_t)(k >> 2) * 128u * 128u + (uint32_t)(k & 3) * 32u;
                const uint32_t off_b = (uint32_t)(k >> 2) * (uint32_t)g.Cout * 128u + (uint32_t)(k & 3) * 32u;
                const uint32_t rb = k < 4 ? 128u : 64u;       // K steps 4, 5 live in the 64-byte-swizzled atom
                umma_i8(tmem, stem_desc(a_u32 + off_a, rb), stem_desc(b_u32 + off_b, rb), idesc, k > 0 ? 1u : 0u);
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
        const float fs_a = (float)s_a;
        {   // bounded wait: a lost MMA completion must not hang the GPU
            uint32_t spins = 0;
            while (!mbar_try_wait(bar, phase)) {
                if (++spins > (1u << 26)) __trap();
            }
            phase ^= 1u;
        }
        tc_fence_after_sync();
        // ---- epilogue: slabs of 32 channels, staged in the 128-byte-swizzled layout of the output tensor map and written
        //      by TMA (a thread's own stores would touch 32 different lines per instruction).  Slabs alternate between
        //      the A rows (free once the MMA has completed) and s_O; ragged tiles are clipped by the TMA unit ----
        const uint32_t sw = (uint32_t)tid & 7u;
        const int nslab = (g.Cout + 31) >> 5;
        for (int sl = 0; sl < nslab; ++sl) {
            const int c0 = sl * 32;
            uint32_t acc[32];
            tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, acc);
            tmem_ld_wait();
#pragma unroll
            for (int u = 0; u < 32; u += 4) {
                const float4 cfa = *reinterpret_cast<const float4*>(s_cf + c0 + u);
                const float4 cfb = *reinterpret_cast<const float4*>(s_cf + c0 + u + 2);
                const float4 rr = *reinterpret_cast<const float4*>(Rrow + c0 + u);
                acc[u] = __float_as_uint(fmaf(cfa.x, (float)(int)acc[u], fmaf(cfa.y, fs_a, rr.x)));
                acc[u + 1] = __float_as_uint(fmaf(cfa.z, (float)(int)acc[u + 1], fmaf(cfa.w, fs_a, rr.y)));
                acc[u + 2] = __float_as_uint(fmaf(cfb.x, (float)(int)acc[u + 2], fmaf(cfb.y, fs_a, rr.z)));
                acc[u + 3] = __float_as_uint(fmaf(cfb.z, (float)(int)acc[u + 3], fmaf(cfb.w, fs_a, rr.w)));
            }
            if (sl >= 1) {                                    // the buffer's previous store (two slabs ago) has been read
                if (tid == 0) tma_store_wait_read<1>();
                __syncthreads();
            }
            unsigned char* buf = (sl & 1) ? s_O : s_A;
            const uint32_t dst = smem_u32(buf) + (uint32_t)tid * 128u;
#pragma unroll
            for (int q4 = 0; q4 < 8; ++q4)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)q4 ^ sw) << 4)), "r"(acc[4 * q4]),
                             "r"(acc[4 * q4 + 1]), "r"(acc[4 * q4 + 2]), "r"(acc[4 * q4 + 3]) : "memory");
            fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                tma_store_4d(&tmap_o, buf, c0, wo0, ho0, (int)b);
                tma_store_commit();
            }
        }
        if (tid == 0) {                                       // the A rows are rewritten by the next tile
            if (nslab & 1) tma_store_wait_read<0>();
            else tma_store_wait_read<1>();
        }
        tc_fence_before_sync();                               // accumulator reads done before the next tile's MMA
        __syncthreads();                                      // ... and the patch / A rows are free again
    }
    if (tid == 0) tma_store_wait<0>();
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) {
        if (g.Cout <= 64) tmem_dealloc<64>(tmem);
        else tmem_dealloc<CS_TMEM_COLS>(tmem);
    }
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
    if (g.rows_in * (g.pitch_in >> 4) > 2 * CS_THREADS) return false;            // two prefetch registers per thread
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
    const size_t smem = 1024 + stem_operand_bytes(128, g.katoms) + CS_SLAB_BYTES + ((wsb + 15) & ~(size_t)15) +
                        (((size_t)g.rows_in * g.pitch_in + 15) & ~(size_t)15) + 16;
    if (smem > 200 * 1024) {
        qb_set_error("qb_conv_stem_fwd: %zu bytes of shared memory needed", smem);
        return QB_ERR_UNSUPPORTED;
    }
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    int per_sm = (int)((size_t)220 * 1024 / smem);
    if (per_sm > 4) per_sm = 4;                               // registers (and 4 x 128 TMEM columns)
    if (per_sm < 1) per_sm = 1;
    const int64_t total = (int64_t)g.B * g.tiles_h * g.tiles_w;
    const unsigned grid = (unsigned)(total < (int64_t)per_sm * sms ? total : (int64_t)per_sm * sms);
    CUtensorMap to;
    {   // output fp32 NHWC: dims {Cout, Wo, Ho, B}; one staged slab = 32 channels x the tile's 4 x 32 pixels
        const uint64_t dims[4] = {(uint64_t)g.Cout, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)g.B};
        const uint64_t str[3] = {(uint64_t)g.Cout * 4, (uint64_t)g.Wo * g.Cout * 4, (uint64_t)g.Ho * g.Wo * g.Cout * 4};
        const uint32_t box[4] = {32u, (uint32_t)CS_TW, (uint32_t)CS_TH, 1u};
        int rc = qb_make_tmap(&to, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, out, dims, str, box, nullptr, 128);
        if (rc) return rc;
    }
    QB_CUDA(cudaFuncSetAttribute(k_conv_stem<KH, WPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_conv_stem<KH, WPR><<<grid, CS_THREADS, smem, st>>>(to, x, ws, g);
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

// ------------------------------------------------------------------------------------------------
// Weight gradient of the stem: dW[n][c][kh][kw] = SUM_pixels g[px][n] * x_hat[px + tap][c]   (quantize.py:143-148: under
// the bi-precision scheme the weight branch sees the UNQUANTIZED gradient, so this is not an integer contraction).
// With x_hat = sa*q + la inside the image and exactly 0 in the padding:
//     dW = sa * SUM g*q  +  la * SUM_{px: tap in bounds} g
// Both sums come from ONE tf32 tensor-core GEMM over the pixels, D[r][n] = SUM_px P[r][px] * G[n][px]:
//     rows r <  KH*kw*Cin : P = the code q of patch byte r (exact in tf32),   rows after them: P = 1 if tap (kh,kw) is in bounds
// Same scheme as the forward kernel, transposed: thread = pixel builds ITS column of both K-major operands (K = pixels, 32
// per 128-byte swizzled row) - the patch codes from the staged input rows, g from a TMA-loaded, swizzled tile of the
// gradient (rounded to tf32, round-to-nearest) - one thread issues the MMAs, the accumulators stay in TMEM over all the
// tiles of the CTA (one CTA per SM), and are written once as a partial result; k_stem_wgrad_reduce adds the partials
// in CTA order (deterministic) and applies sa / la.   cuDNN: 1.6 ms + layout conversions for this layer.
// ------------------------------------------------------------------------------------------------
constexpr int CW_THREADS = 128;

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ uint32_t to_tf32(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ void mbar_wait_bounded(uint64_t* bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > (1u << 26)) __trap();                   // a lost completion must not hang the GPU
    }
}

template <int KH, int WPR>
__global__ void __launch_bounds__(CW_THREADS, 1)
k_stem_wgrad(const __grid_constant__ CUtensorMap tmap_g, const uint8_t* __restrict__ x, float* __restrict__ partial, StemGeom g,
             int mhalves) {
    extern __shared__ unsigned char smem_raw[];
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t a_atom = (uint32_t)mhalves * 128u * 128u, b_atom = (uint32_t)g.Cout * 128u;
    const uint32_t stage_bytes = 128u * (uint32_t)g.Cout * 4u;               // [Cout/32][128 px][128 B]
    unsigned char* s_A = smem;                                               // 4 atoms (one per warp = 32 pixels)
    unsigned char* s_B = s_A + 4 * a_atom;
    unsigned char* s_G = s_B + 4 * b_atom;                                   // TMA destination, 128-byte swizzle
    unsigned char* s_in = s_G + stage_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_in + (((size_t)g.rows_in * g.pitch_in + 15) & ~(size_t)15));
    uint64_t* g_bar = bars;
    uint64_t* mma_bar = bars + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < (int)((4 * a_atom + 4 * b_atom) / 16); i += CW_THREADS) reinterpret_cast<uint4*>(s_A)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (tid == 0) {
        mbar_init(g_bar, 1);
        mbar_init(mma_bar, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<128>(tmem_slot);
    fence_proxy_async();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = *tmem_slot;
    const uint32_t idesc = umma_idesc(/*F32*/ 1, /*tf32*/ 2, /*tf32*/ 2, 128, (uint32_t)g.Cout);
    const uint32_t a_u32 = smem_u32(s_A), b_u32 = smem_u32(s_B);

    const int row_bytes = g.kw * g.Cin;
    const int rows_q = KH * row_bytes;
    const int th = warp, tw = lane;
    const int row_len = g.W * g.Cin, chunks = g.pitch_in >> 4;
    const int total = g.B * g.tiles_h * g.tiles_w;
    const uint32_t last_mask = (row_bytes & 3) ? ((1u << (8 * (row_bytes & 3))) - 1u) : 0xffffffffu;
    const int nhalf = g.Cout >> 5;

    auto tile_coords = [&](int t, int& b, int& ho0, int& wo0) {
        uint32_t q = fdiv((uint32_t)t, g.d_tw);
        const int twi = t - (int)q * g.tiles_w;
        const uint32_t bb = fdiv(q, g.d_th);
        const int thi = (int)q - (int)bb * g.tiles_h;
        b = (int)bb; ho0 = thi * CS_TH; wo0 = twi * CS_TW;
    };
    auto load_chunk = [&](int t, int i) -> uint4 {
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (t >= total || i >= g.rows_in * chunks) return v;
        int b, ho0, wo0;
        tile_coords(t, b, ho0, wo0);
        const int r = (int)fdiv((uint32_t)i, g.d_chunks), c = i - r * chunks;
        const int h = ho0 * g.stride - g.pad + r;
        const long long row_addr = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h) * row_len;
        const long long first = row_addr + (long long)(wo0 * g.stride - g.pad) * g.Cin;
        const long long a0 = (first & ~15LL) + 16LL * c;
        if (h >= 0 && h < g.H) {
            if (a0 >= row_addr && a0 + 16 <= row_addr + row_len) {
                v = __ldg(reinterpret_cast<const uint4*>(a0));
            } else if (a0 + 16 > row_addr && a0 < row_addr + row_len) {
                uint32_t wd[4] = {0u, 0u, 0u, 0u};
                for (int bt = 0; bt < 16; ++bt) {
                    const long long a = a0 + bt;
                    if (a >= row_addr && a < row_addr + row_len)
                        wd[bt >> 2] |= (uint32_t)__ldg(reinterpret_cast<const uint8_t*>(a)) << (8 * (bt & 3));
                }
                v = make_uint4(wd[0], wd[1], wd[2], wd[3]);
            }
        }
        return v;
    };
    auto issue_g = [&](int t) {                               // one thread: the gradient tile of tile t -> s_G
        int b, ho0, wo0;
        tile_coords(t, b, ho0, wo0);
        mbar_arrive_expect_tx(g_bar, stage_bytes);
        for (int hf = 0; hf < nhalf; ++hf) tma_load_4d(s_G + (size_t)hf * 128 * 128, &tmap_g, g_bar, hf * 32, wo0, ho0, b);
    };

    uint4 pre0 = load_chunk((int)blockIdx.x, tid), pre1 = load_chunk((int)blockIdx.x, tid + CW_THREADS);
    if (tid == 0 && (int)blockIdx.x < total) issue_g((int)blockIdx.x);
    uint32_t g_phase = 0, mma_phase = 0;
    bool first = true;
    // this thread's column of the K-major operands: pixel `lane` of atom `warp`
    unsigned char* my_A = s_A + (size_t)warp * a_atom + (size_t)(lane & 3) * 4;
    unsigned char* my_B = s_B + (size_t)warp * b_atom + (size_t)(lane & 3) * 4;
    const uint32_t kchunk = (uint32_t)lane >> 2;
    for (int t = blockIdx.x; t < total; t += gridDim.x) {
        int b, ho0, wo0;
        tile_coords(t, b, ho0, wo0);
        const int h_start = ho0 * g.stride - g.pad, w_start = wo0 * g.stride - g.pad;
        if (tid < g.rows_in * chunks) *reinterpret_cast<uint4*>(s_in + (size_t)tid * 16) = pre0;
        if (tid + CW_THREADS < g.rows_in * chunks) *reinterpret_cast<uint4*>(s_in + (size_t)(tid + CW_THREADS) * 16) = pre1;
        __syncthreads();
        uint32_t p[KH][WPR];
#pragma unroll
        for (int kh = 0; kh < KH; ++kh) {
            const int r = th * g.stride + kh;
            const long long firstb = (long long)reinterpret_cast<uintptr_t>(x) + ((long long)b * g.H + h_start + r) * row_len +
                                     (long long)w_start * g.Cin;
            const int o = tw * g.stride * g.Cin + (int)(firstb & 15LL);
            const int sh = (o & 3) * 8;
            const uint32_t* rowp = reinterpret_cast<const uint32_t*>(s_in + (size_t)r * g.pitch_in + (o & ~3));
            uint32_t prev = rowp[0];
#pragma unroll
            for (int j = 0; j < WPR; ++j) {
                const uint32_t nxt = rowp[j + 1];
                uint32_t v = __funnelshift_r(prev, nxt, sh);
                if (j == WPR - 1) v &= last_mask;
                p[kh][j] = v;
                prev = nxt;
            }
        }
        pre0 = load_chunk(t + (int)gridDim.x, tid);
        pre1 = load_chunk(t + (int)gridDim.x, tid + CW_THREADS);
        // the previous tile's MMAs have read the operands
        if (!first) {
            mbar_wait_bounded(mma_bar, mma_phase);
            mma_phase ^= 1u;
        }
        // ---- P rows: codes, then tap-in-bounds flags ----
        const int ho = ho0 + th, wo = wo0 + tw;
        const int h0 = ho * g.stride - g.pad, w0 = wo * g.stride - g.pad;
#pragma unroll
        for (int kh = 0; kh < KH; ++kh) {
#pragma unroll
            for (int bb = 0; bb < WPR * 4; ++bb) {
                if (bb < row_bytes) {
                    const uint32_t r = (uint32_t)(kh * row_bytes + bb);
                    const float v = (float)((p[kh][bb >> 2] >> (8 * (bb & 3))) & 0xffu);
                    *reinterpret_cast<float*>(my_A + r * 128u + ((kchunk ^ (r & 7u)) << 4)) = v;
                }
            }
        }
        for (int kh = 0; kh < KH; ++kh) {
            const bool vh = (h0 + kh >= 0) && (h0 + kh < g.H);
            for (int kw = 0; kw < g.kw; ++kw) {
                const bool vw = (w0 + kw >= 0) && (w0 + kw < g.W);
                const uint32_t r = (uint32_t)(rows_q + kh * g.kw + kw);
                *reinterpret_cast<float*>(my_A + r * 128u + ((kchunk ^ (r & 7u)) << 4)) = (vh && vw) ? 1.0f : 0.0f;
            }
        }
        // ---- G rows from the TMA tile (pixel tid of the tile = row tid of each 32-channel half) ----
        mbar_wait_bounded(g_bar, g_phase);
        g_phase ^= 1u;
        for (int hf = 0; hf < nhalf; ++hf) {
            const unsigned char* src = s_G + (size_t)hf * 128 * 128 + (size_t)tid * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(src + ((((uint32_t)c) ^ ((uint32_t)tid & 7u)) << 4));
                const uint32_t n = (uint32_t)(hf * 32 + c * 4);
                *reinterpret_cast<uint32_t*>(my_B + (n + 0) * 128u + ((kchunk ^ ((n + 0) & 7u)) << 4)) = to_tf32(v.x);
                *reinterpret_cast<uint32_t*>(my_B + (n + 1) * 128u + ((kchunk ^ ((n + 1) & 7u)) << 4)) = to_tf32(v.y);
                *reinterpret_cast<uint32_t*>(my_B + (n + 2) * 128u + ((kchunk ^ ((n + 2) & 7u)) << 4)) = to_tf32(v.z);
                *reinterpret_cast<uint32_t*>(my_B + (n + 3) * 128u + ((kchunk ^ ((n + 3) & 7u)) << 4)) = to_tf32(v.w);
            }
        }
        fence_proxy_async();
        tc_fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            if (t + (int)gridDim.x < total) issue_g(t + (int)gridDim.x);      // s_G has been consumed
            tc_fence_after_sync();
            for (int w = 0; w < 4; ++w)
                for (int ks = 0; ks < 4; ++ks)
                    for (int h = 0; h < mhalves; ++h)
                        umma_tf32(tmem + (uint32_t)h * 64u,
                                  umma_smem_desc_sw128(a_u32 + (uint32_t)w * a_atom + (uint32_t)h * 128u * 128u + (uint32_t)ks * 32u),
                                  umma_smem_desc_sw128(b_u32 + (uint32_t)w * b_atom + (uint32_t)ks * 32u), idesc,
                                  (first && w == 0 && ks == 0) ? 0u : 1u);
            umma_commit(mma_bar);
        }
        first = false;
    }
    const int rows = rows_q + KH * g.kw;
    if (!first) {
        mbar_wait_bounded(mma_bar, mma_phase);
        tc_fence_after_sync();
        for (int h = 0; h < mhalves; ++h)
            for (int c0 = 0; c0 < g.Cout; c0 += 32) {
                uint32_t acc[32];
                tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(h * 64 + c0), acc);
                tmem_ld_wait();
                const int r = h * 128 + tid;
                if (r < rows) {
                    float* dst = partial + ((size_t)blockIdx.x * rows + r) * g.Cout + c0;
#pragma unroll
                    for (int u = 0; u < 32; u += 4)
                        *reinterpret_cast<uint4*>(dst + u) = make_uint4(acc[u], acc[u + 1], acc[u + 2], acc[u + 3]);
                }
            }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc<128>(tmem);
}

// dW[n][c][kh][kw] = sa * SUM_cta Dq[cta][(kh, kw, c)][n] + la * SUM_cta Dm[cta][(kh, kw)][n]
__global__ void __launch_bounds__(256) k_stem_wgrad_reduce(const float* __restrict__ partial, int nparts, float* __restrict__ dw,
                                                           const float* __restrict__ a_range, int a_bits, int Cout, int Cin, int kh,
                                                           int kw) {
    const int rows_q = kh * kw * Cin, rows = rows_q + kh * kw;
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= rows_q * Cout) return;
    const int n = idx % Cout, r = idx / Cout;                 // r = (kh, kw, c)
    const int c = r % Cin, tap = r / Cin;
    const QParams pa = make_qparams_d((double)__ldg(a_range), (double)__ldg(a_range + 1), a_bits);
    double sq = 0.0, sm = 0.0;
    for (int p = 0; p < nparts; ++p) {
        sq += (double)partial[((size_t)p * rows + r) * Cout + n];
        sm += (double)partial[((size_t)p * rows + rows_q + tap) * Cout + n];
    }
    dw[((size_t)n * Cin + c) * (kh * kw) + tap] = (float)((double)pa.scale * sq + (double)pa.lo * sm);
}

static int stem_wgrad_parts(const StemGeom& g) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    const int64_t total = (int64_t)g.B * g.tiles_h * g.tiles_w;
    return (int)(total < sms ? total : sms);
}

extern "C" size_t qb_conv_stem_wgrad_ws_bytes(int64_t Cout, int64_t Cin, int kh, int kw) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    return (size_t)sms * (size_t)(kh * kw * (Cin + 1)) * (size_t)Cout * 4;
}

template <int KH, int WPR>
static int stem_wgrad_launch(const CUtensorMap& tg, const uint8_t* x, float* partial, const StemGeom& g, int mhalves, int parts,
                             cudaStream_t st) {
    const size_t smem = 1024 + 4 * ((size_t)mhalves * 128 * 128 + (size_t)g.Cout * 128) + (size_t)128 * g.Cout * 4 +
                        (((size_t)g.rows_in * g.pitch_in + 15) & ~(size_t)15) + 32;
    if (smem > 226 * 1024) {
        qb_set_error("qb_conv_stem_wgrad: %zu bytes of shared memory needed", smem);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CUDA(cudaFuncSetAttribute(k_stem_wgrad<KH, WPR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_stem_wgrad<KH, WPR><<<(unsigned)parts, CW_THREADS, smem, st>>>(tg, x, partial, g, mhalves);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* grad_out fp32 [B,Ho,Wo,Cout] (NHWC), x_codes uint8 [B,H,W,Cin] with their quantizer's device range -> dw fp32
 * [Cout][Cin][kh][kw] (contiguous).  Cout 32 or 64.  ws: qb_conv_stem_wgrad_ws_bytes() bytes.                           */
extern "C" int qb_conv_stem_wgrad(const float* grad_out, const uint8_t* x_codes, float* dw, int64_t B, int64_t H, int64_t W,
                                  int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                                  void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(grad_out && x_codes && dw && a_range && ws, "qb_conv_stem_wgrad: null pointer");
    StemGeom g;
    if (!stem_geom(g, B, H, W, Cin, Cout, kh, kw, stride, pad) || !(Cout == 32 || Cout == 64)) {
        qb_set_error("qb_conv_stem_wgrad: unsupported geometry (see qb_conv_stem_fwd; C_out must be 32 or 64)");
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8, "qb_conv_stem_wgrad: codes are 8-bit");
    QB_CHECK_ARG(ws_bytes >= qb_conv_stem_wgrad_ws_bytes(Cout, Cin, kh, kw) && (reinterpret_cast<uintptr_t>(ws) & 15) == 0,
                 "qb_conv_stem_wgrad: workspace too small or not 16-byte aligned");
    QB_CHECK_ARG((reinterpret_cast<uintptr_t>(grad_out) & 15) == 0, "qb_conv_stem_wgrad: grad_out must be 16-byte aligned");
    const int rows = kh * kw * ((int)Cin + 1);
    const int mhalves = (rows + 127) / 128;
    if (mhalves > 2) {
        qb_set_error("qb_conv_stem_wgrad: %d operand rows", rows);
        return QB_ERR_UNSUPPORTED;
    }
    CUtensorMap tg;
    {   // gradient fp32 NHWC: dims {Cout, Wo, Ho, B}; box = 32 channels x the tile's 4 x 32 pixels, zero fill outside
        const uint64_t dims[4] = {(uint64_t)g.Cout, (uint64_t)g.Wo, (uint64_t)g.Ho, (uint64_t)g.B};
        const uint64_t str[3] = {(uint64_t)g.Cout * 4, (uint64_t)g.Wo * g.Cout * 4, (uint64_t)g.Ho * g.Wo * g.Cout * 4};
        const uint32_t box[4] = {32u, (uint32_t)CS_TW, (uint32_t)CS_TH, 1u};
        int rc = qb_make_tmap(&tg, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, grad_out, dims, str, box, nullptr, 128);
        if (rc) return rc;
    }
    const int parts = stem_wgrad_parts(g);
    cudaStream_t st = qb_stream(stream);
    int rc = kh == 7 ? stem_wgrad_launch<7, 6>(tg, x_codes, static_cast<float*>(ws), g, mhalves, parts, st)
                     : stem_wgrad_launch<3, 3>(tg, x_codes, static_cast<float*>(ws), g, mhalves, parts, st);
    if (rc) return rc;
    const int outs = kh * kw * (int)Cin * (int)Cout;
    k_stem_wgrad_reduce<<<(outs + 255) / 256, 256, 0, st>>>(static_cast<const float*>(ws), parts, dw, a_range, a_bits, (int)Cout,
                                                             (int)Cin, kh, kw);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
