// conv_generic.cu - direct convolution / linear kernels for EVERY geometry the tensor-core kernels do not take (channel counts
// that are not multiples of 16, grouped-but-not-depthwise, dilation, rectangular stride / padding, more than 9 taps):
// F.conv2d(x_hat, w_hat, b_hat) and its two gradients (models/modules/quantize.py:142-156, 343) on the fake-quantized fp32
// tensors, on CUDA cores with fp64 accumulation.  They exist so that NO quantized layer is ever routed to cuDNN / cuBLAS
// (north_star: no multi-backend dispatch); the geometries of the reference's models never reach them (tests do).
// One thread per output element (forward, dgrad) / one CTA per weight element with a fixed-order tree reduction (wgrad, bias):
// deterministic.  Tensors are addressed through (n, c, h, w) element strides, so NCHW and channels-last both work.
#include "common.cuh"

struct GcGeom {
    int B, Cin, H, W, Cout, Ho, Wo;
    int kh, kw, sh, sw, ph, pw, dh, dw, groups;
    long long xs[4], os[4];         // element strides (n, c, h, w) of x / dx and of out / grad_out
};

__global__ void __launch_bounds__(256)
k_gc_fwd(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias, float* __restrict__ out, GcGeom g) {
    const long long total = (long long)g.B * g.Cout * g.Ho * g.Wo;
    const int cpg = g.Cin / g.groups, opg = g.Cout / g.groups;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int wo = (int)(i % g.Wo), ho = (int)((i / g.Wo) % g.Ho), n = (int)((i / ((long long)g.Wo * g.Ho)) % g.Cout);
        const int b = (int)(i / ((long long)g.Wo * g.Ho * g.Cout));
        const int grp = n / opg;
        double acc = bias ? (double)bias[n] : 0.0;
        for (int c = 0; c < cpg; ++c)
            for (int r = 0; r < g.kh; ++r) {
                const int hi = ho * g.sh - g.ph + r * g.dh;
                if (hi < 0 || hi >= g.H) continue;
                for (int s = 0; s < g.kw; ++s) {
                    const int wi = wo * g.sw - g.pw + s * g.dw;
                    if (wi < 0 || wi >= g.W) continue;
                    acc += (double)x[b * g.xs[0] + (grp * cpg + c) * g.xs[1] + hi * g.xs[2] + wi * g.xs[3]] *
                           (double)w[(((long long)n * cpg + c) * g.kh + r) * g.kw + s];
                }
            }
        out[b * g.os[0] + n * g.os[1] + ho * g.os[2] + wo * g.os[3]] = (float)acc;
    }
}

__global__ void __launch_bounds__(256)
k_gc_dgrad(const float* __restrict__ go, const float* __restrict__ w, float* __restrict__ dx, GcGeom g) {
    const long long total = (long long)g.B * g.Cin * g.H * g.W;
    const int cpg = g.Cin / g.groups, opg = g.Cout / g.groups;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
        const int wi = (int)(i % g.W), hi = (int)((i / g.W) % g.H), c = (int)((i / ((long long)g.W * g.H)) % g.Cin);
        const int b = (int)(i / ((long long)g.W * g.H * g.Cin));
        const int grp = c / cpg, cl = c - grp * cpg;
        double acc = 0.0;
        for (int r = 0; r < g.kh; ++r) {
            const int hn = hi + g.ph - r * g.dh;
            if (hn < 0 || hn % g.sh != 0) continue;
            const int ho = hn / g.sh;
            if (ho >= g.Ho) continue;
            for (int s = 0; s < g.kw; ++s) {
                const int wn = wi + g.pw - s * g.dw;
                if (wn < 0 || wn % g.sw != 0) continue;
                const int wo = wn / g.sw;
                if (wo >= g.Wo) continue;
                for (int nl = 0; nl < opg; ++nl) {
                    const int n = grp * opg + nl;
                    acc += (double)go[b * g.os[0] + n * g.os[1] + ho * g.os[2] + wo * g.os[3]] *
                           (double)w[(((long long)n * cpg + cl) * g.kh + r) * g.kw + s];
                }
            }
        }
        dx[b * g.xs[0] + c * g.xs[1] + hi * g.xs[2] + wi * g.xs[3]] = (float)acc;
    }
}

__device__ __forceinline__ double gc_block_sum(double v, double* sm) {
    v = warp_sum(v);
    const int wp = threadIdx.x >> 5, ln = threadIdx.x & 31;
    __syncthreads();
    if (ln == 0) sm[wp] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int k = 0; k < 8; ++k) t += sm[k];
    return t;
}

// blockIdx.x < Cout*cpg*kh*kw: one weight element; the remaining Cout blocks (if db): one bias element
__global__ void __launch_bounds__(256)
k_gc_wgrad(const float* __restrict__ go, const float* __restrict__ x, float* __restrict__ dwt, float* __restrict__ db, GcGeom g) {
    __shared__ double sm[8];
    const int cpg = g.Cin / g.groups, opg = g.Cout / g.groups;
    const long long nw = (long long)g.Cout * cpg * g.kh * g.kw;
    const long long px = (long long)g.B * g.Ho * g.Wo;
    double acc = 0.0;
    if (blockIdx.x < nw) {
        const int s = (int)(blockIdx.x % g.kw), r = (int)((blockIdx.x / g.kw) % g.kh), cl = (int)((blockIdx.x / (g.kw * g.kh)) % cpg);
        const int n = (int)(blockIdx.x / ((long long)g.kw * g.kh * cpg));
        const int c = (n / opg) * cpg + cl;
        for (long long p = threadIdx.x; p < px; p += 256) {
            const int wo = (int)(p % g.Wo), ho = (int)((p / g.Wo) % g.Ho), b = (int)(p / ((long long)g.Wo * g.Ho));
            const int hi = ho * g.sh - g.ph + r * g.dh, wi = wo * g.sw - g.pw + s * g.dw;
            if (hi < 0 || hi >= g.H || wi < 0 || wi >= g.W) continue;
            acc += (double)go[b * g.os[0] + n * g.os[1] + ho * g.os[2] + wo * g.os[3]] *
                   (double)x[b * g.xs[0] + c * g.xs[1] + hi * g.xs[2] + wi * g.xs[3]];
        }
        acc = gc_block_sum(acc, sm);
        if (threadIdx.x == 0) dwt[blockIdx.x] = (float)acc;
    } else {
        const int n = (int)(blockIdx.x - nw);
        for (long long p = threadIdx.x; p < px; p += 256) {
            const int wo = (int)(p % g.Wo), ho = (int)((p / g.Wo) % g.Ho), b = (int)(p / ((long long)g.Wo * g.Ho));
            acc += (double)go[b * g.os[0] + n * g.os[1] + ho * g.os[2] + wo * g.os[3]];
        }
        acc = gc_block_sum(acc, sm);
        if (threadIdx.x == 0) db[n] = (float)acc;
    }
}

static int gc_geom(GcGeom& g, const char* who, int64_t B, int64_t Cin, int64_t H, int64_t W, int64_t Cout, const int* conv,
                   const int64_t* x_strides, const int64_t* o_strides) {
    QB_CHECK_ARG(conv && x_strides && o_strides, "%s: null geometry", who);
    g.B = (int)B; g.Cin = (int)Cin; g.H = (int)H; g.W = (int)W; g.Cout = (int)Cout;
    g.kh = conv[0]; g.kw = conv[1]; g.sh = conv[2]; g.sw = conv[3]; g.ph = conv[4]; g.pw = conv[5]; g.dh = conv[6]; g.dw = conv[7];
    g.groups = conv[8];
    QB_CHECK_ARG(B > 0 && Cin > 0 && H > 0 && W > 0 && Cout > 0 && g.kh > 0 && g.kw > 0 && g.sh > 0 && g.sw > 0 && g.ph >= 0 && g.pw >= 0 &&
                 g.dh > 0 && g.dw > 0 && g.groups > 0 && Cin % g.groups == 0 && Cout % g.groups == 0, "%s: bad geometry", who);
    g.Ho = (int)((H + 2 * g.ph - g.dh * (g.kh - 1) - 1) / g.sh + 1);
    g.Wo = (int)((W + 2 * g.pw - g.dw * (g.kw - 1) - 1) / g.sw + 1);
    QB_CHECK_ARG(g.Ho > 0 && g.Wo > 0, "%s: empty output", who);
    QB_CHECK_ARG(B * Cin * H * W < (1LL << 40) && B * Cout * g.Ho * g.Wo < (1LL << 40), "%s: tensor too large", who);
    for (int i = 0; i < 4; ++i) { g.xs[i] = x_strides[i]; g.os[i] = o_strides[i]; }
    return QB_OK;
}

static unsigned gc_grid(long long total) {
    int sms = qb_sm_count();
    if (sms <= 0) sms = 148;
    long long want = (total + 255) / 256;
    const long long cap = (long long)sms * 16;
    return (unsigned)(want < 1 ? 1 : (want < cap ? want : cap));
}

/* conv = {kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, groups}; x_strides / o_strides: element strides (n, c, h, w) of
 * x (and dx) / of out (and grad_out); w [Cout][Cin/groups][kh][kw] contiguous.  Linear layers: H = W = kh = kw = 1. */
extern "C" int qb_conv_generic_fwd(const float* x, const float* w, const float* bias, float* out, int64_t B, int64_t Cin, int64_t H,
                                   int64_t W, int64_t Cout, const int* conv, const int64_t* x_strides, const int64_t* o_strides,
                                   void* stream) {
    QB_CHECK_ARG(x && w && out, "qb_conv_generic_fwd: null pointer");
    GcGeom g;
    int rc = gc_geom(g, "qb_conv_generic_fwd", B, Cin, H, W, Cout, conv, x_strides, o_strides);
    if (rc) return rc;
    k_gc_fwd<<<gc_grid((long long)B * Cout * g.Ho * g.Wo), 256, 0, qb_stream(stream)>>>(x, w, bias, out, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_generic_dgrad(const float* grad_out, const float* w, float* dx, int64_t B, int64_t Cin, int64_t H, int64_t W,
                                     int64_t Cout, const int* conv, const int64_t* x_strides, const int64_t* o_strides, void* stream) {
    QB_CHECK_ARG(grad_out && w && dx, "qb_conv_generic_dgrad: null pointer");
    GcGeom g;
    int rc = gc_geom(g, "qb_conv_generic_dgrad", B, Cin, H, W, Cout, conv, x_strides, o_strides);
    if (rc) return rc;
    k_gc_dgrad<<<gc_grid((long long)B * Cin * H * W), 256, 0, qb_stream(stream)>>>(grad_out, w, dx, g);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

extern "C" int qb_conv_generic_wgrad(const float* grad_out, const float* x, float* dw, float* db, int64_t B, int64_t Cin, int64_t H,
                                     int64_t W, int64_t Cout, const int* conv, const int64_t* x_strides, const int64_t* o_strides,
                                     void* stream) {
    QB_CHECK_ARG(grad_out && x && (dw || db), "qb_conv_generic_wgrad: null pointer");
    GcGeom g;
    int rc = gc_geom(g, "qb_conv_generic_wgrad", B, Cin, H, W, Cout, conv, x_strides, o_strides);
    if (rc) return rc;
    const long long nw = dw ? (long long)Cout * (Cin / g.groups) * g.kh * g.kw : 0;
    QB_CHECK_ARG(nw + Cout < 2147483647LL, "qb_conv_generic_wgrad: weight too large");
    if (!dw) {          // bias only: shift the block index past an empty weight range
        GcGeom gb = g;
        gb.kh = gb.kw = 0;
        k_gc_wgrad<<<(unsigned)Cout, 256, 0, qb_stream(stream)>>>(grad_out, x, dw, db, gb);
    } else {
        k_gc_wgrad<<<(unsigned)(nw + (db ? Cout : 0)), 256, 0, qb_stream(stream)>>>(grad_out, x, dw, db, g);
    }
    QB_LAUNCH_CHECK();
    return QB_OK;
}
