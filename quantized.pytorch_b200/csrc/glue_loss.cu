// glue_loss.cu - the tail of the reference's training step that sits between the last quantized layer and the optimiser:
//   nn.AvgPool2d(k) with stride k   (models/resnet_quantized.py:153,189; models/mobilenet_quantized.py:114,124)
//   nn.CrossEntropyLoss()           (main.py:362-380: criterion(output, target), loss.backward())
// Row (f) N1 of SURVEY section 8 ("glue either side of the path").  Plain fp32, channels-last for the pooling; the arithmetic
// follows ATen's CUDA kernels step by step (window sums in raster order then one IEEE division; log-softmax as
// x - max - log(sum exp(x - max)) with a lane-strided sum and a xor-butterfly), so results agree with torch to the last bit
// or two and the model-level parity tests see no difference.
#include <math.h>

#include "common.cuh"

// ---------------------------------------------------------------------------------------------------
// average pooling, kernel k, stride k, no padding (the global 7x7 pool of the ImageNet models is the k = H = W case)
// ---------------------------------------------------------------------------------------------------
struct AvgGeom {
    int64_t B, H, W, C;
    int Ho, Wo, k;
    float div;
};

__global__ void __launch_bounds__(256) k_avgpool_fwd_nhwc(const float* __restrict__ x, float* __restrict__ out, AvgGeom g, int64_t total) {
    const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int64_t c4 = g.C >> 2;
    const int64_t q = t % c4, pix = t / c4;
    const int wo = (int)(pix % g.Wo);
    const int ho = (int)((pix / g.Wo) % g.Ho);
    const int64_t b = pix / ((int64_t)g.Wo * g.Ho);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = 0; i < g.k; ++i) {
        const float* row = x + (((b * g.H + (int64_t)ho * g.k + i) * g.W + (int64_t)wo * g.k) * g.C) + 4 * q;
        for (int j = 0; j < g.k; ++j) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(row + (int64_t)j * g.C));
            acc.x = __fadd_rn(acc.x, v.x); acc.y = __fadd_rn(acc.y, v.y); acc.z = __fadd_rn(acc.z, v.z); acc.w = __fadd_rn(acc.w, v.w);
        }
    }
    float4 o;
    o.x = __fdiv_rn(acc.x, g.div); o.y = __fdiv_rn(acc.y, g.div); o.z = __fdiv_rn(acc.z, g.div); o.w = __fdiv_rn(acc.w, g.div);
    reinterpret_cast<float4*>(out)[t] = o;
}

__global__ void __launch_bounds__(256) k_avgpool_bwd_nhwc(const float* __restrict__ gout, float* __restrict__ dx, AvgGeom g, int64_t total) {
    const int64_t t = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (t >= total) return;
    const int64_t c4 = g.C >> 2;
    const int64_t q = t % c4, pix = t / c4;
    const int w = (int)(pix % g.W);
    const int h = (int)((pix / g.W) % g.H);
    const int64_t b = pix / (g.W * g.H);
    const int ho = h / g.k, wo = w / g.k;
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ho < g.Ho && wo < g.Wo) {                                            // rows / columns past the last whole window get 0
        const float4 v = __ldg(reinterpret_cast<const float4*>(gout + (((b * g.Ho + ho) * g.Wo + wo) * g.C)) + q);
        o.x = __fdiv_rn(v.x, g.div); o.y = __fdiv_rn(v.y, g.div); o.z = __fdiv_rn(v.z, g.div); o.w = __fdiv_rn(v.w, g.div);
    }
    reinterpret_cast<float4*>(dx)[t] = o;
}

static int avg_geom(AvgGeom& g, const char* who, int64_t B, int64_t H, int64_t W, int64_t C, int k, const void* a, const void* b) {
    if (!a || !b || B <= 0 || H <= 0 || W <= 0 || C <= 0 || k <= 0 || k > H || k > W) {
        qb_set_error("%s: bad arguments (B=%lld H=%lld W=%lld C=%lld k=%d)", who, (long long)B, (long long)H, (long long)W, (long long)C, k);
        return QB_ERR_ARG;
    }
    if (C % 4 != 0 || ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) != 0) {
        qb_set_error("%s: C %% 4 == 0 and 16-byte aligned tensors required", who);
        return QB_ERR_UNSUPPORTED;
    }
    g = AvgGeom{B, H, W, C, (int)(H / k), (int)(W / k), k, (float)(k * k)};
    return QB_OK;
}

/* out [B, H/k, W/k, C] = nn.AvgPool2d(k)(x) for channels-last x [B, H, W, C] (stride k, no padding, floor mode). */
extern "C" int qb_avgpool2d_nhwc_fwd(const float* x, float* out, int64_t B, int64_t H, int64_t W, int64_t C, int k, void* stream) {
    AvgGeom g;
    if (int rc = avg_geom(g, "qb_avgpool2d_nhwc_fwd", B, H, W, C, k, x, out)) return rc;
    const int64_t total = B * g.Ho * g.Wo * (C / 4);
    k_avgpool_fwd_nhwc<<<(unsigned)((total + 255) / 256), 256, 0, qb_stream(stream)>>>(x, out, g, total);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* dx [B, H, W, C] from grad_out [B, H/k, W/k, C]: every element of a window receives g / k^2. */
extern "C" int qb_avgpool2d_nhwc_bwd(const float* grad_out, float* dx, int64_t B, int64_t H, int64_t W, int64_t C, int k, void* stream) {
    AvgGeom g;
    if (int rc = avg_geom(g, "qb_avgpool2d_nhwc_bwd", B, H, W, C, k, grad_out, dx)) return rc;
    const int64_t total = B * H * W * (C / 4);
    k_avgpool_bwd_nhwc<<<(unsigned)((total + 255) / 256), 256, 0, qb_stream(stream)>>>(grad_out, dx, g, total);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------
// cross entropy, mean over the rows whose target is not ignore_index: loss = -1/N SUM_r log_softmax(x_r)[t_r]
// One warp per row.  Forward keeps (max_r, log SUM exp(x_r - max_r)) so that backward rebuilds log_softmax bit for bit
// without a [B, C] intermediate: dlogits = go - exp(lsm) * SUM(go), go = -(g / N) at the target, 0 elsewhere.
// ---------------------------------------------------------------------------------------------------
constexpr int CE_WARPS = 8;

__global__ void __launch_bounds__(CE_WARPS * 32)
k_ce_rows_fwd(const float* __restrict__ x, const int64_t* __restrict__ target, int64_t B, int64_t C, int64_t ignore_index,
              float* __restrict__ row_state, float* __restrict__ row_loss) {
    const int64_t r = (int64_t)blockIdx.x * CE_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= B) return;
    const float* row = x + r * C;
    float mx = -QB_INF;
    for (int64_t c = lane; c < C; c += 32) mx = fmaxf(mx, __ldg(row + c));
    mx = warp_max(mx);
    float sum = 0.0f;
    for (int64_t c = lane; c < C; c += 32) sum += expf(__ldg(row + c) - mx);
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float lse = logf(sum);
    if (lane == 0) {
        const int64_t t = target[r];
        row_state[2 * r] = mx;
        row_state[2 * r + 1] = lse;
        // an out-of-range target is a caller error torch asserts on; here it contributes like an ignored row and is flagged NaN
        float l = 0.0f;
        if (t != ignore_index) l = (t >= 0 && t < C) ? -(__ldg(row + t) - mx - lse) : __int_as_float(0x7fc00000);
        row_loss[r] = l;
    }
}

__global__ void __launch_bounds__(256) k_ce_mean(const float* __restrict__ row_loss, const int64_t* __restrict__ target, int64_t B,
                                                 int64_t ignore_index, float* loss, float* n_valid) {
    __shared__ double ssum[256];
    __shared__ int64_t scnt[256];
    double s = 0.0;
    int64_t n = 0;
    for (int64_t r = threadIdx.x; r < B; r += 256) {
        if (target[r] != ignore_index) { s += (double)row_loss[r]; ++n; }
    }
    ssum[threadIdx.x] = s;
    scnt[threadIdx.x] = n;
    __syncthreads();
    for (int o = 128; o; o >>= 1) {                                         // fixed tree: deterministic
        if ((int)threadIdx.x < o) { ssum[threadIdx.x] += ssum[threadIdx.x + o]; scnt[threadIdx.x] += scnt[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *n_valid = (float)scnt[0];
        *loss = (float)(ssum[0] / (double)scnt[0]);                         // no valid row: 0 / 0 = nan, as torch
    }
}

__global__ void __launch_bounds__(CE_WARPS * 32)
k_ce_rows_bwd(const float* __restrict__ x, const int64_t* __restrict__ target, const float* __restrict__ row_state,
              const float* __restrict__ n_valid, const float* __restrict__ grad_out, int64_t B, int64_t C, int64_t ignore_index,
              float* __restrict__ dx) {
    const int64_t r = (int64_t)blockIdx.x * CE_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (r >= B) return;
    const float* row = x + r * C;
    float* out = dx + r * C;
    const int64_t t = target[r];
    if (t == ignore_index) {
        for (int64_t c = lane; c < C; c += 32) out[c] = 0.0f;
        return;
    }
    const float w = __fdiv_rn(__ldg(grad_out), __ldg(n_valid));
    const float mx = row_state[2 * r], lse = row_state[2 * r + 1];
    const float gsum = -w;                                                  // SUM over the row of go
    for (int64_t c = lane; c < C; c += 32) {
        const float lsm = __ldg(row + c) - mx - lse;
        const float go = (c == t) ? -w : 0.0f;
        out[c] = go - expf(lsm) * gsum;
    }
}

extern "C" size_t qb_cross_entropy_ws_bytes(int64_t B) { return B < 0 ? 0 : (size_t)B * sizeof(float); }

/* nn.CrossEntropyLoss() (mean reduction, no class weights, no label smoothing) on logits [B, C] and int64 targets [B]
 * (main.py:362-380).  loss: device float; row_state: float[2*B] kept for backward; n_valid: device float (rows counted). */
extern "C" int qb_cross_entropy_fwd(const float* logits, const int64_t* target, int64_t B, int64_t C, int64_t ignore_index,
                                    float* loss, float* row_state, float* n_valid, void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(logits && target && loss && row_state && n_valid && ws, "qb_cross_entropy_fwd: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0, "qb_cross_entropy_fwd: B=%lld C=%lld", (long long)B, (long long)C);
    if (ws_bytes < qb_cross_entropy_ws_bytes(B)) {
        qb_set_error("qb_cross_entropy_fwd: workspace %zu < %zu bytes", ws_bytes, qb_cross_entropy_ws_bytes(B));
        return QB_ERR_WORKSPACE;
    }
    cudaStream_t st = qb_stream(stream);
    float* row_loss = static_cast<float*>(ws);
    k_ce_rows_fwd<<<(unsigned)((B + CE_WARPS - 1) / CE_WARPS), CE_WARPS * 32, 0, st>>>(logits, target, B, C, ignore_index, row_state, row_loss);
    QB_LAUNCH_CHECK();
    k_ce_mean<<<1, 256, 0, st>>>(row_loss, target, B, ignore_index, loss, n_valid);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* dlogits [B, C] = d loss / d logits * grad_out (device float, the gradient arriving at the loss). */
extern "C" int qb_cross_entropy_bwd(const float* logits, const int64_t* target, const float* row_state, const float* n_valid,
                                    const float* grad_out, float* dlogits, int64_t B, int64_t C, int64_t ignore_index, void* stream) {
    QB_CHECK_ARG(logits && target && row_state && n_valid && grad_out && dlogits, "qb_cross_entropy_bwd: null pointer");
    QB_CHECK_ARG(B > 0 && C > 0, "qb_cross_entropy_bwd: B=%lld C=%lld", (long long)B, (long long)C);
    k_ce_rows_bwd<<<(unsigned)((B + CE_WARPS - 1) / CE_WARPS), CE_WARPS * 32, 0, qb_stream(stream)>>>(logits, target, row_state, n_valid,
                                                                                                   grad_out, B, C, ignore_index, dlogits);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
