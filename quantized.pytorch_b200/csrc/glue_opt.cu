// glue_opt.cu - SURVEY.md section 8(f) N1: the optimizer / regulariser glue of the training step (main.py:362-380) as
// multi-tensor kernels: ONE launch for the L2 regulariser 0.5*wd*SUM w^2 of every conv / linear weight and bias
// (models/resnet_quantized.py:157-165; 66 pow / sum / add launches in ATen), ONE launch for its gradient, ONE launch for the
// SGD-with-momentum update of every parameter (torch.optim.SGD, dampening 0, no Nesterov).  Tensor lists travel by value in
// the kernel parameters (no device-side table to keep in sync with autograd's freshly allocated gradients).
#include "common.cuh"

constexpr int MT_MAX = 96;            // tensors per launch (kernel parameter space: 96 * 28 B + offsets < 4 KB)
constexpr int MT_CHUNK = 8192;        // elements per CTA
constexpr int MT_THREADS = 256;

struct MtList {
    const float* a[MT_MAX];           // read-only operand (weights / gradients)
    float* b[MT_MAX];                 // first in/out operand
    float* c[MT_MAX];                 // second in/out operand
    int n[MT_MAX];
    int first_block[MT_MAX + 1];      // CTA range of tensor t: [first_block[t], first_block[t+1])
    int count;
};

__device__ __forceinline__ int mt_find(const MtList& l, int blk) {
    int lo = 0, hi = l.count - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (l.first_block[mid] <= blk) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// partial[blk] = SUM over this CTA's chunk of w^2 (fp64, fixed order); k_mt_sumsq_final adds the partials in order
__global__ void __launch_bounds__(MT_THREADS) k_mt_sumsq(MtList l, double* __restrict__ partial) {
    __shared__ double sm[MT_THREADS / 32];
    const int t = mt_find(l, blockIdx.x);
    const int base = (blockIdx.x - l.first_block[t]) * MT_CHUNK;
    const float* w = l.a[t];
    const int end = min(base + MT_CHUNK, l.n[t]);
    double acc = 0.0;
    for (int i = base + threadIdx.x; i < end; i += MT_THREADS) {
        const float v = w[i];
        acc += (double)v * (double)v;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int k = 0; k < MT_THREADS / 32; ++k) s += sm[k];
        partial[blockIdx.x] = s;
    }
}
__global__ void k_mt_sumsq_final(const double* __restrict__ partial, int nblocks, double half_wd, float* __restrict__ out, int accumulate) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double s = 0.0;
    for (int k = 0; k < nblocks; ++k) s += partial[k];
    const float v = (float)(half_wd * s);
    out[0] = accumulate ? out[0] + v : v;
}

// grad_w = (upstream * wd) * w for every tensor (the gradient of 0.5*wd*SUM w^2 scaled by the incoming scalar gradient)
__global__ void __launch_bounds__(MT_THREADS) k_mt_l2_grad(MtList l, const float* __restrict__ upstream, float wd) {
    const int t = mt_find(l, blockIdx.x);
    const int base = (blockIdx.x - l.first_block[t]) * MT_CHUNK;
    const int end = min(base + MT_CHUNK, l.n[t]);
    const float s = (upstream ? __ldg(upstream) : 1.0f) * wd;
    const float* w = l.a[t];
    float* g = l.b[t];
    for (int i = base + threadIdx.x; i < end; i += MT_THREADS) g[i] = s * w[i];
}

// torch.optim.SGD (momentum > 0, dampening 0, nesterov False, maximize False), one rounding per torch kernel:
//   g' = g + wd * p  (only where wd != 0)      buf = first ? g' : buf * momentum + g'      p = p - lr * buf
__global__ void __launch_bounds__(MT_THREADS)
k_mt_sgd(MtList l, float lr, float momentum, float wd, int first) {
    const int t = mt_find(l, blockIdx.x);
    const int base = (blockIdx.x - l.first_block[t]) * MT_CHUNK;
    const int end = min(base + MT_CHUNK, l.n[t]);
    const float* g = l.a[t];
    float* p = l.b[t];
    float* buf = l.c[t];
    for (int i = base + threadIdx.x; i < end; i += MT_THREADS) {
        float gi = g[i];
        const float pi = p[i];
        if (wd != 0.0f) gi = fmaf(wd, pi, gi);                       // grad.add(param, alpha=wd): ATen's a + alpha*b contracts to one FMA
        float b = gi;
        if (momentum != 0.0f) {
            if (!first) b = __fadd_rn(__fmul_rn(buf[i], momentum), gi);
            buf[i] = b;
        }
        p[i] = fmaf(-lr, b, pi);                                     // param.add_(buf, alpha=-lr)
    }
}

static int mt_fill(MtList& l, int count, const void* const* a, void* const* b, void* const* c, const int64_t* sizes, const char* who) {
    QB_CHECK_ARG(count > 0 && count <= MT_MAX, "%s: %d tensors per call (1..%d)", who, count, MT_MAX);
    l.count = count;
    int blocks = 0;
    for (int t = 0; t < count; ++t) {
        QB_CHECK_ARG(sizes[t] > 0 && sizes[t] < 2147483647LL, "%s: tensor %d has %lld elements", who, t, (long long)sizes[t]);
        l.a[t] = static_cast<const float*>(a[t]);
        l.b[t] = b ? static_cast<float*>(b[t]) : nullptr;
        l.c[t] = c ? static_cast<float*>(c[t]) : nullptr;
        l.n[t] = (int)sizes[t];
        l.first_block[t] = blocks;
        blocks += (int)((sizes[t] + MT_CHUNK - 1) / MT_CHUNK);
    }
    l.first_block[count] = blocks;
    return QB_OK;
}

extern "C" size_t qb_mt_ws_bytes(int count, const int64_t* sizes) {
    size_t blocks = 0;
    for (int t = 0; t < count; ++t) blocks += (size_t)((sizes[t] + MT_CHUNK - 1) / MT_CHUNK);
    return blocks * sizeof(double) + 64;
}

/* out[0] (= or +=) 0.5 * wd * SUM_t SUM_i w_t[i]^2 over `count` (<= 96) fp32 tensors; deterministic.  ws: qb_mt_ws_bytes(). */
extern "C" int qb_l2_sumsq(const void* const* w, const int64_t* sizes, int count, double weight_decay, float* out, int accumulate,
                           void* ws, size_t ws_bytes, void* stream) {
    QB_CHECK_ARG(w && sizes && out && ws, "qb_l2_sumsq: null pointer");
    MtList l;
    int rc = mt_fill(l, count, w, nullptr, nullptr, sizes, "qb_l2_sumsq");
    if (rc) return rc;
    const int blocks = l.first_block[count];
    QB_CHECK_ARG(ws_bytes >= (size_t)blocks * sizeof(double), "qb_l2_sumsq: workspace too small");
    cudaStream_t st = qb_stream(stream);
    k_mt_sumsq<<<blocks, MT_THREADS, 0, st>>>(l, static_cast<double*>(ws));
    QB_LAUNCH_CHECK();
    k_mt_sumsq_final<<<1, 32, 0, st>>>(static_cast<const double*>(ws), blocks, 0.5 * weight_decay, out, accumulate);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* grad_t = (upstream[0] * wd) * w_t: the gradient of the regulariser (upstream = device scalar gradient of the loss, or NULL = 1). */
extern "C" int qb_l2_grad(const void* const* w, void* const* grad, const int64_t* sizes, int count, double weight_decay,
                          const float* upstream, void* stream) {
    QB_CHECK_ARG(w && grad && sizes, "qb_l2_grad: null pointer");
    MtList l;
    int rc = mt_fill(l, count, w, grad, nullptr, sizes, "qb_l2_grad");
    if (rc) return rc;
    k_mt_l2_grad<<<l.first_block[count], MT_THREADS, 0, qb_stream(stream)>>>(l, upstream, (float)weight_decay);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

/* One SGD step (torch.optim.SGD semantics, see the kernel) over `count` (<= 96) parameters: grad (read), param and momentum
 * buffer (updated in place).  first != 0: the momentum buffers are initialised with the gradient (torch's first step). */
extern "C" int qb_sgd_step(const void* const* grad, void* const* param, void* const* buf, const int64_t* sizes, int count, double lr,
                           double momentum, double weight_decay, int first, void* stream) {
    QB_CHECK_ARG(grad && param && sizes && (buf || momentum == 0.0), "qb_sgd_step: null pointer");
    MtList l;
    int rc = mt_fill(l, count, grad, param, buf, sizes, "qb_sgd_step");
    if (rc) return rc;
    k_mt_sgd<<<l.first_block[count], MT_THREADS, 0, qb_stream(stream)>>>(l, (float)lr, (float)momentum, (float)weight_decay, first);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

// ---------------------------------------------------------------------------------------------------------------------------
// qb_l2_sumsq_noalloc: qb_l2_sumsq for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The per-CTA partial sums
// live in a file-scope device array instead of a caller workspace: QB_L2_NOALLOC_MAX_BLOCKS doubles = 64 KB held by every GPU
// that loads the library until exit, good for 8192 CTAs * 8192 elements = 2^26 weights per call (ResNet-18: 1.2e7).
// In host code `g_l2_scratch` is only a stand-in for the symbol: its device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and handed to the kernels.
// Every call on a device shares the array.  Nothing in it has to start at zero: each CTA of k_mt_sumsq overwrites its own
// slot before k_mt_sumsq_final reads slots [0, blocks) on the same stream, and stale slots beyond `blocks` are never read.
// Two calls that overlap on the GPU (different streams, same device) race on the array - serialise them.
constexpr int QB_L2_NOALLOC_MAX_BLOCKS = 8192;
constexpr int QB_MAX_DEVICES = 64;
static __device__ __align__(256) double g_l2_scratch[QB_L2_NOALLOC_MAX_BLOCKS];
static double* g_l2_scratch_addr[QB_MAX_DEVICES];          // device address of g_l2_scratch per device ordinal (0 = not yet set up)

static int l2_scratch_for_current_device(double** out) {
    int dev = -1;
    QB_CUDA(cudaGetDevice(&dev));
    QB_CHECK_ARG(dev >= 0 && dev < QB_MAX_DEVICES, "qb_l2_sumsq_noalloc: device ordinal %d (max %d)", dev, QB_MAX_DEVICES - 1);
    double* p = __atomic_load_n(&g_l2_scratch_addr[dev], __ATOMIC_ACQUIRE);
    if (!p) {                                               // set-up, once per device (racing threads look up the same address)
        QB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_l2_scratch));
        QB_CHECK_ARG(p != nullptr, "qb_l2_sumsq_noalloc: cudaGetSymbolAddress returned a null address");
        __atomic_store_n(&g_l2_scratch_addr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return QB_OK;
}

extern "C" int64_t qb_l2_sumsq_noalloc_max_blocks(void) { return QB_L2_NOALLOC_MAX_BLOCKS; }

/* Look the scratch address up for the CURRENT device now (optional: the first qb_l2_sumsq_noalloc call does it otherwise; a
 * caller that captures graphs calls this once per device before capturing). */
extern "C" int qb_l2_sumsq_noalloc_setup(void) {
    double* p = nullptr;
    return l2_scratch_for_current_device(&p);
}

/* out[0] (= or +=) 0.5 * wd * SUM_t SUM_i w_t[i]^2, as qb_l2_sumsq, without a workspace argument and without any allocation.
 * Rejects SUM_t ceil(sizes[t] / 8192) > 8192 (QB_ERR_ARG, message in qb_last_error()). */
extern "C" int qb_l2_sumsq_noalloc(const void* const* w, const int64_t* sizes, int count, double weight_decay, float* out,
                                   int accumulate, void* stream) {
    QB_CHECK_ARG(w && sizes && out, "qb_l2_sumsq_noalloc: null pointer");
    MtList l;
    int rc = mt_fill(l, count, w, nullptr, nullptr, sizes, "qb_l2_sumsq_noalloc");
    if (rc) return rc;
    const int blocks = l.first_block[count];
    QB_CHECK_ARG(blocks <= QB_L2_NOALLOC_MAX_BLOCKS,
                 "qb_l2_sumsq_noalloc: %d chunks of %d elements exceed the static scratch (%d chunks = %lld elements); use qb_l2_sumsq",
                 blocks, MT_CHUNK, QB_L2_NOALLOC_MAX_BLOCKS, (long long)QB_L2_NOALLOC_MAX_BLOCKS * MT_CHUNK);
    double* scratch = nullptr;
    rc = l2_scratch_for_current_device(&scratch);
    if (rc) return rc;
    cudaStream_t st = qb_stream(stream);
    k_mt_sumsq<<<blocks, MT_THREADS, 0, st>>>(l, scratch);
    QB_LAUNCH_CHECK();
    k_mt_sumsq_final<<<1, 32, 0, st>>>(scratch, blocks, 0.5 * weight_decay, out, accumulate);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
