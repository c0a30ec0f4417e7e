// gemm_i8.cu - QLinear's contraction as an int8 tcgen05 GEMM with the asymmetric dequantisation in the epilogue.
//
// Reference: F.linear(x_hat, w_hat, b_hat) on DEQUANTISED fp32 tensors (models/modules/quantize.py:424, :152-154).
// Both operands lie on <=256-level affine grids, x_hat = qa*sa + la (per tensor), w_hat[n,:] = qw*sw[n] + lw[n]
// (per output row), so with integer code sums
//     out[m,n] = sa*sw[n]*SUM_k qa qw + sa*lw[n]*SUM_k qa[m,k] + la*sw[n]*SUM_k qw[n,k] + K*la*lw[n] + bias[n]
// The SUM_k qa qw term is an exact uint8 x uint8 -> int32 GEMM on the 5th-generation tensor cores:
//   warp 0   : TMA producer (cp.async.bulk.tensor, 128-byte swizzle) into a multi-stage shared-memory ring
//   warp 1   : TMEM allocation + single-thread tcgen05.mma.kind::i8 issue, accumulator in TMEM (128 lanes x 128 cols)
//   warps 2-5: epilogue - tcgen05.ld the accumulator, apply the formula in fp64, store fp32
#include <stdlib.h>

#include "tc_common.cuh"

// ------------------------------------------------------------------------------------------------
// host: driver entry point + tensor-map helper
// ------------------------------------------------------------------------------------------------
qb_encode_tiled_fn qb_get_encode_tiled() {
    static qb_encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
        if (e == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = reinterpret_cast<qb_encode_tiled_fn>(p);
        else (void)cudaGetLastError();
    }
    if (!fn) qb_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return fn;
}

int qb_make_tmap(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides, int swizzle_bytes) {
    qb_encode_tiled_fn enc = qb_get_encode_tiled();
    if (!enc) return QB_ERR_CUDA;
    cuuint64_t gdim[5], gstr[5];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        es[i] = elem_strides ? elem_strides[i] : 1;
        if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
    }
    CUresult r = enc(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B),
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        qb_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rank %d, dims %llu x %llu, box %u x %u)", (int)r, rank,
                     (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0);
        return QB_ERR_CUDA;
    }
    return QB_OK;
}

// ------------------------------------------------------------------------------------------------
// kernel
// ------------------------------------------------------------------------------------------------
constexpr int G_BM = 128, G_BN = 128, G_BK = 128;      // tile: 128 x 128 outputs, 128 bytes of K per stage
constexpr int G_STAGES = 6;                            // 6 x 32 KB = 192 KB of operand staging
constexpr int G_UMMA_K = 32;                           // K per tcgen05.mma for 8-bit operands
constexpr int G_THREADS = 192;                         // 6 warps
constexpr uint32_t G_TMEM_COLS = 128;
constexpr uint32_t G_STAGE_BYTES = (G_BM + G_BN) * G_BK;
constexpr size_t G_SMEM_BYTES = (size_t)G_STAGES * G_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;

struct GemmEpilogue {
    const float* a_range;      // device {min,max} of the activation quantizer (QB_SCALE_DOUBLE)
    int a_bits;
    const float* b_row_min;    // per output row (QB_SCALE_FLOAT)
    const float* b_row_max;
    int b_bits;
    const int32_t* a_rowsum;   // SUM_k qa[m,k]
    const int32_t* b_rowsum;   // SUM_k qw[n,k]
    const float* bias;         // already fake-quantized bias or nullptr
    int b_offset;              // 0: B holds raw unsigned codes; 2^(bits-1): B holds centred signed codes (q - offset)
};

__global__ void __launch_bounds__(G_THREADS, 1)
k_gemm_i8(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, float* __restrict__ C,
          int M, int N, int K, int64_t ldc, GemmEpilogue ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;                                   // [stage][128 rows][128 B]
    uint8_t* smem_b = smem + (size_t)G_STAGES * G_BM * G_BK;  // [stage][128 rows][128 B]
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + (size_t)G_STAGES * G_STAGE_BYTES);
    uint64_t* empty_bar = full_bar + G_STAGES;
    uint64_t* accum_bar = empty_bar + G_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * G_BM, n0 = blockIdx.x * G_BN;
    const int num_kb = (K + G_BK - 1) / G_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        for (int s = 0; s < G_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(accum_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<G_TMEM_COLS>(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (elect_one()) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % G_STAGES;
                const uint32_t ph = (kb / G_STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);                       // slot free (first lap passes immediately)
                mbar_arrive_expect_tx(&full_bar[s], G_STAGE_BYTES);
                tma_load_2d(smem_a + (size_t)s * G_BM * G_BK, &tmap_a, &full_bar[s], kb * G_BK, m0);
                tma_load_2d(smem_b + (size_t)s * G_BN * G_BK, &tmap_b, &full_bar[s], kb * G_BK, n0);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        const uint32_t idesc = umma_idesc(/*S32*/ 2, /*A u8*/ 0, /*B*/ ep.b_offset ? 1u : 0u, G_BM, G_BN);
        for (int kb = 0; kb < num_kb; ++kb) {
            const int s = kb % G_STAGES;
            const uint32_t ph = (kb / G_STAGES) & 1;
            mbar_wait(&full_bar[s], ph);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t a_addr = smem_u32(smem_a + (size_t)s * G_BM * G_BK);
                const uint32_t b_addr = smem_u32(smem_b + (size_t)s * G_BN * G_BK);
#pragma unroll
                for (int k = 0; k < G_BK / G_UMMA_K; ++k) {
                    const uint64_t da = umma_smem_desc_sw128(a_addr + k * G_UMMA_K);
                    const uint64_t db = umma_smem_desc_sw128(b_addr + k * G_UMMA_K);
                    umma_i8(tmem_base, da, db, idesc, (kb | k) != 0);
                }
                umma_commit(&empty_bar[s]);                             // frees the smem slot when the MMAs retire
                if (kb == num_kb - 1) umma_commit(accum_bar);           // accumulator complete
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: warp w may touch TMEM lanes [32*(w%4), +32) =====
        const int sub = warp & 3;
        const int m = m0 + sub * 32 + lane;
        mbar_wait(accum_bar, 0);
        tc_fence_after_sync();
        const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
        const double sa = (double)pa.scale, la = (double)pa.lo;
        const double rs_a = (m < M) ? (double)__ldg(ep.a_rowsum + m) : 0.0;
#pragma unroll 1
        for (int c = 0; c < G_BN; c += 32) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(sub * 32) << 16) + (uint32_t)c, r);
            tmem_ld_wait();
            if (m < M) {
                float* dst = C + (int64_t)m * ldc + n0 + c;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = n0 + c + j;
                    if (n < N) {
                        const QParams pb = make_qparams_f(__ldg(ep.b_row_min + n), __ldg(ep.b_row_max + n), ep.b_bits);
                        const double sw = (double)pb.scale, lw = (double)pb.lo + (double)ep.b_offset * (double)pb.scale;
                        double v = sa * sw * (double)(int32_t)r[j] + sa * lw * rs_a + la * sw * (double)__ldg(ep.b_rowsum + n) +
                                   (double)K * la * lw;
                        if (ep.bias) v += (double)__ldg(ep.bias + n);
                        dst[j] = (float)v;
                    }
                }
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) tmem_dealloc<G_TMEM_COLS>(tmem_base);
}

// SUM over each row of uint8 codes (the epilogue's correction terms).  Short rows (a pixel's channels): one thread per
// row - consecutive threads read consecutive rows, so a warp still covers one contiguous span; long rows: one warp per row.
__device__ __forceinline__ uint32_t sum16(const uint4 v, uint32_t acc) {
    acc = __dp4a(v.x, 0x01010101u, acc);
    acc = __dp4a(v.y, 0x01010101u, acc);
    acc = __dp4a(v.z, 0x01010101u, acc);
    return __dp4a(v.w, 0x01010101u, acc);
}

__global__ void __launch_bounds__(256) k_rowsum_u8_thread(const uint8_t* __restrict__ codes, int64_t rows, int cols16, int32_t* out) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const uint4* p = reinterpret_cast<const uint4*>(codes + row * (int64_t)cols16 * 16);
    uint32_t acc = 0;
    for (int j = 0; j < cols16; ++j) acc = sum16(__ldcs(p + j), acc);
    out[row] = (int32_t)acc;
}

__global__ void k_rowsum_u8(const uint8_t* __restrict__ codes, int64_t rows, int64_t cols, int32_t* out) {
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const uint8_t* p = codes + row * cols;
    uint32_t acc = 0;
    if ((cols & 15) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        for (int64_t j = lane; j < (cols >> 4); j += 32) acc = sum16(__ldg(reinterpret_cast<const uint4*>(p) + j), acc);
    } else {
        for (int64_t j = lane; j < cols; j += 32) acc += p[j];
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[row] = (int32_t)acc;
}

extern "C" int qb_rowsum_u8(const uint8_t* codes, int64_t rows, int64_t cols, int32_t* out, void* stream) {
    QB_CHECK_ARG(codes && out && rows > 0 && cols > 0, "qb_rowsum_u8: bad arguments");
    if ((cols & 15) == 0 && cols <= 512 && (reinterpret_cast<uintptr_t>(codes) & 15) == 0)
        k_rowsum_u8_thread<<<(unsigned)((rows + 255) / 256), 256, 0, qb_stream(stream)>>>(codes, rows, (int)(cols >> 4), out);
    else
        k_rowsum_u8<<<(unsigned)((rows * 32 + 255) / 256), 256, 0, qb_stream(stream)>>>(codes, rows, cols, out);
    QB_LAUNCH_CHECK();
    return QB_OK;
}

int qb_gemm_i8_pair_launch(const uint8_t* A, const uint8_t* B, float* C, int64_t M, int64_t N, int64_t K, int64_t ldc,
                           const float* a_range, int a_bits, const float* b_row_min, const float* b_row_max, int b_bits,
                           const int32_t* a_rowsum, const int32_t* b_rowsum, const float* bias, int b_offset, cudaStream_t st);

// QB_GEMM_KERNEL=single|pair pins the kernel choice (tests, sweeps); default: CTA pairs once there are >= 8 pair tiles
static int gemm_kernel_override() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("QB_GEMM_KERNEL");
        v = !e ? 0 : (e[0] == 's' ? 1 : (e[0] == 'p' ? 2 : 0));
    }
    return v;
}

extern "C" int qb_gemm_i8(const uint8_t* A, const uint8_t* B, float* C, int64_t M, int64_t N, int64_t K, int64_t ldc,
                          const float* a_range, int a_bits, const float* b_row_min, const float* b_row_max, int b_bits,
                          const int32_t* a_rowsum, const int32_t* b_rowsum, const float* bias, int b_offset, void* stream) {
    QB_CHECK_ARG(A && B && C && a_range && b_row_min && b_row_max && a_rowsum && b_rowsum, "qb_gemm_i8: null pointer");
    QB_CHECK_ARG(M > 0 && N > 0 && K > 0 && ldc >= N, "qb_gemm_i8: bad shape M=%lld N=%lld K=%lld ldc=%lld", (long long)M,
                 (long long)N, (long long)K, (long long)ldc);
    QB_CHECK_ARG(a_bits >= 1 && a_bits <= 8 && b_bits >= 1 && b_bits <= 8, "qb_gemm_i8: codes are uint8 (bits<=8)");
    if ((K & 15) != 0 || ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(B)) & 15) != 0) {
        qb_set_error("qb_gemm_i8: K=%lld must be a multiple of 16 and A/B 16-byte aligned (TMA row pitch)", (long long)K);
        return QB_ERR_UNSUPPORTED;
    }
    QB_CHECK_ARG(M < 2147483647LL && N < 2147483647LL && K < 2147483647LL, "qb_gemm_i8: dimension exceeds int32");
    QB_CHECK_ARG(b_offset == 0 || b_offset == (1 << (b_bits - 1)), "qb_gemm_i8: b_offset must be 0 or 2^(b_bits-1)");
    {
        const bool pair_ok = (ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0;
        const int64_t pair_tiles = ((M + 255) / 256) * ((N + 255) / 256);
        const int ov = gemm_kernel_override();
        if (pair_ok && (ov == 2 || (ov == 0 && pair_tiles >= 8)))
            return qb_gemm_i8_pair_launch(A, B, C, M, N, K, ldc, a_range, a_bits, b_row_min, b_row_max, b_bits, a_rowsum,
                                          b_rowsum, bias, b_offset, qb_stream(stream));
    }
    CUtensorMap ta, tb;
    const uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, db[2] = {(uint64_t)K, (uint64_t)N}, st[1] = {(uint64_t)K};
    const uint32_t box[2] = {(uint32_t)G_BK, (uint32_t)G_BM};
    int rc = qb_make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, A, da, st, box, nullptr);
    if (rc) return rc;
    rc = qb_make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, B, db, st, box, nullptr);
    if (rc) return rc;
    QB_CUDA(cudaFuncSetAttribute(k_gemm_i8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G_SMEM_BYTES));   // per device: not cached
    GemmEpilogue ep{a_range, a_bits, b_row_min, b_row_max, b_bits, a_rowsum, b_rowsum, bias, b_offset};
    dim3 grid((unsigned)((N + G_BN - 1) / G_BN), (unsigned)((M + G_BM - 1) / G_BM));
    k_gemm_i8<<<grid, G_THREADS, G_SMEM_BYTES, qb_stream(stream)>>>(ta, tb, C, (int)M, (int)N, (int)K, ldc, ep);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
