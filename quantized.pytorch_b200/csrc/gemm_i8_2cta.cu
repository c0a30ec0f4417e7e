// gemm_i8_2cta.cu - the large-shape QLinear contraction: a persistent int8 tcgen05 GEMM in which the two SMs of a TPC
// work as ONE CTA pair (cta_group::2).  Same algebra and arguments as k_gemm_i8 (gemm_i8.cu; reference
// models/modules/quantize.py:424, :152-154); this file only changes how the integer term SUM_k qa*qw is produced.
//
// Why a pair: at M=N=K=8192 a 128x128 tile moves 256 operand bytes per 128x128 MACs - 64 MAC/B - and 148 SMs at full
// tensor rate would need ~35 TB/s from L2 (the chip delivers ~12).  A pair computes a 256 x 256 tile with
// tcgen05.mma.cta_group::2 (UMMA M=256, N=256): each CTA stages only ITS 128 rows of A and ITS 128 rows of B, the
// tensor cores of both SMs read both halves, so a CTA moves 32 KB per 128x256x128 MACs - 128 MAC/B.
//
// Roles per CTA (320 threads):
//   warp 0    TMA producer: its own A/B halves into its own shared memory; the transaction bytes of BOTH CTAs are
//             counted on the LEADER's (cluster rank 0) full barrier
//   warp 1    leader only: single-thread MMA issue; tcgen05.commit multicasts "stage free" / "accumulator ready" to the
//             barriers of both CTAs.  (In both CTAs this warp owns the pair-wide TMEM allocation.)
//   warps 2-9 epilogue: each CTA drains ITS 128 accumulator rows (two TMEM buffers of 256 columns, so tile j+1's MMAs
//             overlap tile j's epilogue), dequantises with per-column coefficients, stages [128][32] fp32 slabs in
//             swizzled shared memory and writes them with TMA stores (ragged edges clipped by the hardware)
#include "tc_common.cuh"

constexpr int G2_THREADS = 320;
constexpr int G2_BM = 128;                 // A rows per CTA (pair tile: 256)
constexpr int G2_BNH = 128;                // B rows per CTA
constexpr int G2_BN = 256;                 // accumulator columns (pair tile N)
constexpr int G2_BK = 128;                 // bytes of K per stage
constexpr int G2_STAGES = 5;
constexpr uint32_t G2_A_BYTES = G2_BM * G2_BK, G2_B_BYTES = G2_BNH * G2_BK, G2_STAGE_BYTES = G2_A_BYTES + G2_B_BYTES;
constexpr int G2_SLAB_BYTES = 128 * 128;   // [128 rows][32 fp32]
constexpr int G2_STAGING_BYTES = 2 * G2_SLAB_BYTES;      // one slab per epilogue group
constexpr int G2_TAIL_BYTES = 256 + G2_BN * 16;          // barriers + float4 coefficients
constexpr size_t G2_SMEM_BYTES = (size_t)G2_STAGES * G2_STAGE_BYTES + G2_STAGING_BYTES + G2_TAIL_BYTES + 1024;
constexpr uint32_t G2_TMEM_COLS = 512;
constexpr int G2_GROUP_M = 8;              // tile raster: bands of 8 pair-tiles in M, N fastest inside a band

struct Gemm2Epilogue {
    const float* a_range; int a_bits;
    const float* b_row_min; const float* b_row_max; int b_bits;
    const int32_t* a_rowsum; const int32_t* b_rowsum; const float* bias; int b_offset;
};

// ---- cluster / cta_group::2 primitives -------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t local_smem_addr, uint32_t rank) {     // address of the same object in CTA `rank`
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t leader_bar_cluster_addr, int c0,
                                                 int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(leader_bar_cluster_addr), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* slot_in_smem) {        // one full warp in EACH CTA of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_in_smem)), "n"(G2_TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(G2_TMEM_COLS));
}
__device__ __forceinline__ void umma_i8_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {                // arrives on `bar` in BOTH CTAs
    const uint16_t mask = 3;
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void g2_bar(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void g2_bar_all() { asm volatile("bar.sync 3, 256;" ::: "memory"); }
__device__ __forceinline__ float4 g2_lds128(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__device__ __forceinline__ void g2_tile(int t, int tiles_m, int tiles_n, int& tm, int& tn) {
    const int per_band = G2_GROUP_M * tiles_n;
    const int band = t / per_band, within = t - band * per_band;
    int rows = tiles_m - band * G2_GROUP_M;
    if (rows > G2_GROUP_M) rows = G2_GROUP_M;
    tn = within / rows;
    tm = band * G2_GROUP_M + (within - tn * rows);
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G2_THREADS, 1)
k_gemm_i8_pair(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_c, int M, int N, int K, Gemm2Epilogue ep) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* staging = smem + (size_t)G2_STAGES * G2_STAGE_BYTES;
    uint8_t* tail = staging + G2_STAGING_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(tail);
    uint64_t* empty_bar = full_bar + G2_STAGES;
    uint64_t* tfull_bar = empty_bar + G2_STAGES;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
    float4* s_coef = reinterpret_cast<float4*>(tail + 256);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int num_pairs = gridDim.x >> 1, pair = blockIdx.x >> 1;
    const int tiles_m = (M + 2 * G2_BM - 1) / (2 * G2_BM), tiles_n = (N + G2_BN - 1) / G2_BN;
    const int total_tiles = tiles_m * tiles_n;
    const int num_kb = (K + G2_BK - 1) / G2_BK;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmap_a);
        tma_prefetch_desc(&tmap_b);
        tma_prefetch_desc(&tmap_c);
        for (int s = 0; s < G2_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);               // the leader's own arrive.expect_tx (used in the leader only)
            mbar_init(&empty_bar[s], 1);              // one multicast commit
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tfull_bar[i], 1);              // one multicast commit
            mbar_init(&tempty_bar[i], 16);            // 8 epilogue warps x 2 CTAs (used in the leader only)
        }
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc_pair(tmem_slot);
    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();                               // the peer's barriers exist before anything is signalled on them
    tc_fence_after_sync();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (elect_one()) {
            uint32_t it = 0;
            for (int t = pair; t < total_tiles; t += num_pairs) {
                int tm, tn;
                g2_tile(t, tiles_m, tiles_n, tm, tn);
                const int m_row = tm * 2 * G2_BM + (int)rank * G2_BM, n_row = tn * G2_BN + (int)rank * G2_BNH;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % G2_STAGES;
                    const uint32_t ph = (it / G2_STAGES) & 1;
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    if (leader) mbar_arrive_expect_tx(&full_bar[s], 2 * G2_STAGE_BYTES);
                    const uint32_t bar = mapa_rank(smem_u32(&full_bar[s]), 0);
                    uint8_t* sa = smem + (size_t)s * G2_STAGE_BYTES;
                    tma_load_2d_pair(sa, &tmap_a, bar, kb * G2_BK, m_row);
                    tma_load_2d_pair(sa + G2_A_BYTES, &tmap_b, bar, kb * G2_BK, n_row);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader) {
            const uint32_t idesc = umma_idesc(/*S32*/ 2, /*A u8*/ 0, /*B*/ ep.b_offset ? 1u : 0u, 2 * G2_BM, G2_BN);
            uint32_t it = 0;
            int j = 0;
            for (int t = pair; t < total_tiles; t += num_pairs, ++j) {
                const int buf = j & 1;
                mbar_wait(&tempty_bar[buf], ((j >> 1) & 1) ^ 1);
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem_base + (uint32_t)(buf * G2_BN);
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const int s = it % G2_STAGES;
                    const uint32_t ph = (it / G2_STAGES) & 1;
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint32_t a_addr = smem_u32(smem + (size_t)s * G2_STAGE_BYTES);
                        const uint32_t b_addr = a_addr + G2_A_BYTES;
#pragma unroll
                        for (int k = 0; k < G2_BK / 32; ++k)
                            umma_i8_pair(d_tmem, umma_smem_desc_sw128(a_addr + k * 32), umma_smem_desc_sw128(b_addr + k * 32),
                                         idesc, (kb | k) != 0);
                        umma_commit_pair(&empty_bar[s]);
                        if (kb == num_kb - 1) umma_commit_pair(&tfull_bar[buf]);
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ===== epilogue (both CTAs): 2 groups x 4 warps, group `grp` takes the column slabs cs with cs % 2 == grp =====
        const int e_all = threadIdx.x - 64, grp = e_all >> 7, et = e_all & 127;
        const int sub = warp & 3;
        const int i = sub * 32 + lane;                                  // accumulator row = TMEM lane = staging row
        const QParams pa = make_qparams_d((double)__ldg(ep.a_range), (double)__ldg(ep.a_range + 1), ep.a_bits);
        const double sa = (double)pa.scale, la = (double)pa.lo;
        const int32_t half = 1 << (ep.b_bits - 1);
        uint8_t* my_staging = staging + grp * G2_SLAB_BYTES;
        const uint32_t dst = smem_u32(my_staging) + (uint32_t)i * 128u, sw = (uint32_t)(i & 7);
        const uint32_t coef_u32 = smem_u32(s_coef);
        const uint32_t tempty_leader0 = mapa_rank(smem_u32(&tempty_bar[0]), 0);
        const int bar_id = 1 + grp;
        int j = 0;
        for (int t = pair; t < total_tiles; t += num_pairs, ++j) {
            int tm, tn;
            g2_tile(t, tiles_m, tiles_n, tm, tn);
            const int m0 = tm * 2 * G2_BM + (int)rank * G2_BM, n0 = tn * G2_BN;
            const int m = m0 + i;
            const int32_t rs_ai = (m < M) ? __ldg(ep.a_rowsum + m) : 0;
            const float rs_a = (float)rs_ai;
            // raw unsigned B codes are centred HERE, exactly, in integers: SUM qa*(qw - h) = acc - h*SUM qa.  The fp32
            // formula then has no large cancelling terms (weights are nearly symmetric about the middle code).
            const int32_t centre = ep.b_offset ? 0 : half * rs_ai;
            g2_bar_all();                                               // everyone is done with the previous coefficients
            {
                const int n = n0 + e_all;
                float4 cf = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n < N) {
                    const QParams pb = make_qparams_f(__ldg(ep.b_row_min + n), __ldg(ep.b_row_max + n), ep.b_bits);
                    const double swd = (double)pb.scale, lw = (double)pb.lo + (double)half * (double)pb.scale;
                    const double rs_b = (double)__ldg(ep.b_rowsum + n) - (ep.b_offset ? 0.0 : (double)half * (double)K);
                    cf.x = (float)(sa * swd);
                    cf.y = (float)(sa * lw);
                    cf.z = (float)(la * swd * rs_b + (double)K * la * lw +
                                   (ep.bias ? (double)__ldg(ep.bias + n) : 0.0));
                }
                s_coef[e_all] = cf;
            }
            g2_bar_all();
            const int buf = j & 1;
            mbar_wait(&tfull_bar[buf], (j >> 1) & 1);
            tc_fence_after_sync();
            const uint32_t t_addr = tmem_base + (uint32_t)(buf * G2_BN) + ((uint32_t)(sub * 32) << 16);
            const int ncols = (N - n0 < G2_BN) ? N - n0 : G2_BN;
            const int nslabs = (ncols + 31) >> 5;
            const int last_cs = ((nslabs - 1 - grp) & ~1) + grp;
            const bool rows_live = m0 < M;                              // a CTA whose 128 rows are all padding stores nothing
            if (nslabs <= grp || !rows_live) {
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(tempty_leader0 + (uint32_t)buf * 8u);
            }
            if (!rows_live) continue;
#pragma unroll 1
            for (int cs = grp; cs < nslabs; cs += 2) {
                const int c = cs * 32;
                uint32_t r[32];
                tmem_ld_32x32(t_addr + (uint32_t)c, r);
                tmem_ld_wait();
                if (cs == last_cs) {                                    // this warp's last read of the buffer
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(tempty_leader0 + (uint32_t)buf * 8u);
                }
                const uint32_t cf_addr = coef_u32 + (uint32_t)c * 16u;
#pragma unroll
                for (int q = 0; q < 32; ++q) {
                    const float4 cf = g2_lds128(cf_addr + q * 16);
                    r[q] = __float_as_uint(fmaf(cf.x, (float)((int32_t)r[q] - centre), fmaf(cf.y, rs_a, cf.z)));
                }
                if (et == 0) tma_store_wait_read<0>();                  // the group's previous store has left the slab
                g2_bar(bar_id);
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst + (((uint32_t)q ^ sw) << 4)),
                                 "r"(r[4 * q]), "r"(r[4 * q + 1]), "r"(r[4 * q + 2]), "r"(r[4 * q + 3]) : "memory");
                fence_proxy_async();
                g2_bar(bar_id);
                if (et == 0) {
                    tma_store_2d(&tmap_c, my_staging, n0 + c, m0);
                    tma_store_commit();
                }
            }
        }
        if (et == 0) tma_store_wait<0>();
    }
    tc_fence_before_sync();
    __syncthreads();
    cluster_sync_all();                               // no CTA of the pair leaves while the other may still touch it
    if (warp == 1) tmem_dealloc_pair(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
int qb_gemm_i8_pair_launch(const uint8_t* A, const uint8_t* B, float* C, int64_t M, int64_t N, int64_t K, int64_t ldc,
                           const float* a_range, int a_bits, const float* b_row_min, const float* b_row_max, int b_bits,
                           const int32_t* a_rowsum, const int32_t* b_rowsum, const float* bias, int b_offset, cudaStream_t st) {
    CUtensorMap ta, tb, tc;
    const uint64_t da[2] = {(uint64_t)K, (uint64_t)M}, db[2] = {(uint64_t)K, (uint64_t)N}, sk[1] = {(uint64_t)K};
    const uint32_t box[2] = {(uint32_t)G2_BK, (uint32_t)G2_BM};
    int rc = qb_make_tmap(&ta, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, A, da, sk, box, nullptr);
    if (rc) return rc;
    rc = qb_make_tmap(&tb, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, B, db, sk, box, nullptr);
    if (rc) return rc;
    const uint64_t dc[2] = {(uint64_t)N, (uint64_t)M}, sc[1] = {(uint64_t)ldc * 4};
    const uint32_t boxc[2] = {32u, 128u};
    rc = qb_make_tmap(&tc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, C, dc, sc, boxc, nullptr);
    if (rc) return rc;
    static int max_pairs = 0;
    if (!max_pairs) {
        QB_CUDA(cudaFuncSetAttribute(k_gemm_i8_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G2_SMEM_BYTES));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * 74);
        cfg.blockDim = dim3(G2_THREADS);
        cfg.dynamicSmemBytes = G2_SMEM_BYTES;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at;
        cfg.numAttrs = 1;
        int n = 0;
        QB_CUDA(cudaOccupancyMaxActiveClusters(&n, k_gemm_i8_pair, &cfg));
        if (n < 1) {
            qb_set_error("qb_gemm_i8: no CTA pair of k_gemm_i8_pair can be resident on this device");
            return QB_ERR_CUDA;
        }
        max_pairs = n;
    }
    const int64_t tiles = ((M + 2 * G2_BM - 1) / (2 * G2_BM)) * ((N + G2_BN - 1) / G2_BN);
    const int pairs = (int)(tiles < max_pairs ? tiles : max_pairs);
    Gemm2Epilogue ep{a_range, a_bits, b_row_min, b_row_max, b_bits, a_rowsum, b_rowsum, bias, b_offset};
    k_gemm_i8_pair<<<2 * pairs, G2_THREADS, G2_SMEM_BYTES, st>>>(ta, tb, tc, (int)M, (int)N, (int)K, ep);
    QB_LAUNCH_CHECK();
    return QB_OK;
}
