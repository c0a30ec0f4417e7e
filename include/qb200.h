/* qb200.h - C ABI of libqb200.so: the B200 (sm_100a) fake-quantization hot path.
 *
 * The reference (amishacorns/quantized.pytorch) is pure Python on ATen and has NO FFI of its own;
 * each entry point below names the reference code (models/modules/quantize.py:LINE) it replaces.
 * INTEGRATION.md shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is DEVICE memory unless it says "host";
 *  - the caller owns every buffer; the library keeps no reference past the stream work it enqueues;
 *  - every call takes the CUDA stream explicitly (void* = cudaStream_t) and never synchronises;
 *  - returns 0 on success, a negative qb_status otherwise; qb_last_error() gives the message of the
 *    last failure on the calling thread.  There is no CPU fallback: no device => QB_ERR_CUDA.
 *  - "range" is float[2] = {min,max} in device memory, so no value ever visits the host on the step path
 *    (the reference does float(tensor) twice per quantizer, quantize.py:126,132,249,486,493).
 */
#ifndef QB200_H
#define QB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    QB_OK = 0,
    QB_ERR_ARG = -1,         /* null pointer, non-positive size, unsupported bit width ...   */
    QB_ERR_UNSUPPORTED = -2, /* a shape/option this build has no kernel for                  */
    QB_ERR_CUDA = -3,        /* a CUDA runtime/driver call failed (message has the cudaError) */
    QB_ERR_WORKSPACE = -4    /* workspace too small (see the *_ws_bytes helpers)             */
} qb_status;

/* How the scale is derived from {min,max} - both exist in the reference (quantize.py:71-75):     */
#define QB_SCALE_DOUBLE 0 /* Python-float range: s = max((double)(max-min)/(2^b-1), 1e-8) -> fp32  */
#define QB_SCALE_FLOAT 1  /* tensor range:       s = clamp_min((max-min)/(float)(2^b-1), 1e-8f)    */

const char* qb_last_error(void);
int qb_version(void);
/* device properties the host side sizes launches with; fails with QB_ERR_CUDA if no sm_100 device */
int qb_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* l2_bytes);

/* ---------------------------------------------------------------------------------------------------
 * QuantMeasure statistics - quantize.py:226-235 (per-sample min/max averaged over the batch, global
 * mean and unbiased std; four ATen reductions + 2 host syncs there, one pass here).
 *   x        [B, inner] fp32 contiguous
 *   stats    double[8] out: {sum_b min_b, sum_b max_b, sum x, sum x^2, global min, global max, n, B}
 *   ws       workspace, qb_measure_ws_bytes(B) bytes, zero-filled ONCE by the caller before first use
 *            (the kernel leaves it clean); one workspace per concurrently running call.
 * Sums are fp64 and combined in a fixed order: results are run-to-run deterministic.            */
size_t qb_measure_ws_bytes(int64_t B);
int qb_measure_stats(const float* x, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes,
                     void* stream);

/* QuantMeasure buffer update + range selection - quantize.py:216-219,228-246,256-268.
 *   stats      double[8] from qb_measure_stats (possibly all-reduced over ranks: entries 0-3 SUM,
 *              4 MIN, 5 MAX, 6-7 SUM); ignored when training == 0
 *   running    float[5] views, in the order running_min, running_max, num_measurements, running_var,
 *              running_mean (each the reference's [1] buffer); updated in place when training
 *   momentum   < 0 => cumulative average n/(n+1) (momentum=None in the reference)
 *   method     0 'avg', 1 'aciq' (Laplace clipping, alpha table quantize.py:214; needs 2<=bits<=8)
 *   range_out  float[2] the {min,max} this call quantizes with                                     */
int qb_measure_finalize(const double* stats, float* running_min, float* running_max, float* num_measurements,
                        float* running_var, float* running_mean, double momentum, int training, int method,
                        int bits, float* range_out, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * UniformQuantize.forward, min-anchored deterministic branch - quantize.py:66-75,90,95,100
 * (what every quantize() call reaches, because of the argument shuffle at :159-160).
 *   y = rne(clamp((x - min) / s, 0, 2^bits-1)) * s + min, true fp32 division, unfused mul/add.
 *   range_dev  float[2] device {min,max}; if NULL the host doubles lo,hi are used instead
 *   codes      optional uint8 (bits<=8) integer codes, same indexing as x; may be NULL
 *   y          may alias x (in place) or be NULL when only codes are wanted                      */
int qb_fakequant_affine(const float* x, float* y, uint8_t* codes, int64_t n, int bits, const float* range_dev,
                        double lo, double hi, int scale_mode, void* stream);
/* Same, plus codes16 (optional): the integer codes as bf16 (exact for bits <= 8), the x operand of qb_conv_wgrad. */
int qb_fakequant_affine16(const float* x, float* y, uint8_t* codes, uint16_t* codes16, int64_t n, int bits,
                          const float* range_dev, double lo, double hi, int scale_mode, void* stream);
/* The activation quantizer of a network's FIRST QConv2d (quantize.py:339-343 on models/resnet_quantized.py:147,
 * models/mobilenet_quantized.py:66) fused with the layout change: x is the fp32 image batch as torch delivers it, NCHW
 * [B, 3, H*W]; codes are uint8 [B, H*W, 3] (channels-last) for qb_conv_stem_fwd.  Same codes as qb_fakequant_affine with a
 * device range on the channels-last copy; saves that copy (8 B/elem of the batch).  H*W % 4 == 0, bits <= 8. */
int qb_fakequant_affine_nchw3(const float* x, uint8_t* codes, int64_t B, int64_t HW, int bits, const float* range_dev, void* stream);
/* The activation quantizer of a tensor-core contraction (quantize.py:339-343) together with the per-row code sums its
 * zero-point correction needs: qb_fakequant_affine16(y = NULL, device range) + qb_rowsum_u8 in one pass over x [rows, cols]
 * (the sums are taken from the registers that hold the codes).  codes16 optional.  cols a power of two in 16..256, bits <= 8. */
int qb_fakequant_affine_rowsum_supported(int64_t cols, int bits);
int qb_fakequant_affine_rowsum(const float* x, uint8_t* codes, uint16_t* codes16, int64_t rows, int64_t cols, int bits,
                               const float* range_dev, int32_t* rowsum, void* stream);

/* The min-anchored branch with stochastic rounding (quantize.py:90-100, enforce_true_zero=False, stochastic=True; reached
 * through quantize(x, num_chunks=<truthy>), :159-160): y = rne(clamp((x - min)/s + u, 0, 2^bits-1)) * s + min, u = noise
 * (injected tensor, or in-kernel Philox as in qb_fakequant_grad).  y may alias x. */
int qb_fakequant_affine_stochastic(const float* x, float* y, int64_t n, int bits, const float* range_dev, double lo, double hi,
                                   int scale_mode, const float* noise, uint64_t seed, uint64_t offset, const uint64_t* step_dev,
                                   void* stream);

/* Self-test hook (no reference counterpart): the quantizers divide by the scale with an FMA sequence that must equal
 * IEEE division (ATen's div_, quantize.py:90,92); this evaluates that sequence and __fdiv_rn on the same operands
 * so a test can compare them bit for bit.  a, out_fast, out_ieee: fp32[n].                                          */
int qb_selftest_div(const float* a, float b, float* out_fast, float* out_ieee, int64_t n, void* stream);

/* Per-row variant for weights - quantize.py:318-323,332-334: row_min/row_max over flatten(1) are
 * computed here (and written out, they are the module's weight_min/weight_max buffers), then each
 * row is quantized with its own QB_SCALE_FLOAT scale.  If compute_range == 0 the given row_min/max
 * are used as is (frozen ranges / overwrite_params, quantize.py:301-312).                          */
int qb_fakequant_rows(const float* w, float* wq, uint8_t* codes, int64_t rows, int64_t cols, int bits,
                      float* row_min, float* row_max, int compute_range, void* stream);

/* Global min/max of a tensor -> range_out float[2].  UniformQuantizeGrad.backward's
 * float(g.min()), float(g.max()) (quantize.py:126,132) and the range-from-data case (:45-55).
 *   ws: qb_minmax_ws_bytes() bytes, zero-filled once by the caller.                               */
size_t qb_minmax_ws_bytes(void);
int qb_minmax(const float* x, int64_t n, float* range_out, void* ws, size_t ws_bytes, void* stream);

/* UniformQuantize.forward, zero-point + stochastic branch - quantize.py:77-88,92-98 (reached only
 * from UniformQuantizeGrad.backward :137-138):
 *   zp = trunc(clamp(-min/s, 0, 2^b-1)) in double;  q = rne(clamp(g/s + zp + u, 0, 2^b-1));  out = (q-zp)*s
 *   noise   optional fp32 tensor of u in [-0.5,0.5) to inject (parity tests replay the reference's
 *           uniform_ draws); if NULL u comes from Philox4x32-10 keyed by (seed, element index + offset);
 *           stochastic == 0 => u = 0
 *   step_dev optional device uint64 mixed into the Philox key (seed + step*golden) so that a captured
 *           CUDA graph draws fresh noise on every replay; may be NULL
 *   codes   optional uint8 codes q (bits<=8)                                                     */
int qb_fakequant_grad(const float* g, float* out, uint8_t* codes, int64_t n, int bits, const float* range_dev,
                      double lo, double hi, const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                      const uint64_t* step_dev, void* stream);

/* Fused single-launch "range from the same tensor, then quantize" (cooperative grid sync):
 * replaces qb_minmax + qb_fakequant_grad for the gradient path (12 B/elem: two reads, one write).
 *   range_out float[2] receives the {min,max} used.  ws as for qb_minmax.                          */
int qb_fakequant_grad_fused(const float* g, float* out, uint8_t* codes, int64_t n, int bits, float* range_out,
                            const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                            const uint64_t* step_dev, void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * RangeBN - quantize.py:461-505.  x is the ALREADY fake-quantized NCHW input [B,C,HW] (2-D inputs:
 * HW = 1).  Chunking follows :468-469: channel c's elements in (b,h,w) order cut into num_chunks
 * runs of L = B*HW/num_chunks (B*HW % num_chunks must be 0, as in the reference).
 *
 * qb_rangebn_stats: per (c,k) max / min and the flat (b*HW+p) index of their FIRST occurrence
 * (ATen's tie-break, SURVEY.md KAT 8), and the per-channel reduction
 *   red double[3*C] = {sum_k max_ck | sum_k min_ck | sum x_c}   (three planes of C)
 * which is what crosses ranks under batch sharding (all-reduce SUM; each rank owns whole chunks).
 *   amax,amin  int32[C*num_chunks] out (may be NULL in eval/no-grad)
 *   ws         float[2*C*num_chunks] + double[C*num_chunks] scratch for the per-chunk values:
 *              qb_rangebn_ws_bytes(C, num_chunks)                                                  */
size_t qb_rangebn_ws_bytes(int64_t C, int num_chunks);
int qb_rangebn_stats(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                     int32_t* amin, double* red, void* ws, size_t ws_bytes, void* stream);

/* Per-channel finalize (:470-498): mean, scale = 1/((mean_max-mean_min)*scale_fix+eps), running
 * buffer update (running <- running*momentum + new*(1-momentum)), then the three 8-bit vector
 * quantizers: scale and gamma with QB_SCALE_DOUBLE over their own min/max, beta with
 * QB_SCALE_FLOAT over its data range.  training == 0: mean/scale come from running_mean/var.
 *   red       double[3*C] from qb_rangebn_stats (summed over ranks); num_chunks is the GLOBAL chunk count
 *   count     global element count per channel (B*HW summed over ranks)
 *   chunk_len L used in scale_fix (global)
 *   outputs   mean[C], scale_raw[C] (unquantized 1/range, needed by backward), qscale[C], qgamma[C], qbeta[C]
 *   gamma/beta may be NULL (affine=False): qgamma/qbeta are then not written.
 * Single CTA; C up to 65536.                                                                      */
int qb_rangebn_finalize(const double* red, int64_t C, int num_chunks,
                        double count, double chunk_len, float eps, double momentum, int training, int bits,
                        const float* gamma, const float* beta, float* running_mean, float* running_var,
                        float* mean, float* scale_raw, float* qscale, float* qgamma, float* qbeta, void* stream);

/* out = ((x - mean_c) * qscale_c) * qgamma_c + qbeta_c, four separately rounded fp32 ops (:488-499).
 * qgamma / qbeta may be NULL.  out may alias x.                                                   */
int qb_rangebn_apply(const float* x, float* out, int64_t B, int64_t C, int64_t HW, const float* mean,
                     const float* qscale, const float* qgamma, const float* qbeta, void* stream);

/* Backward of RangeBN given the (already quantized) output gradient gq - autograd of :467-499:
 *   reduce: sg[c] = sum gq, sgx[c] = sum gq * x   (fp64)
 *   apply : dx = gq*(qgamma_c*qscale_c) + a_c  and  dx[first argmax of chunk] += e_c, dx[first argmin] -= e_c
 * with a_c, e_c produced by qb_rangebn_bwd_coeffs from the reduced sums:
 *   dbeta = sg, dgamma = qscale*(sgx - mean*sg), dscale = qgamma*(sgx - mean*sg),
 *   e = -scale_fix*scale_raw^2*dscale/num_chunks, a = -(qscale*qgamma*sg)/count.                    */
size_t qb_rangebn_bwd_ws_bytes(int64_t C);
int qb_rangebn_bwd_reduce(const float* gq, const float* x, int64_t B, int64_t C, int64_t HW, double* sg,
                          double* sgx, void* ws, size_t ws_bytes, void* stream);
int qb_rangebn_bwd_coeffs(const double* sg, const double* sgx, int64_t C, int num_chunks, double count,
                          double chunk_len, int training, const float* mean, const float* scale_raw,
                          const float* qscale, const float* qgamma, float* dgamma, float* dbeta, float* coef_a,
                          float* coef_e, void* stream);
int qb_rangebn_bwd_apply(const float* gq, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                         const float* qscale, const float* qgamma, const float* coef_a, const float* coef_e,
                         const int32_t* amax, const int32_t* amin, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Channels-last (NHWC) variants of the RangeBN kernels: x is [B, H*W, C] in memory (torch.channels_last),
 * the native activation layout of this library (the int8 implicit GEMM reads and writes it).  Same
 * semantics, outputs and index conventions (amax/amin hold the flat b*HW+p index) as the NCHW entry
 * points above; finalize / bwd_coeffs are layout independent.                                        */
size_t qb_rangebn_nhwc_ws_bytes(int64_t C, int num_chunks);
/* Fusion hooks of the channels-last path (all optional, NULL = plain behaviour):
 *   q_range/q_bits   x is the RAW RangeBN input and the QuantMeasure fake-quant of quantize.py:463 (range float[2] on
 *                    the device, QB_SCALE_DOUBLE) is applied on the fly - x_hat is never materialised;
 *   amax_kc/amin_kc  additional [chunk][channel] copies of the arg tables for the fused backward;
 *   g_range/g_bits/noise/stochastic/seed/offset/step_dev   gq is the RAW output gradient and UniformQuantizeGrad's
 *                    backward (quantize.py:124-139, arguments as qb_fakequant_grad) is applied on the fly; the noise of
 *                    an element is a function of (seed, element index) so both backward passes see the same draw.
 *                    Needs C % 8 == 0 and 32-byte aligned tensors.
 *   x_codes_out / x_codes   (q_bits <= 8, C % 8 == 0) the statistics pass also WRITES x_hat as uint8 codes [B, H*W, C];
 *                    apply and the backward reduce then read those codes (1 B/elem) instead of re-reading and
 *                    re-quantizing the fp32 input (x may be NULL there): the raw input need not be kept for backward. */
int qb_rangebn_stats_nhwc(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax,
                          int32_t* amin, int32_t* amax_kc, int32_t* amin_kc, double* red, const float* q_range,
                          int q_bits, uint8_t* x_codes_out, void* ws, size_t ws_bytes, void* stream);
/* qb_rangebn_stats_nhwc followed by qb_rangebn_finalize(training = 1) (quantize.py:467-498) in one call: where the
 * statistics pass ends in a single-CTA kernel the per-channel finalize runs at its tail (one launch less per layer).  Same
 * results as the two calls.  Single-process use: across ranks `red` is exchanged between the two calls, which then stay apart. */
int qb_rangebn_stats_finalize_nhwc(const float* x, int64_t B, int64_t C, int64_t HW, int num_chunks, int32_t* amax, int32_t* amin,
                                   int32_t* amax_kc, int32_t* amin_kc, double* red, const float* q_range, int q_bits,
                                   uint8_t* x_codes_out, float eps, double momentum, int bits, const float* gamma,
                                   const float* beta, float* running_mean, float* running_var, float* mean, float* scale_raw,
                                   float* qscale, float* qgamma, float* qbeta, void* ws, size_t ws_bytes, void* stream);
int qb_rangebn_apply_nhwc(const float* x, const uint8_t* x_codes, float* out, int64_t B, int64_t C, int64_t HW,
                          const float* mean, const float* qscale, const float* qgamma, const float* qbeta,
                          const float* q_range, int q_bits, void* stream);
size_t qb_rangebn_bwd_nhwc_ws_bytes(int64_t C);
int qb_rangebn_bwd_reduce_nhwc(const float* gq, const float* x, const uint8_t* x_codes, int64_t B, int64_t C, int64_t HW,
                               double* sg, double* sgx, const float* q_range, int q_bits, const float* g_range, int g_bits,
                               const float* noise, int stochastic, uint64_t seed, uint64_t offset,
                               const uint64_t* step_dev, uint8_t* gq_codes_out, const uint8_t* relu_bits, void* ws,
                               size_t ws_bytes, void* stream);
/* qb_rangebn_bwd_reduce_nhwc followed by qb_rangebn_bwd_coeffs in one call (autograd of quantize.py:467-499): the warp that
 * folds a channel's partial sums derives its coefficients on the spot (one launch less per layer; same arithmetic).
 * Single-process use: across ranks the sums are exchanged between the two calls. */
int qb_rangebn_bwd_reduce_coeffs_nhwc(const float* gq, const float* x, const uint8_t* x_codes, int64_t B, int64_t C, int64_t HW,
                                      double* sg, double* sgx, const float* q_range, int q_bits, const float* g_range, int g_bits,
                                      const float* noise, int stochastic, uint64_t seed, uint64_t offset, const uint64_t* step_dev,
                                      uint8_t* gq_codes_out, const uint8_t* relu_bits, int num_chunks, double count,
                                      double chunk_len, int training, const float* mean, const float* scale_raw,
                                      const float* qscale, const float* qgamma, float* dgamma, float* dbeta, float* coef_a,
                                      float* coef_e, void* ws, size_t ws_bytes, void* stream);
/* relu_bits (optional): gq is the gradient w.r.t. the OUTPUT of the nn.ReLU that follows this RangeBN and relu_bits the
 * bit mask qb_rangebn_apply_relu_measure wrote; the ReLU backward (threshold_backward) then happens on the fly and the
 * masked gradient is never written.  g_range must be the range of the MASKED gradient (qb_relu_bwd_minmax_bits, dx = NULL). */
/* dx = (Q(g)*qgamma)*qscale + a, +-e at the chunk arg-max/arg-min rows, in ONE pass, plus min/max of the final dx
 * (gq_codes: optional uint8 codes of Q(g) written by qb_rangebn_bwd_reduce_nhwc's gq_codes_out - the pass then reads
 * 1 B/elem and only dequantizes instead of reading g and regenerating the noise;
 * dx_range_out float[2], optional; mm_ws = qb_minmax_ws_bytes() bytes zero-filled once): the range the gradient
 * quantizer of the preceding QConv2d needs (quantize.py:126,132) without another read of dx. */
int qb_rangebn_bwd_apply_fused_nhwc(const float* g, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                                    const float* qscale, const float* qgamma, const float* coef_a,
                                    const float* coef_e, const int32_t* amax_kc, const int32_t* amin_kc,
                                    const float* g_range, int g_bits, const float* noise, int stochastic,
                                    uint64_t seed, uint64_t offset, const uint64_t* step_dev, const uint8_t* gq_codes,
                                    float* dx_range_out, void* mm_ws, size_t mm_ws_bytes, void* stream);
int qb_rangebn_bwd_apply_nhwc(const float* gq, float* dx, int64_t B, int64_t C, int64_t HW, int num_chunks,
                              const float* qscale, const float* qgamma, const float* coef_a, const float* coef_e,
                              const int32_t* amax, const int32_t* amin, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Contraction - F.linear / F.conv2d on the dequantised tensors (quantize.py:143-156, 343, 424) as an int8
 * tcgen05 GEMM on the integer codes with the asymmetric dequantisation folded into the epilogue:
 *   C[m,n] = sa*sw[n]*SUM_k A[m,k]B[n,k] + sa*lw[n]*a_rowsum[m] + la*sw[n]*b_rowsum[n] + K*la*lw[n] + bias[n]
 *   A uint8 [M,K] codes of the activation quantizer with device range a_range {la, max} (QB_SCALE_DOUBLE)
 *   B uint8 [N,K] codes of the per-row weight quantizer with ranges b_row_min/max[N] (QB_SCALE_FLOAT)
 *   a_rowsum int32[M], b_rowsum int32[N]: row sums of the codes (qb_rowsum_u8)
 *   bias: fp32 [N] (already fake-quantized) or NULL;  C fp32 [M, ldc]
 *   b_offset: 0 if B holds the raw unsigned codes; 2^(b_bits-1) if B holds the CENTRED signed codes q - 2^(b_bits-1)
 *             that qb_fakequant_rows_conv emits (b_rowsum then sums the centred codes) - better conditioned
 * K must be a multiple of 16 and A, B 16-byte aligned (TMA).                                          */
int qb_rowsum_u8(const uint8_t* codes, int64_t rows, int64_t cols, int32_t* out, void* stream);
int qb_gemm_i8(const uint8_t* A, const uint8_t* B, float* C, int64_t M, int64_t N, int64_t K, int64_t ldc,
               const float* a_range, int a_bits, const float* b_row_min, const float* b_row_max, int b_bits,
               const int32_t* a_rowsum, const int32_t* b_rowsum, const float* bias, int b_offset, void* stream);

/* Weight preparation for the implicit GEMM: per-output-channel range + fake-quant of a [Cout, Cin, kh*kw] fp32 weight
 * (quantize.py:318-334) with, besides the fp32 result wq (same layout; may be NULL), the uint8 codes re-ordered to
 * [Cout][kh*kw][Cin] (K-major, channel innermost like the NHWC activations), CENTRED: int8 q - 2^(bits-1), and
 * tapsum[Cout][kh*kw] = SUM_c of the centred codes.  With kh*kw == 1 this is the [N,K] operand of qb_gemm_i8.          */
int qb_fakequant_rows_conv(const float* w, float* wq, uint8_t* codes_khwc, int32_t* tapsum, int64_t Cout, int64_t Cin,
                           int taps, int bits, float* row_min, float* row_max, int compute_range, void* stream);

/* QConv2d forward contraction (quantize.py:143-148, 343): int8 tcgen05 implicit GEMM over channels-last codes.
 *   x_codes uint8 [B,H,W,Cin] (NHWC), a_range device {min,max} of its quantizer; pixsum int32[B*H*W] = SUM_c codes
 *   w_codes_khwc uint8 [Cout][kh*kw][Cin], per-row ranges, tapsum (from qb_fakequant_rows_conv); bias fp32[Cout] or NULL
 *   boxsum_ws int32[B*Ho*Wo] scratch: receives the per-output-pixel sums of pixsum over the in-bounds taps
 *   out_stats (optional) double[8]: the QuantMeasure statistics block of the OUTPUT (same layout as qb_measure_stats,
 *   stats_ws = qb_measure_ws_bytes(B) bytes zero-filled once): the range reduction of the layer that follows
 *   (quantize.py:223-235) fused into the epilogue, so the output is not read again for it
 *   out fp32 [B,Ho,Wo,Cout] (NHWC), written by TMA stores.  Zero padding is exact for any input range (border-aware
 *   epilogue).
 * Supported: Cin % 32 == 0, Cout % 16 == 0, kh*kw <= 9, stride 1 or 2, dilation 1, groups 1; anything else returns
 * QB_ERR_UNSUPPORTED (the caller must then use another contraction; nothing is computed).                            */
int qb_conv_i8_fwd(const uint8_t* x_codes, const uint8_t* w_codes_khwc, float* out, int64_t B, int64_t H, int64_t W,
                   int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                   const float* w_row_min, const float* w_row_max, int w_bits, const int32_t* pixsum,
                   int32_t* boxsum_ws, const int32_t* tapsum, const float* bias, double* out_stats, void* stats_ws,
                   size_t stats_ws_bytes, void* stream);

/* QConv2d forward contraction of a network's FIRST layer (models/resnet_quantized.py:147 conv1 7x7/2 over 3 input
 * channels; models/mobilenet_quantized.py first 3x3/2 conv): tcgen05 kind::i8 over a thread-built im2col operand, same
 * operands as qb_conv_i8_fwd (x_codes NHWC uint8, centred weight codes [Cout][kh*kw][Cin] + tapsum from
 * qb_fakequant_rows_conv, device-side ranges).  No pixsum / boxsum scratch: the activation code sum of each window is
 * taken in the kernel; zero padding is exact (per-border-case constants).
 * Supported (qb_conv_stem_supported): 7x7 filters with 21..24-byte rows (kw*Cin) or 3x3 with 9..12-byte rows,
 * Cout % 16 == 0 and <= 128, stride 1 or 2, dilation 1, groups 1, at most 8 distinct border cases per axis; otherwise
 * QB_ERR_UNSUPPORTED.  ws: qb_conv_stem_ws_bytes(Cout) bytes of scratch, 16-byte aligned.  out fp32 [B,Ho,Wo,Cout].   */
int qb_conv_stem_supported(int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad);
size_t qb_conv_stem_ws_bytes(int64_t Cout);
int qb_conv_stem_fwd(const uint8_t* x_codes, const int8_t* w_codes_khwc, float* out, int64_t B, int64_t H, int64_t W,
                     int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                     const float* w_row_min, const float* w_row_max, int w_bits, const int32_t* tapsum,
                     const float* bias, void* ws, size_t ws_bytes, void* stream);

/* QConv2d forward contraction of a DEPTHWISE 3x3 convolution (groups == C_in == C_out; models/mobilenet_quantized.py:21-60
 * through quantize.py:143-148, 343) on the integer codes.  18 operations per 5 bytes of traffic: a streaming kernel
 * bounded by HBM (no tensor cores), one thread per output pixel x 4 channels, IDP.4A products, packed code sums.
 *   x_codes uint8 [B,H,W,C] (NHWC) with device range a_range {min,max}; w_codes int8 [C][9] CENTRED codes (what
 *   qb_fakequant_rows_conv emits for a [C,1,3,3] weight) with per-channel ranges; bias fp32[C] or NULL
 *   out fp32 [B,Ho,Wo,C] (NHWC); zero padding exact for any input range
 *   out_stats / stats_ws: as in qb_conv_i8_fwd (QuantMeasure statistics of the output fused into the kernel)
 * Supported (qb_conv_dw_supported): 3x3, C % 4 == 0, stride 1 or 2, pad 0..2, dilation 1; else QB_ERR_UNSUPPORTED.     */
int qb_conv_dw_supported(int64_t C, int kh, int kw, int stride, int pad);
int qb_conv_dw_fwd(const uint8_t* x_codes, const int8_t* w_codes, float* out, int64_t B, int64_t H, int64_t W, int64_t C,
                   int kh, int kw, int stride, int pad, const float* a_range, int a_bits, const float* w_row_min,
                   const float* w_row_max, int w_bits, const float* bias, double* out_stats, void* stats_ws,
                   size_t stats_ws_bytes, void* stream);

/* Backward of the depthwise 3x3 convolution under conv2d_biprec (quantize.py:142-148; aten::convolution_backward with
 * groups == C in the reference): fp32 streaming kernels over channels-last tensors, same geometry limits as
 * qb_conv_dw_fwd.
 *   qb_conv_dw_dgrad: dx[B,H,W,C] = transposed depthwise convolution of grad_q [B,Ho,Wo,C] (the gradient the caller has
 *     already quantized to num_bits_grad) with the fake-quantized weights w_hat fp32 [C][9].
 *   qb_conv_dw_wgrad: dw[C][9] = SUM over output pixels of x_hat (zero outside the image) * grad_out, and db[C] = SUM
 *     grad_out (db may be NULL); deterministic (fixed-order partial blocks in ws: qb_conv_dw_wgrad_ws_bytes() bytes).
 *     Needs C/4 to divide 256.                                                                                        */
int qb_conv_dw_dgrad(const float* grad_q, const float* w_hat, float* dx, int64_t B, int64_t H, int64_t W, int64_t C, int kh,
                     int kw, int stride, int pad, void* stream);
size_t qb_conv_dw_wgrad_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t C, int stride, int pad);
int qb_conv_dw_wgrad(const float* grad_out, const float* x_hat, float* dw, float* db, int64_t B, int64_t H, int64_t W,
                     int64_t C, int kh, int kw, int stride, int pad, void* ws, size_t ws_bytes, void* stream);

/* Weight gradient of that first layer (quantize.py:143-148: the weight branch of the bi-precision backward sees the
 * unquantized gradient): dW = x_hat^T * g over all output pixels, from the input CODES (x_hat = sa*q + la inside the
 * image, 0 in the padding) and the fp32 NHWC gradient, on tf32 tensor cores (g rounded to tf32, codes exact),
 * accumulated per CTA in TMEM and reduced in a fixed order.  C_out 32 or 64.  dw fp32 [C_out][C_in][kh][kw].
 * ws: qb_conv_stem_wgrad_ws_bytes() bytes, 16-byte aligned.                                                           */
size_t qb_conv_stem_wgrad_ws_bytes(int64_t Cout, int64_t Cin, int kh, int kw);
int qb_conv_stem_wgrad(const float* grad_out, const uint8_t* x_codes, float* dw, int64_t B, int64_t H, int64_t W,
                       int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* a_range, int a_bits,
                       void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * quantize2.py flavour: QParams (range, zero_point) are device tensors, one per tensor or one per
 * row of dim 0; no 1e-8 floor on the scale; optional signed codes, optional raw-code output, stochastic rounding on
 * the same branch.  models/modules/quantize2.py:41-76 (UniformQuantize.forward).
 *   x, y       fp32 [rows, cols] (y may alias x); zero_point, range: fp32[per_row ? rows : 1] on the device
 *   dequantize 0 -> y holds the integer codes as fp32 (quantize2.py:72 skipped)
 *   noise / stochastic / seed / offset / step_dev as in qb_fakequant_grad                                           */
int qb_fakequant2(const float* x, float* y, int64_t rows, int64_t cols, const float* zero_point, const float* range,
                  int per_row, int bits, int is_signed, int dequantize, const float* noise, int stochastic,
                  uint64_t seed, uint64_t offset, const uint64_t* step_dev, void* stream);
/* Per-tensor, unsigned, deterministic form of the same quantizer that also emits the integer codes as uint8 and bf16 (bits <= 8):
 * feeds quantize2.QConv2d / QLinear (quantize2.py:204-272) to qb_conv_i8_fwd / qb_gemm_i8 / qb_conv_wgrad.  y optional. */
int qb_fakequant2_codes(const float* x, float* y, uint8_t* codes, uint16_t* codes16, int64_t n, const float* zero_point,
                        const float* range, int bits, void* stream);
/* calculate_qparams with flatten_dims=(1,-1) (quantize2.py:19-28): min / max of every row of x [rows, cols].         */
int qb_row_minmax(const float* x, int64_t rows, int64_t cols, float* row_min, float* row_max, void* stream);
/* quantize2.RangeBN (quantize2.py:275-349): red = double[3C] from qb_rangebn_stats[_nhwc]; scale is neither inverted
 * nor quantized, running_var tracks the scale, the normalisation divides by den = scale + eps.  rinv = 1/den feeds the
 * backward kernels (qb_rangebn_bwd_*: pass rinv as scale_raw and qscale, gamma as qgamma).                            */
int qb_rangebn2_finalize(const double* red, int64_t C, int num_chunks, double count, double chunk_len, float eps,
                         double momentum, int training, float* running_mean, float* running_var, float* mean,
                         float* scale, float* den, float* rinv, void* stream);
int qb_rangebn2_apply(const float* x, float* out, int64_t B, int64_t C, int64_t HW, int nhwc, const float* mean,
                      const float* den, const float* gamma, const float* beta, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Glue between the quantized modules: nn.MaxPool2d over channels-last fp32
 * (models/resnet_quantized.py:150).  Same values, tie-breaking (first maximal element of the window, rows first) and
 * backward accumulation order as ATen's max_pool2d; idx holds the window-relative position kh*k + kw of the maximum as
 * one byte per output element.  Needs C % 8 == 0, k <= 15, 2*pad <= k, 32-byte aligned x/out/grad/dx, 8-byte aligned idx. */
/* nn.ReLU(inplace) between the modules (models/resnet_quantized.py:52-68,147): y = relu(x) (y may alias x) and, in the same
 * pass, the QuantMeasure statistics block (as qb_measure_stats) of y for the quantized layer that consumes it. */
int qb_relu_measure(const float* x, float* y, int64_t B, int64_t inner, double* stats, void* ws, size_t ws_bytes, void* stream);
/* Backward of that ReLU with the range of its result: dx = relu_out <= 0 ? 0 : grad_out (aten::threshold_backward) and
 * range_out float[2] = {min(dx), max(dx)}, which the gradient quantizer of the layer that receives dx asks for next
 * (quantize.py:126,132).  dx may alias grad_out.  ws = qb_minmax_ws_bytes() bytes zero-filled once. */
int qb_relu_bwd_minmax(const float* grad_out, const float* relu_out, float* dx, int64_t n, float* range_out, void* ws,
                       size_t ws_bytes, void* stream);
/* The residual add and ReLU that close a ResNet block (`out += residual; out = relu(out)`, models/resnet_quantized.py:62-68,
 * 105-112) in one pass: y = relu(x + r) (y may alias x, r must not alias y; x + r rounded to fp32 before the clamp, as the
 * two torch ops do) plus the statistics block of y. */
int qb_add_relu_measure(const float* x, const float* r, float* y, int64_t B, int64_t inner, double* stats, void* ws,
                        size_t ws_bytes, void* stream);
/* qb_relu_bwd_minmax with the ReLU output given as a bit mask (bit e%8 of byte e/8 set = the gradient passes; written by
 * qb_rangebn_apply_relu_measure): 8.125 B/elem instead of 12.  dx == NULL: only range_out = {min, max} of the masked
 * gradient (4.125 B/elem), for a consumer that applies the mask itself (qb_rangebn_bwd_reduce_nhwc's relu_bits).
 * n % 8 == 0, 32-byte aligned tensors. */
int qb_relu_bwd_minmax_bits(const float* grad_out, const uint8_t* relu_bits, float* dx, int64_t n, float* range_out, void* ws,
                            size_t ws_bytes, void* stream);
/* RangeBN normalisation (models/modules/quantize.py:488-499) + the nn.ReLU that follows it in every block of
 * models/resnet_quantized.py:47-75 / models/mobilenet_quantized.py:21-60 [+ the residual add in front of that ReLU, residual
 * != NULL] + the statistics block of the result, in ONE pass over the uint8 codes of RangeBN's fake-quantized input that
 * qb_rangebn_stats_nhwc wrote:  y = relu(((deq(code) - mean_c) * qscale_c) * qgamma_c + qbeta_c [+ residual]).
 * Bit-identical to qb_rangebn_apply_nhwc followed by qb_relu_measure / qb_add_relu_measure (values and statistics); moves
 * 5 (+4) B/elem instead of 13 (+4): the normalised tensor is never written.  Channels-last, C % 4 == 0, q_bits <= 8,
 * 16-byte aligned fp32 tensors; qgamma / qbeta may be NULL; ws = qb_measure_ws_bytes(B).  relu_bits_out (optional, H*W*C % 8
 * == 0): bit e%8 of byte e/8 = "y[e] is not <= 0" - all the ReLU backward needs, so y need not be kept for it. */
int qb_rangebn_apply_relu_measure(const uint8_t* x_codes, const float* residual, float* y, int64_t B, int64_t C, int64_t HW,
                                  const float* mean, const float* qscale, const float* qgamma, const float* qbeta,
                                  const float* q_range, int q_bits, uint8_t* relu_bits_out, double* stats, void* ws,
                                  size_t ws_bytes, void* stream);
int qb_maxpool2d_nhwc_fwd(const float* x, float* out, uint8_t* idx, int64_t B, int64_t H, int64_t W, int64_t C, int k,
                          int stride, int pad, void* stream);
/* maxpool(RangeBN-normalise(x)) from the uint8 codes of RangeBN's fake-quantized input (x_codes_out of
 * qb_rangebn_stats_nhwc) and the vectors of qb_rangebn_finalize: the stem tail conv1 -> bn1 -> relu -> maxpool
 * (models/resnet_quantized.py:147-150, quantize.py:488-499) without ever writing the normalised tensor.  out / idx equal
 * qb_maxpool2d_nhwc_fwd on the output of qb_rangebn_apply_nhwc, bit for bit. */
int qb_rangebn_apply_maxpool_nhwc(const uint8_t* x_codes, float* out, uint8_t* idx, int64_t B, int64_t H, int64_t W, int64_t C, int k,
                                  int stride, int pad, const float* mean, const float* qscale, const float* qgamma,
                                  const float* qbeta, const float* q_range, int q_bits, void* stream);
int qb_maxpool2d_nhwc_bwd(const float* grad_out, const uint8_t* idx, float* dx, int64_t B, int64_t H, int64_t W, int64_t C,
                          int k, int stride, int pad, void* stream);
/* Backward of ReLU -> MaxPool2d as one pass (models/resnet_quantized.py:147-150, bn1 -> relu -> maxpool): dx is the gradient
 * w.r.t. the ReLU's input.  pool_out = the forward output of the pooling (its window maxima): a window whose maximum is not
 * positive contributes nothing, which is what aten::threshold_backward after max_pool2d_backward yields.  range_out
 * float[2] = {min(dx), max(dx)} for the gradient quantizer that consumes dx.  ws = qb_maxpool2d_relu_bwd_ws_bytes() bytes
 * zero-filled once. */
size_t qb_maxpool2d_relu_bwd_ws_bytes(void);
int qb_maxpool2d_relu_nhwc_bwd(const float* grad_out, const uint8_t* idx, const float* pool_out, float* dx, int64_t B, int64_t H,
                               int64_t W, int64_t C, int k, int stride, int pad, float* range_out, void* ws, size_t ws_bytes,
                               void* stream);

/* ---- the tail of the training step (glue_loss.cu) ----------------------------------------------------------------------
 * nn.AvgPool2d(k) with stride k, no padding (models/resnet_quantized.py:153,189; models/mobilenet_quantized.py:114,124) on
 * channels-last fp32 x [B, H, W, C], C % 4 == 0: out [B, H/k, W/k, C]; window sums in raster order, one IEEE division. */
int qb_avgpool2d_nhwc_fwd(const float* x, float* out, int64_t B, int64_t H, int64_t W, int64_t C, int k, void* stream);
int qb_avgpool2d_nhwc_bwd(const float* grad_out, float* dx, int64_t B, int64_t H, int64_t W, int64_t C, int k, void* stream);
/* nn.CrossEntropyLoss() (main.py:362-380: mean reduction, no class weights, no label smoothing) on fp32 logits [B, C] and
 * int64 targets [B].  fwd: loss (device float), n_valid (device float: rows whose target != ignore_index), row_state
 * float[2*B] = (row max, log-sum-exp) kept for bwd; ws = qb_cross_entropy_ws_bytes(B).  bwd: dlogits [B, C] =
 * (softmax - onehot) * grad_out / n_valid, rebuilt from row_state (no [B, C] log-softmax intermediate). */
size_t qb_cross_entropy_ws_bytes(int64_t B);
int qb_cross_entropy_fwd(const float* logits, const int64_t* target, int64_t B, int64_t C, int64_t ignore_index, float* loss,
                         float* row_state, float* n_valid, void* ws, size_t ws_bytes, void* stream);
int qb_cross_entropy_bwd(const float* logits, const int64_t* target, const float* row_state, const float* n_valid,
                         const float* grad_out, float* dlogits, int64_t B, int64_t C, int64_t ignore_index, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Dense backward contractions of QConv2d / QLinear (conv_bwd.cu) - replace aten::convolution_backward / matmul under
 * conv2d_biprec / linear_biprec (models/modules/quantize.py:142-156) and the non-biprec branch (:342-346).
 * tcgen05 kind::f16 (bf16 operands, fp32 accumulation) fed by TMA.  One operand of each contraction is an integer code,
 * exact in bf16; the other is split into two bf16 planes hi + lo (relative operand error 2^-17).
 *
 * qb_bwd_pack: ONE pass over the fp32 output gradient g (n elements, any dense layout):
 *   gq16 (optional)      bf16 integers (q - zp) of UniformQuantizeGrad.backward's quantizer (:124-139: zero-point,
 *                        stochastic; range_dev = device {min,max}, bits <= 8; noise as in qb_fakequant_grad)
 *   g_hi, g_lo (optional, together)   bf16 split of the RAW g
 * qb_wt_pack: w_hat fp32 [C_out][C_in][taps] (fake-quantized weight) -> wt bf16 [2][C_in][taps][C_out] (hi, lo planes).
 * qb_conv_dgrad: dx fp32 [B,H,W,C_in] (channels-last) = scale * SUM_{tap,n} a[b,ho,wo,n] * w_hat[n,c,tap];  a = a0 (+ a1):
 *   bf16 planes [B,Ho,Wo,C_out]; scale = scale_mul * (scale_range ? step of the zero-point quantizer over scale_range at
 *   scale_bits : 1).  Stride s is evaluated as s*s unit-stride parity classes of dx.  Linear: H = W = kh = kw = 1.
 * qb_conv_wgrad: dw fp32 [C_out][C_in][kh][kw] = sa * SUM_pixels g * q_a + la * SUM_{pixels whose tap is in bounds} g, and
 *   db = SUM_pixels g, from the bf16 planes of g, the fp32 g (column sums only) and x16 = bf16 integer codes q_a
 *   [B,H,W,C_in] of the input quantizer (device range a_range, a_bits).  Split over the pixels, fixed-order reduction
 *   (deterministic).  ws: qb_conv_wgrad_ws_bytes() bytes, 16-byte aligned.
 * Unsupported geometries return QB_ERR_UNSUPPORTED (there is no library fallback). */
int qb_bwd_pack(const float* g, int64_t n, uint16_t* gq16, uint16_t* g_hi, uint16_t* g_lo, int bits, const float* range_dev,
                const float* noise, int stochastic, uint64_t seed, uint64_t offset, const uint64_t* step_dev, void* stream);
int qb_wt_pack(const float* w_hat, uint16_t* wt, int64_t Cout, int64_t Cin, int taps, void* stream);
int qb_conv_dgrad(const uint16_t* a0, const uint16_t* a1, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W,
                  int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad, const float* scale_range, int scale_bits,
                  float scale_mul, void* stream);
/* qb_conv_dgrad that ACCUMULATES: dx += dX.  dx already holds the other gradient that reaches the layer's input - the skip
 * connection of a residual block (`out += residual`, models/resnet_quantized.py:62-68) - so autograd's separate accumulation
 * pass (12 B/elem) becomes TMA reduce-add stores.  One gradient plane (integer codes); every pixel parity class of the stride
 * must be reached by the filter and taken by the flat-strip kernel (checked first); otherwise QB_ERR_UNSUPPORTED and dx is left
 * untouched.  With a downsample branch (models/resnet_quantized.py:58-60) dx holds that branch's input gradient. */
int qb_conv_dgrad_acc(const uint16_t* a0, const uint16_t* wt, float* dx, int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout,
                      int kh, int kw, int stride, int pad, const float* scale_range, int scale_bits, float scale_mul, void* stream);
/* Weight gradient of the FIRST layer (3 input channels; 7x7/stride 2/pad 3 or 3x3/stride 2/pad 1, even W) on the qb_conv_wgrad
 * kernel: the uint8 input codes [B,H,W,3] are expanded to zero-padded 4-channel bf16 and read through an OVERLAPPING tensor
 * map as a virtual operand whose "channels" are a filter row's window (conv_bwd.cu).  g_hi / g_lo: bf16 planes of the output
 * gradient [B,Ho,Wo,C_out]; dw fp32 [C_out][3][k][k] = sa * SUM g*q + la * SUM_{tap in bounds} g.  ws_bytes() == 0: geometry
 * not taken (qb_conv_stem_wgrad remains). */
size_t qb_conv_stem_wgrad16_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t Cout, int k, int stride, int pad);
int qb_conv_stem_wgrad16(const uint16_t* g_hi, const uint16_t* g_lo, const uint8_t* x_codes, float* dw, int64_t B, int64_t H, int64_t W,
                         int64_t Cout, int k, int stride, int pad, const float* a_range, int a_bits, void* ws, size_t ws_bytes,
                         void* stream);

/* Direct convolution / linear kernels for every geometry the tensor-core kernels do not take (conv_generic.cu): the
 * reference's F.conv2d(x_hat, w_hat, b_hat) / F.linear and their gradients (models/modules/quantize.py:142-156, 343, 424) on
 * the fake-quantized fp32 tensors, CUDA cores, fp64 accumulation, deterministic.  They keep EVERY quantized layer off cuDNN /
 * cuBLAS.  conv = {kh, kw, stride_h, stride_w, pad_h, pad_w, dil_h, dil_w, groups}; x_strides / o_strides = element strides
 * (n, c, h, w) of x (dx) and out (grad_out); w [C_out][C_in/groups][kh][kw].  Linear: H = W = kh = kw = 1. */
int qb_conv_generic_fwd(const float* x, const float* w, const float* bias, float* out, int64_t B, int64_t Cin, int64_t H, int64_t W,
                        int64_t Cout, const int* conv, const int64_t* x_strides, const int64_t* o_strides, void* stream);
int qb_conv_generic_dgrad(const float* grad_out, const float* w, float* dx, int64_t B, int64_t Cin, int64_t H, int64_t W, int64_t Cout,
                          const int* conv, const int64_t* x_strides, const int64_t* o_strides, void* stream);
int qb_conv_generic_wgrad(const float* grad_out, const float* x, float* dw, float* db, int64_t B, int64_t Cin, int64_t H, int64_t W,
                          int64_t Cout, const int* conv, const int64_t* x_strides, const int64_t* o_strides, void* stream);
/* Optimizer / regulariser glue of the training step as multi-tensor kernels (glue_opt.cu; SURVEY.md section 8f N1;
 * main.py:362-380, models/resnet_quantized.py:157-165).  Up to 96 fp32 tensors per call, passed as arrays of pointers and
 * element counts.
 *   qb_l2_sumsq: out[0] (= or +=) 0.5 * wd * SUM_t SUM_i w_t[i]^2 (fp64 accumulation, fixed order); ws: qb_mt_ws_bytes().
 *   qb_l2_grad : grad_t = (upstream[0] * wd) * w_t (upstream NULL = 1).
 *   qb_sgd_step: torch.optim.SGD step (dampening 0, no Nesterov): g' = g + wd*p; buf = first ? g' : buf*momentum + g';
 *                p -= lr*buf, every operation rounded as torch's kernels round it. */
size_t qb_mt_ws_bytes(int count, const int64_t* sizes);
int qb_l2_sumsq(const void* const* w, const int64_t* sizes, int count, double weight_decay, float* out, int accumulate, void* ws,
                size_t ws_bytes, void* stream);
int qb_l2_grad(const void* const* w, void* const* grad, const int64_t* sizes, int count, double weight_decay, const float* upstream,
               void* stream);
int qb_sgd_step(const void* const* grad, void* const* param, void* const* buf, const int64_t* sizes, int count, double lr,
                double momentum, double weight_decay, int first, void* stream);
/* qb_l2_sumsq without a workspace and without ANY allocation (no cudaMalloc / cudaFree / allocating library call), for
 * callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.  This entry
 * point has no counterpart in the reference.  The per-CTA partial sums
 * live in a static device array inside the library (64 KB per GPU that loads it, held until exit), which limits one call to
 * SUM_t ceil(sizes[t] / 8192) <= qb_l2_sumsq_noalloc_max_blocks() (= 8192, i.e. up to 2^26 elements); larger inputs are
 * rejected with QB_ERR_ARG and a message in qb_last_error() - use qb_l2_sumsq with a caller workspace for those.  Result is
 * bit-identical to qb_l2_sumsq.  All calls on a device share the array (it keeps the previous call's partial sums; none of
 * it needs to start at zero because every slot that is read was written by the same call): calls that OVERLAP on the GPU
 * (two streams of one device) race and must be serialised by the caller.  qb_l2_sumsq_noalloc_setup() resolves the array's
 * device address for the current device (cudaGetSymbolAddress, once per device); the first call does this itself, a caller
 * that captures graphs calls it once per device before the capture. */
int64_t qb_l2_sumsq_noalloc_max_blocks(void);
int qb_l2_sumsq_noalloc_setup(void);
int qb_l2_sumsq_noalloc(const void* const* w, const int64_t* sizes, int count, double weight_decay, float* out, int accumulate,
                        void* stream);
/* Hardware probe used by tools/umma_shift_probe.py (UMMA descriptors that start inside an 8-row swizzle atom); not on any
 * product path. */
int qb_debug_umma_shift(const uint16_t* a, const uint16_t* b, float* out, int mn, int r0, int use_boff, void* stream);
/* Tensor-pipe rate probe (tools/umma_rate_probe.py): `blocks` CTAs each issue n back-to-back kind::f16 MMAs (M 128, N, K 16) on
 * zero-filled shared memory and report elapsed cycles; mode bits: 1 A MN-major, 2 B MN-major, 4 walk a K window, 8 A starts
 * inside a swizzle atom.  Not on any product path. */
int qb_debug_umma_rate(long long* cycles, int blocks, int n, int N, int mode, void* stream);
size_t qb_conv_wgrad_ws_bytes(int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad);
int qb_conv_wgrad(const uint16_t* g_hi, const uint16_t* g_lo, const float* g_f32, const uint16_t* x16, float* dw, float* db,
                  int64_t B, int64_t H, int64_t W, int64_t Cin, int64_t Cout, int kh, int kw, int stride, int pad,
                  const float* a_range, int a_bits, void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------------
 * Cross-rank exchange of the statistic blocks of batch-sharded training: a one-shot all-reduce
 * over NVLink peer memory for latency-bound messages.  The reference has no counterpart (nn.DataParallel keeps
 * per-replica statistics, main.py:344-345).  Every rank passes the peer-mapped addresses of all ranks' exchange buffers
 * (qb_allreduce_small_buffer_bytes(world, slot_bytes) bytes each, zero-filled once); data is reduced in place, in rank
 * order (bit-identical on every rank).  dtype 0 = double, 1 = float; mode 0 = SUM, 1 = QuantMeasure block (element 4 MIN,
 * 5 MAX, others SUM), 2 = {MIN, MAX} pair.  CUDA-graph safe (the call counter lives in the buffer). */
size_t qb_allreduce_small_buffer_bytes(int world, size_t slot_bytes);
int qb_allreduce_small(void* data, int64_t n, int dtype, int mode, const uint64_t* peer_bufs, int rank, int world,
                       size_t slot_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* QB200_H */
