/*
 * nanofed_b200 — C ABI of the sm_100a kernel library (libnanofed_b200.so).
 *
 * Conventions
 *   - every function returns 0 on success, a negative NFB_E* code on a rejected
 *     argument, or a positive cudaError_t if the launch failed; nfb_last_error()
 *     returns a thread-local, human-readable reason for the last non-zero return.
 *   - pointers are DEVICE pointers unless the name says host; `stream` is a
 *     cudaStream_t passed as void* (0 = legacy default stream).  Nothing here
 *     synchronises the device.
 *   - dtype codes: 0 = float32, 1 = bfloat16, 2 = float16.
 *   - unless noted the library allocates nothing: outputs and scratch are
 *     caller-provided.  Scratch buffers must be zero-filled ONCE by whoever
 *     allocates them; kernels leave them zeroed again on exit ("self-cleaning"),
 *     so one scratch buffer can be reused by consecutive launches on one stream.
 *     Two launches that overlap on the GPU must not share a scratch buffer.
 *
 * Reference parity (camille-004/nanofed, paths relative to its root) is cited per
 * entry point; the reference itself has no native code, each
 * entry point replaces a span of Python/torch-library execution.
 */
#ifndef NANOFED_B200_H_
#define NANOFED_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NFB_DT_F32 0
#define NFB_DT_BF16 1
#define NFB_DT_F16 2

#define NFB_EINVAL (-1)   /* bad argument (null pointer, bad dtype, misalignment) */
#define NFB_ERANGE (-2)   /* size beyond what the entry point supports          */
#define NFB_ESTATE (-3)   /* wrong device / library state                       */

int nfb_version(void);
const char* nfb_last_error(void);

/* ---- K-RNG -----------------------------------------------------------------
 * Counter-based Philox4x32-10 stream (see nanofed_b200/ops/philox.py for the
 * exact stream definition; host and device produce the same integers).
 * kind: 0 = normal(0, scale^2) via Box-Muller, 1 = Laplace(0, scale), 2 = uniform(0,1).
 * Replaces nanofed/privacy/noise/generators.py:49-67 (CPU torch.Generator).
 */
int nfb_philox_fill(void* out, int out_dtype, int64_t n, int kind, float scale,
                    uint64_t seed, uint64_t offset, void* stream);

/* x[i] += m, cumulative[i] += m with m ~ U(0,1) from the stream: the masking
 * arithmetic of nanofed/server/aggregator/secure.py:230-237.                       */
int nfb_mask_add(float* x, float* cumulative, int64_t n, uint64_t seed, uint64_t offset,
                 void* stream);

/* ---- K-STATS ---------------------------------------------------------------
 * Deterministic two-stage reductions (fixed summation order, no float atomics).
 * nfb_stats_scratch_bytes(rows) is the scratch size needed for `rows` reduction
 * rows (K for update_stats, num_tensors for multi_sumsq, K*S for segment_sumsq).
 */
size_t nfb_stats_scratch_bytes(int64_t rows);

/* sumsq[k] = sum_j x[k,j]^2 ; nonfinite[k] = any(!isfinite(x[k,:])).
 * Replaces the concatenated-copy norm of nanofed/privacy/mechanisms.py:89-91 and the
 * finite/norm checks of nanofed/server/validation.py:88-95,111-127.                */
int nfb_update_stats(const void* x, int x_dtype, int64_t K, int64_t N, int64_t row_stride,
                     float* sumsq, int* nonfinite, void* scratch, size_t scratch_bytes,
                     void* stream);

/* Same result as nfb_update_stats for callers that cannot allocate device memory (plugins
 * that only receive buffers, fixed-memory services, code running after its pool is sealed or
 * on an out-of-memory path).  No counterpart in the reference.  The scratch is a 260 KiB
 * array built into the library image (held by every GPU that loads the library until exit),
 * so K is limited to nfb_stats_static_max_rows() = 1024; larger K returns NFB_ERANGE.  The
 * first call on each device looks up the array's device address — make it before any stream
 * capture.  All calls on a GPU share the array: calls that can overlap on the GPU race, so
 * keep them ordered.  Allocates nothing.                                                    */
int nfb_stats_static_max_rows(void);
int nfb_update_stats_static(const void* x, int x_dtype, int64_t K, int64_t N, int64_t row_stride,
                            float* sumsq, int* nonfinite, void* stream);

/* Per-tensor sum of squares over a table of T device pointers/lengths (all one dtype).
 * ptrs/lens are device arrays.  Replaces clip_grad_norm_'s norm
 * (nanofed/trainer/private.py:62-63) and per-tensor norms of validation.py:93-95.  */
int nfb_multi_sumsq(const void* const* ptrs, const int64_t* lens, int T, int dtype,
                    float* sumsq, int* nonfinite, void* scratch, size_t scratch_bytes,
                    void* stream);

/* out[k,s] = sum x[k, seg[s].offset : +seg[s].len]^2 ; segs = int64 [S,2] on device.  */
int nfb_segment_sumsq(const void* x, int x_dtype, int64_t K, int64_t row_stride,
                      const int64_t* segs, int S, float* out, void* scratch,
                      size_t scratch_bytes, void* stream);

/* ---- K-AGG -----------------------------------------------------------------
 * The fused FedAvg reduction:
 *   w_k   = counts[k]*mask[k] / sum(counts*mask)        (flags&1: normalise)
 *   c_k   = min(clip_norm / (sqrt(sumsq[k]) + 1e-6), 1) (clip_norm>0 and sumsq!=NULL)
 *   out[j] (+)= sum_k w_k c_k x[k,j] + noise_std*sqrt(sum_k w_k^2) * z[j] * noise_mask[j]
 *   out_lowp[j] = cast(out[j])                           (optional bf16/fp16 copy)
 * z = unit_noise[j] if given, else the Philox normal stream at (seed, offset+j).
 * flags: bit0 normalise weights, bit1 accumulate into out (streaming waves).
 * fp32 accumulation; x may be fp32/bf16/fp16; rows with w_k c_k == 0 are skipped,
 * so a masked-out client may hold NaNs.  One launch replaces upload + FedAvg + DP
 * clip + noise + cast: nanofed/communication/http/client.py:174-195, server.py:228-275,
 * orchestration/coordinator.py:307-322, server/aggregator/fedavg.py:53-63,
 * privacy/mechanisms.py:85-129 (via aggregator/privacy.py:179-194,322-335,88-92).
 */
int nfb_fedavg_reduce(const void* x, int x_dtype, int64_t K, int64_t N, int64_t row_stride,
                      const float* counts, const float* sumsq, const float* mask,
                      float clip_norm, float noise_std, const uint8_t* noise_mask,
                      const float* unit_noise, uint64_t seed, uint64_t offset, float* out,
                      void* out_lowp, int lowp_dtype, int flags, void* reserved, void* stream);

/* ---- K-BCAST ---------------------------------------------------------------
 * Replicate the fp32 global vector into K client slots (fp32 and/or bf16 working
 * copies) and zero the K momentum rows in the same pass: GET /model + load_state_dict +
 * fresh optimizer per round (nanofed/communication/http/server.py:191-208,
 * client.py:119-135, examples/mnist/run_experiment.py:65-73).  Strides in elements.  */
int nfb_bcast_global(const float* master, int64_t N, int64_t K, float* slots,
                     int64_t slots_stride, void* slots_bf16, int64_t lowp_stride,
                     float* momentum, int64_t momentum_stride, void* stream);

/* ---- K-OPT -----------------------------------------------------------------
 * torch.optim.SGD / Adam(W) semantics on flat buffers, optional DP-SGD front end
 * (g' = g*grad_scale[0] + noise_std*z) and bf16 working-copy write-back.
 * Replaces the user-supplied optimizer (examples/mnist/run_experiment.py:70-73) and
 * clip + per-param noise of nanofed/trainer/private.py:54-86.
 * row_len / zero_prefix: when zero_prefix > 0 the buffers are [n/row_len][row_len] and
 * g[r][0:zero_prefix) is set to zero after it has been consumed (gradient entries that the next
 * step accumulates into with atomics); pass 0, 0 otherwise.                               */
int nfb_fused_sgd(float* p, void* g, int g_dtype, float* momentum_buf, void* lowp_out,
                  int64_t n, float lr, float momentum, float dampening, float weight_decay,
                  int nesterov, int first_step, const float* grad_scale, float noise_std,
                  uint64_t seed, uint64_t offset, int64_t row_len, int64_t zero_prefix,
                  void* stream);
/* Same update, but columns [skip_lo, skip_hi) of every row_len row are neither read nor written:
 * they belong to a parameter whose gradient was applied by the kernel that produced it
 * (nfb_tc_linear_wgrad_sgd in csrc/tc_gemm.cu); pass skip_lo == skip_hi for none.  The bounds
 * must be multiples of 4.  zero_from > 0: g[r][zero_from:row_len) is zeroed after use too.  */
int nfb_fused_sgd_skip(float* p, void* g, int g_dtype, float* momentum_buf, void* lowp_out,
                       int64_t n, float lr, float momentum, float dampening, float weight_decay,
                       int nesterov, int first_step, const float* grad_scale, float noise_std,
                       uint64_t seed, uint64_t offset, int64_t row_len, int64_t zero_prefix,
                       int64_t skip_lo, int64_t skip_hi, int64_t zero_from, void* stream);
int nfb_fused_adam(float* p, const void* g, int g_dtype, float* exp_avg, float* exp_avg_sq,
                   void* lowp_out, int64_t n, float lr, float beta1, float beta2, float eps,
                   float weight_decay, float bias_corr1, float bias_corr2, int decoupled,
                   const float* grad_scale, void* stream);

/* ---- K-XENT ----------------------------------------------------------------
 * For G groups of [B, C] logits: log-softmax, NLL sum, #correct, and
 * dlogits = (softmax - onehot)*grad_scale in one pass; stats[g][0..2] +=
 * (loss_sum, correct, B) when stats != NULL (device-side accumulators that replace the
 * two .item() syncs per batch of nanofed/trainer/base.py:153-159).  Covers
 * nanofed/models/mnist.py:28 + trainer/torch.py:14,20-22.  targets_i64: 1 = int64.   */
int nfb_xent_fused(const void* logits, int dtype, const void* targets, int targets_i64,
                   int64_t G, int B, int C, void* dlogits, int dlogits_dtype, float grad_scale,
                   float* loss_sum, float* correct, float* stats, void* stream);

/* ---- K-PSGN ----------------------------------------------------------------
 * Ghost norms of a linear layer: out[i] = |a_i|^2 |d_i|^2 (+ |d_i|^2 with bias).
 * North-star extension of nanofed/trainer/private.py:54-63 (batch-level clip only).  */
int nfb_linear_psgn(const void* act, int act_dtype, const void* dout, int dout_dtype,
                    int64_t B, int64_t in_features, int64_t out_features, int bias,
                    float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NANOFED_B200_H_ */
