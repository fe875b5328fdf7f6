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
/* Optional K-BCAST target of nfb_fedavg_reduce (host struct passed as `bcast`): the finished
 * columns are also written into `rows` client rows (fp32 and/or bf16) and `rows` momentum rows are
 * zeroed, in the same pass - the next round's broadcast + optimizer reset
 * (examples/mnist/run_experiment.py:65-73) without a second kernel.  The rows may be x itself.  */
typedef struct {
  float* slots;
  int64_t slots_stride;
  void* lowp;               /* bf16 rows */
  int64_t lowp_stride;
  float* momentum;
  int64_t momentum_stride;
  int64_t rows;
} nfb_bcast_target;
int nfb_fedavg_reduce(const void* x, int x_dtype, int64_t K, int64_t N, int64_t row_stride,
                      const float* counts, const float* sumsq, const float* mask,
                      float clip_norm, float noise_std, const uint8_t* noise_mask,
                      const float* unit_noise, uint64_t seed, uint64_t offset, float* out,
                      void* out_lowp, int lowp_dtype, int flags, void* bcast, void* stream);

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
/* nfb_fused_sgd_skip plus a permuted bf16 side output: the parameter stored as
 * [perm_rows][perm_inner][perm_taps] at columns perm_lo.. of every row is also written as
 * [perm_rows][perm_taps][perm_inner] to perm_out + row * perm_group_stride (tap-major conv weights
 * for the implicit-GEMM kernels; replaces nfb_mnist_prep_conv2 after the first step).      */
int nfb_fused_sgd_perm(float* p, void* g, int g_dtype, float* momentum_buf, void* lowp_out,
                       int64_t n, float lr, float momentum, float dampening, float weight_decay,
                       int nesterov, int first_step, const float* grad_scale, float noise_std,
                       uint64_t seed, uint64_t offset, int64_t row_len, int64_t zero_prefix,
                       int64_t skip_lo, int64_t skip_hi, int64_t zero_from, void* perm_out,
                       int64_t perm_group_stride, int64_t perm_lo, int64_t perm_rows,
                       int perm_inner, int perm_taps, void* stream);
/* DP-SGD for a whole cohort: row r (client) is clipped by min(1, clip_norm / sqrt(row_sumsq[r]))
 * and gets N(0, noise_std^2) on the entries noise_mask marks (nanofed/trainer/private.py:54-86:
 * clip_grad_norm_ + per-parameter Gaussian noise); step_ctr (device, optional) advances the noise
 * stream between the steps of a replayed CUDA graph; perm_out optional.                        */
int nfb_fused_sgd_dp(float* p, void* g, int g_dtype, float* momentum_buf, void* lowp_out, int64_t n,
                     float lr, float momentum, float dampening, float weight_decay, int nesterov,
                     int first_step, float noise_std, uint64_t seed, uint64_t offset,
                     int64_t row_len, int64_t zero_prefix, int64_t skip_lo, int64_t skip_hi,
                     int64_t zero_from, const float* row_sumsq, float clip_norm, const uint8_t* noise_mask,
                     const unsigned long long* step_ctr, void* perm_out, int64_t perm_group_stride,
                     int64_t perm_lo, int64_t perm_rows, int perm_inner, int perm_taps,
                     float* sumsq_out, void* stream);
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

/* ---- K-AGG + K-BCAST across GPUs (csrc/fedavg_allreduce.cu) -------------------------------
 * ONE cooperative kernel per rank: local weighted reduce of this GPU's K client rows, chunk by
 * chunk; reduce-scatter of each chunk over peer memory (P2P loads, or multimem.ld_reduce when the
 * mc_* pointers are given); FedAvg epilogue (1/sum n, DP clip factors, folded Philox noise, bf16
 * cast); all-gather by stores into every rank's model buffers (P2P or multimem.st).
 * views_dev: device array of G rows of 11 int64 pointers
 *   {slots, counts, mask, sumsq, partial, scal, flags, master, lowp, counters, stats_all}
 * (entries of other ranks are peer-mapped addresses of their symmetric buffers; slots / counts /
 * mask / sumsq / counters only for my_rank).  ranks_in_grid = G runs all ranks in one launch on
 * one GPU (test emulation, blockIdx.y = rank).  epoch must increase by one per call; producers =
 * CTAs that reduce local clients (the rest consume), chunk_elems = pipeline granule (multiple of
 * 4), grid_blocks <= nfb_allreduce_max_blocks().  Waits are bounded by spin_limit_ns and set
 * *status instead of hanging.  flags bit0: add this call's clients to the partial / weight sums
 * left by earlier calls (streaming waves), bit1: local phase only (no flag, no peer traffic) -
 * every wave but the last passes 3 (the first: 2), the last passes 1.  my_stats (optional,
 * stats_len floats): this rank's per-client round metrics, copied into every rank's
 * stats_all[my_rank] before the completion flag (job-wide metrics_round_N.json,
 * nanofed/orchestration/coordinator.py:324-340).  When every client of every rank is masked out
 * the previous global model is kept.  bcast (optional, host nfb_bcast_target): K-BCAST as the tail
 * of the same kernel - after the completion barrier every CTA replicates the new model into the
 * local client rows and zeroes the momentum rows (counters needs 2 + 64 words then).
 * dbg_timestamps (optional, 8 words): %globaltimer at the kernel's phase boundaries.  Replaces
 * nanofed/server/aggregator/fedavg.py:53-63 +
 * communication/http upload/download (client.py:174-195, server.py:191-208).              */
int nfb_allreduce_max_blocks(int x_dtype, int64_t K);
int nfb_fedavg_allreduce(const void* views_dev, int my_rank, int G, int ranks_in_grid,
                         int x_dtype, int64_t K, int64_t N, int64_t row_stride, float clip_norm,
                         float noise_std, const uint8_t* noise_mask, uint64_t seed,
                         uint64_t offset, uint32_t epoch, int producers, int64_t chunk_elems,
                         int grid_blocks, const float* mc_partial, float* mc_master,
                         void* mc_lowp, int* status, uint64_t spin_limit_ns, int flags,
                         const float* my_stats, int stats_len, const void* bcast,
                         unsigned long long* dbg_timestamps, void* stream);

/* ---- K-GEMM / K-CONV: tcgen05 kernels (csrc/tc_gemm.cu, tc_fc.cu, tc_conv.cu, tc_conv1.cu) ---
 * Operands are bf16 [groups][rows][inner] views with unit inner stride (group = client); every
 * function launches ONE kernel for all groups.  `status` is a device int set to 1 if a bounded
 * pipeline wait expires.  Together they replace nn.Conv2d / nn.Linear forward and autograd
 * backward of nanofed/models/mnist.py:9-14,17-27 as run by trainer/base.py:145-150.        */
typedef struct {
  const void* ptr;
  long long inner, rows, groups, row_stride, group_stride; /* elements */
} nfb_tc_operand;
typedef struct {
  void* out;
  int out_bf16, atomic;
  long long out_gs, out_ms, out_ns, out_ts; /* element strides: group, m, n, tap */
  const float* bias;
  long long bias_gs;
  int bias_on_m, relu;
  const void* mask_src; /* bf16, same indexing as out: out = 0 where mask <= 0 */
  int m_valid, n_valid;
} nfb_tc_epilogue;
/* D_t[g][m,n] = sum_k A[g][m + a_mn(t), k + a_k(t)] * B[g][n + b_mn(t), k + b_k(t)].
 * variant: 0 conv_fwd, 1 conv_dgrad, 2 conv_wgrad (9 accumulators), 3 fc_fwd, 4 fc_dgrad,
 * 5 fc_wgrad, 6 fc_dgrad with M = batch.  tap_table: taps x {a_k, a_mn, b_k, b_mn}.          */
int nfb_tc_gemm(int variant, const nfb_tc_operand* A, const nfb_tc_operand* B, int m_tiles,
                int n_tiles, int groups, int taps, int inner, int k_split, const int* tap_table,
                const nfb_tc_epilogue* epi, int* status, void* stream);
/* w[g][j][i] -= lr * (sum_b dy[g][b][j] * x[g][b][i] + wd * w), w_lowp = bf16(w): the weight
 * gradient is applied in the GEMM epilogue and never written.                              */
int nfb_tc_linear_wgrad_sgd(const nfb_tc_operand* dy, const nfb_tc_operand* x, int groups, int j,
                            int i, int inner, float* w, long long w_gs, void* w_lowp,
                            long long lowp_gs, float lr, float weight_decay, float* sumsq_out,
                            int* status, void* stream);
/* out[g][b,i] = scale * sum_j dy[g][b,j] * w[g][j,i]  (persistent; j = 128, b <= 128).      */
/* K-PSGN on tensor cores: out[i] += ||sum_p dy[i][p][:]^T u[i][p][:]||_F^2 per sample (group) i;
 * dy [samples][P][cout], u [samples][P][kdim] bf16 (unfolded patches, optionally with a ones
 * column for the bias term); inner = ceil(P / 64); out zero on entry.  Replaces the whole-batch
 * clip_grad_norm_ of nanofed/trainer/private.py:62-63 with true per-sample norms.           */
/* Batch ghost norm: out[c] = |sum_b d[c][b] (x) x[c][b]|_F^2 + add_a[c] + add_b[c] from gram_x[c] =
 * x x^T ([B][B] fp32, consumed and zeroed) and d [C][B][J] bf16 - the squared norm of a linear
 * layer's batch weight gradient without materialising it (DP-SGD clip of the grouped engine;
 * reference nanofed/trainer/private.py:62-63 computes it from the materialised gradient).    */
int nfb_batch_ghost_norm(float* gram_x, const void* d, int clients, int batch, int J,
                         const float* add_a, const float* add_b, float* out, void* stream);
/* nfb_tc_linear_wgrad_sgd + DP-SGD front end in the epilogue (clip factor from row_sumsq, Philox
 * noise at the weight's position in the [groups][row_len] slab: same stream as nfb_fused_sgd_dp). */
int nfb_tc_linear_wgrad_sgd_dp(const nfb_tc_operand* dy, const nfb_tc_operand* x, int groups, int j,
                               int i, int inner, float* w, long long w_gs, void* w_lowp,
                               long long lowp_gs, float lr, float weight_decay,
                               const float* row_sumsq, float clip_norm, float noise_std,
                               uint64_t seed, uint64_t offset, const unsigned long long* step_ctr,
                               long long row_len, long long w_off, int* status, void* stream);
int nfb_tc_psgn(const nfb_tc_operand* dy, const nfb_tc_operand* u, int samples, int cout, int kdim,
                int inner, float* out, int* status, void* stream);
int nfb_tc_fc_dgrad(const nfb_tc_operand* dy, const nfb_tc_operand* w, void* out, long long out_rs,
                    long long out_gs, float scale, int* status, void* stream);
/* 3x3 convolutions on the row-flattened NHWC grid (activation = [clients][B*W*W rows][ch], tap
 * (kh,kw) = row shift kh*W + kw); persistent kernels of csrc/tc_conv.cu.  w = bf16 [C][64][288]
 * (tap-major, channel-minor; see nfb_mnist_prep_conv2).
 * mode 0: out[q,co] = relu(sum x[q+s_t, ci] w[co, t*32+ci] + bias[co])          (32 -> 64 ch)
 * mode 1: out[q,ci] = (sum dy[q-s_t, co] w[co, t*32+ci]) * (mask[q,ci] > 0)     (64 -> 32 ch) */
int nfb_tc_conv3x3(int mode, const void* a, long long rows, long long a_group_stride,
                   const void* w, long long w_group_stride, int clients, int width, void* out,
                   long long out_group_stride, const float* bias, long long bias_group_stride,
                   const void* mask, int* status, void* stream);
/* mode 0's convolution with bias + ReLU + 2x2 max-pool + dropout in the epilogue: writes only
 * pool_out [C][B][64*12*12] (PyTorch flatten order) and one code byte per pooled element
 * (bits 0-1 argmax slot, bit 2 ReLU passed, bit 3 kept by dropout); width must be 26.        */
int nfb_tc_conv3x3_pool_fwd(const void* x, long long rows, long long x_group_stride, const void* w,
                            long long w_group_stride, int clients, int images, int width,
                            const float* bias, long long bias_group_stride, void* pool_out,
                            uint8_t* pool_idx, float drop_p, const void* rng_state,
                            unsigned long long seed, int* status, void* stream);
/* grad[c][co][ci][kh][kw] += sum_q dy[c,q,co] * x[c, q + kh*W + kw, ci]   (fp32 atomics).     */
int nfb_tc_conv3x3_wgrad(const void* dy, const void* x, long long rows, long long dy_group_stride,
                         long long x_group_stride, int clients, int width, float* grad,
                         long long grad_group_stride, int* status, void* stream);
/* mode 1 fused with the weight/bias gradient of the 1 -> 32 convolution in front of it: the
 * masked data gradient feeds a second MMA against the 3x3 windows of the layer-1 input x1
 * (fp32 [C][B][(W+2)^2]) and is never stored.  grad1[c][w1_off + ci*9 + t], grad1[c][b1_off + ci]
 * are accumulated with atomics.                                                             */
int nfb_tc_conv3x3_dgrad_wgrad1(const void* dy, long long rows, long long dy_group_stride,
                                const void* w, long long w_group_stride, int clients, int width,
                                const void* mask, long long mask_group_stride, const float* x1,
                                long long x1_group_stride, float* grad1,
                                long long grad1_group_stride, long long w1_off, long long b1_off,
                                int* status, void* stream);
/* a1[c][q][ch] = relu(conv1(x)[q][ch] + bias[ch]), 1 -> 32 channels, one tcgen05.mma pair per
 * 128 positions with bf16 hi/lo split operands; weights/bias read from the fp32 parameter rows. */
int nfb_tc_conv1_fwd(const float* x, long long x_group_stride, const float* slots,
                     long long slot_group_stride, long long w_off, long long b_off, int clients,
                     int images, int width, void* a1, long long a1_group_stride, int* status,
                     void* stream);

/* ---- K-ELT / K-XENT pieces of the MNIST CNN cohort engine (csrc/mnist_ops.cu) ----------------
 * slots / grads are the [C][row_stride] fp32 parameter and gradient rows of the arena; *_off are
 * element offsets of a state_dict entry inside a row (runtime/flat.py).                     */
int nfb_mnist_prep_conv2(const float* slots, int64_t row_stride, int64_t w_off, int C, void* w2p,
                         void* stream);  /* OIHW fp32 -> [C][64][tap*32+ci] bf16 */
int nfb_mnist_conv1_fwd(const float* x, int64_t x_cstride, const float* slots, int64_t row_stride,
                        int64_t w_off, int64_t b_off, int C, int B, void* a1, void* stream);
int nfb_mnist_pool_fwd(const void* a2, int C, int B, float drop_p, const void* rng_state,
                       uint64_t seed, void* p, uint8_t* pidx, void* stream);
/* dp [C][B][9216] + code bytes -> da2 [C][B*676][64] (zero outside the 24x24 window) and the
 * conv2 bias gradient; drop_p = 0 when the producer of dp already applied 1/(1-p).           */
int nfb_mnist_pool_bwd(const void* dp, const uint8_t* pidx, int C, int B, float drop_p, void* da2,
                       float* grads, int64_t row_stride, int64_t b2_off, void* stream);
/* fc1 bias + ReLU + dropout, fc2, log-softmax/NLL/accuracy into stats[c][0..2], dlogits, fc2 and
 * fc1-bias gradients (fp32 atomics when the batch is split over row blocks: targets must be
 * zero), dh = gradient w.r.t. the fc1 pre-activation (bf16).  hpre is consumed and left zeroed.
 * rng_state: 2 x uint64 {dropout step, block arrival counter}.                              */
int nfb_mnist_fc_head(float* hpre, const int64_t* labels, int64_t label_stride, int C, int B,
                      const float* slots, float* grads, int64_t row_stride, int64_t b1_off,
                      int64_t w2_off, int64_t b2_off, float drop_p, void* rng_state,
                      int advance_rng, uint64_t seed, void* dh, float* stats, void* stream);
int nfb_mnist_conv1_wgrad(const void* da1, const float* x, int64_t x_cstride, int C, int B,
                          float* grads, int64_t row_stride, int64_t w_off, int64_t b_off,
                          void* stream);

/* ---- ResNet cohort engine: generic K-CONV (csrc/rn_conv.cu) and K-ELT (csrc/rn_elt.cu) ---------
 * No counterpart in the reference: these replace the torch library calls (Conv2d, BatchNorm2d,
 * ReLU, AdaptiveAvgPool2d, Linear, CrossEntropyLoss, SGD.step) that a ResNet-18 executes inside
 * nanofed/trainer/base.py:134-160, for all clients of a GPU per launch.
 * An activation is bf16 NHWC [imgs][H][W][C] with imgs = clients * batch, C a multiple of 64 and a
 * 128-byte aligned base; `parity` != 0 addresses it through the stride-2 view [2C, W/2, 2, H/2].  */
typedef struct {
  const void* ptr;
  int imgs, H, W, C, parity;
} nfb_rn_act;
/* taps: n_taps x {dh, dw, ph, pw, wrow, src, cls}: row/column shift of the A box, parity of the A
 * view, first weight row of the tap, operand pair (0 | 1), output parity class 2*ph_out + pw_out the
 * tap contributes to (used with n_cls = 4: the four classes of a stride-2 data gradient in one
 * launch; -1 / n_cls = 1: every tap counts and out_ph / out_pw select the class).  dgrad = 0: out[pos, n] = sum A[pos+tap, k] *
 * W[wrow + n][k] (weights K-major); dgrad = 1: ... * W[wrow + k][n].  w_dims = {inner, rows, client
 * stride} per pair.  H x W is the position grid of the output; stats (optional) receives per
 * client the per-channel sum and sum of squares of the stored bf16 values ([clients][2][n_out],
 * fp32 atomics); residual (+ optional ReLU mask) is added before the store; out_mask (dense
 * tensor of the output's shape) zeroes the stored value where it is not positive (the ReLU
 * backward of the layer the gradient flows into).  bn_x (dgrad, with stats): the raw conv output
 * under that activation's BatchNorm - stats then receives sum(g) and sum(g * x) of the stored
 * gradient (the reduction pass of the BatchNorm backward, folded into this epilogue).
 * gn_stats (forward, dense output): [images][8][2] GroupNorm(8) sums of the stored values,
 * accumulated (the statistics pass of nfb_rn_gn_stats folded into this epilogue).             */
int nfb_rn_conv_gemm(int dgrad, const nfb_rn_act* a0, const nfb_rn_act* a1, const void* w0,
                     const void* w1, const int* w_dims, int clients, int imgs_per_client, int H,
                     int W, int n_out, int n_taps, const int* taps, const nfb_rn_act* out,
                     int out_ph, int out_pw, float* stats, const void* residual,
                     const void* res_mask, const void* out_mask, const void* bn_x,
                     float* gn_stats, int bn_tile, int halves, int n_cls, int* status,
                     void* stream);
/* out[client][tap][co][ci] (+)= sum_pos x[pos + tap][ci] * dy[pos][co]; k_split > 1 adds with fp32
 * atomics (the target must then be zero).  With k_split == 1 and sgd_w != NULL the gradient is not
 * written at all: the epilogue applies w -= lr * (dW + wd * w) to the tap-major fp32 master sgd_w
 * (same layout and client stride as out) and refreshes the bf16 copy sgd_wt.                  */
int nfb_rn_conv_wgrad(const nfb_rn_act* x, const nfb_rn_act* dy, int clients, int imgs_per_client,
                      int n_taps, const int* taps, float* out, long long out_gs, int k_split,
                      int bn_tile, int posk, float* sgd_w, void* sgd_wt, float lr, float wd,
                      int* status, void* stream);
int nfb_rn_im2col3(const float* x, long long x_gs, int clients, int batch, void* out, void* stream);
/* bn_offs / res_offs: {weight, bias, running_mean, running_var, num_batches_tracked} offsets in a
 * client's fp32 row; stats as written by nfb_rn_conv_gemm.                                    */
int nfb_rn_bn_apply(const void* x, const void* res, void* y, float* slots, long long row_stride,
                    const float* stats, const long long* bn_offs, const float* res_stats,
                    const long long* res_offs, int clients, int P, int C, int relu, float eps,
                    float momentum, void* stream);
/* apply = 0: accumulate dgamma / dbeta (fp32 atomics into sgrad rows); apply = 1: write dx (dx2).
 * raw_sums (apply only): [clients][2][C] sum(g), sum(g * x) left by the dgrad kernel that produced
 * dout - no reduce pass is needed then, dgamma / dbeta are derived and stored by this launch.    */
int nfb_rn_bn_bwd(int apply, const void* dout, const void* out, const void* x, const void* x2,
                  void* dx, void* dx2, const float* slots, long long row_stride, float* sgrad,
                  long long sgrad_stride, const long long* grad_offs, const float* stats,
                  const long long* bn_offs, const float* stats2, const long long* bn2_offs,
                  int clients, int P, int C, float eps, const float* raw_sums, void* stream);
/* GroupNorm(8) of the ResNet engine (csrc/rn_gn.cu).  Activations [imgs][HW][C] bf16; stats and
 * bsums are [imgs][8][2] fp32 and must be zero before the launch that accumulates into them
 * (gn_stats / the reduce pass of gn_bwd).  gn_offs = {weight, bias} offsets in a client's row;
 * image i belongs to client i / imgs_per_client.                                               */
int nfb_rn_gn_stats(const void* x, int imgs, int HW, int C, float* stats, void* stream);
int nfb_rn_gn_apply(const void* x, const void* res, void* y, const float* slots,
                    long long row_stride, const float* stats, const long long* gn_offs,
                    const float* res_stats, const long long* res_offs, int imgs,
                    int imgs_per_client, int HW, int C, int relu, float eps, void* stream);
int nfb_rn_gn_bwd(int apply, const void* dout, const void* x, const void* x2, void* dx, void* dx2,
                  const float* slots, long long row_stride, float* sgrad, long long sgrad_stride,
                  const long long* grad_offs, const float* stats, const long long* gn_offs,
                  const float* stats2, const long long* gn2_offs, float* bsums, float* bsums2,
                  int imgs, int imgs_per_client, int HW, int C, float eps, void* stream);
int nfb_rn_head(const void* feat, int clients, int batch, int positions, const int64_t* labels,
                long long label_gs, const float* slots, long long row_stride, long long w_off,
                long long b_off, float* sgrad, long long sgrad_stride, long long dw_off,
                long long db_off, void* dfeat, float* stats, void* stream);
/* conv weights: arena rows (torch layout) <-> tap-major fp32 master wm + bf16 copy wt.       */
int nfb_rn_wm_convert(const void* table, int n_layers, long long total_cols, float* slots,
                      long long row_stride, float* wm, void* wt, long long t_stride, int clients,
                      int to_slots, void* stream);
/* ranges: n_ranges x {t_off, len} (int64, device, len % 4 == 0).                              */
/* Client-side DP-SGD inside the ResNet engine's update kernels (NULL or row_sumsq == NULL: plain
 * SGD): g' = g * min(1, clip_norm / (||g_client|| + 1e-6)) + noise_std * z.  row_sumsq [clients of
 * the launch] is the squared norm of the client's WHOLE gradient; the Philox stream position is
 * ((*step_base + step) * clients_total + client0 + g) * row_groups + (idx_base + index) / 4.   */
typedef struct {
  const float* row_sumsq;
  float clip_norm, noise_std;
  unsigned long long seed;
  const unsigned long long* step_base;
  int step, client0, clients_total;
  long long row_groups, idx_base;
} nfb_rn_dp;
int nfb_rn_tap_sgd(const void* ranges, int n_ranges, long long max_len, float* wm, float* dwt,
                   float* momentum, void* wt, long long t_stride, int clients, float lr, float mu,
                   float wd, const nfb_rn_dp* dp, void* stream);
int nfb_rn_small_sgd(const void* table, int n_entries, float* slots, long long row_stride,
                     float* momentum, float* sgrad, long long sgrad_stride, int clients, float lr,
                     float mu, float wd, float* zero_ptr, long long zero_len, const nfb_rn_dp* dp,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NANOFED_B200_H_ */
