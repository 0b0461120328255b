/* commefficient_b200 -- C ABI of the native sm_100a library (libceb200.so).
 *
 * Conventions
 *   - every function returns 0 on success; on failure a non-zero code, and
 *     ceb_last_error() describes what went wrong (thread-local string);
 *   - pointers are device pointers unless a comment says "host";
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - compute entry points never allocate or free device memory: callers pass every
 *     buffer, including scratch (sizes via the *_bytes() queries), so the library can be
 *     driven from PyTorch's caching allocator without fragmenting it (the ceb_peer_alloc
 *     family, which creates IPC-exportable buffers, is the exception);
 *   - top-k results are deterministic: the entries above the k-th largest magnitude in ascending
 *     index order, then the lowest-index entries equal to it (replicas that select on the same
 *     vector pick the same set; torch.topk leaves this unspecified).
 *
 * "Reference" below is pursueorigin/CommEfficient (paths relative to
 * CommEfficient/ in that repository).
 */
#ifndef COMMEFFICIENT_B200_H
#define COMMEFFICIENT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- library ---------------------------------------------------------------- */
const char* ceb_last_error(void);
int ceb_version(void);
/* Kernel launches issued by this library since the last reset (host counter). */
unsigned long long ceb_launch_count(void);
void ceb_reset_launch_count(void);
/* Credit n launches replayed from a captured CUDA graph (they bypass the library's host code). */
void ceb_add_launch_count(unsigned long long n);

/* ---- exact top-k by magnitude (reference: utils.py:232-252 `_topk`) ---------- */
/* Bytes of zero-initialisable scratch every top-k style call needs (~34 KB). */
int64_t ceb_select_scratch_bytes(void);
/* out_idx/out_val receive the k largest-|v| entries of vec[0..n). 0 < k <= n < 2^31. */
int ceb_topk_select(const float* vec, int64_t n, int64_t k, void* scratch, int64_t scratch_bytes,
                    int32_t* out_idx, float* out_val, void* stream);

/* ---- count sketch (reference: external csvec.CSVec; fed_worker.py:312-320,
 *      fed_aggregator.py:568-613).  packed: [r][n_tokens] int32, blk: [num_blocks]
 *      int32 -- layout documented in commefficient_b200/ops/hashing.py ---------- */
/* table[r][bucket(i)] += sign(i) * scale * (g[i] + wd * w[i]); w may be NULL. */
int ceb_sketch_accumulate(const float* g, const float* w, float wd, float scale, int64_t d,
                          const int32_t* packed, int64_t n_tokens, const int32_t* blk, int num_blocks,
                          float* table, int r, int64_t c, void* stream);
/* Same restricted to coordinates lo <= i < hi (g and w are indexed by the global coordinate): sketch the
 * part of a gradient that is already final while the rest is still being computed (second stream). */
int ceb_sketch_accumulate_range(const float* g, const float* w, float wd, float scale, int64_t d, int64_t lo, int64_t hi,
                                const int32_t* packed, int64_t n_tokens, const int32_t* blk, int num_blocks,
                                float* table, int r, int64_t c, void* stream);
/* out2[1] = l2 estimate sqrt(median_r sum_c T^2); out2[0] = clip/l2 if l2 >= clip > 0 else 1.
 * rowsq_tmp: r floats of scratch.  (reference: utils.py:305-313 on a CSVec) */
int ceb_sketch_l2_scale(const float* table, int r, int64_t c, float clip, float* rowsq_tmp, float* out2,
                        void* stream);
/* y += alpha * (dev_scale ? *dev_scale : 1) * x; zero_x != 0 clears x afterwards. */
int ceb_axpy_dev(float* y, float* x, int64_t n, float alpha, const float* dev_scale, int zero_x,
                 void* stream);
/* Vv = rho*Vv + g*inv_n; error_mode 2: Ve += Vv, 1: Ve = Vv, 0: Ve untouched.
 * (reference: fed_aggregator.py:578-582) */
int ceb_sketch_server_update(float* Vv, float* Ve, const float* g, float inv_n, float rho, int error_mode,
                             int64_t n, void* stream);
/* est[i] = median_r(table[r][bucket(i)]*sign(i)) for all i<d, then exact top-k of |est|.
 * (reference: CSVec.unSketch(k), fed_aggregator.py:584-590) */
int ceb_sketch_unsketch_topk(const float* table, const int32_t* packed, int64_t n_tokens,
                             const int32_t* blk, int num_blocks, int r, int64_t c, int64_t d, int64_t k,
                             float* est, void* scratch, int64_t scratch_bytes, int32_t* out_idx,
                             float* out_val, void* stream);
/* Allocation-free variant of ceb_sketch_unsketch_topk (no counterpart in the reference:
 * it exists for hosts of this library that may not allocate device memory -- a service whose
 * GPU memory was partitioned once at start-up, a plug-in inside another framework's pool, a
 * latency-critical loop that must avoid cudaMalloc's device-wide synchronisation).  Scratch
 * lives in a device array built into the library: CEB_STATIC_UNSKETCH_MAX_D floats + ~34 KB
 * on every GPU that loads the library, held until exit.  d above the maximum is an error.
 * All calls on a device share that array (overlapping calls race; serialise them), and the
 * first call on a device looks up its address -- do that call outside stream capture. */
#define CEB_STATIC_UNSKETCH_MAX_D (8 * 1024 * 1024)
int ceb_sketch_unsketch_topk_static(const float* table, const int32_t* packed, int64_t n_tokens,
                                    const int32_t* blk, int num_blocks, int r, int64_t c, int64_t d,
                                    int64_t k, int32_t* out_idx, float* out_val, void* stream);
/* out[j] = estimate of coordinate idx[j] (idx NULL: coordinates 0..count-1). */
int ceb_sketch_query(const float* table, const int32_t* packed, int64_t n_tokens, const int32_t* blk,
                     int num_blocks, int r, int64_t c, int64_t d, const int32_t* idx, int64_t count,
                     float* out, void* stream);
/* For every selected (idx,val) with val != 0: zero its r buckets in Vv / Ve (either may be
 * NULL) and apply w[idx] -= lr*val (lr_vec[idx] when non-NULL); changed[idx] = 1 when
 * non-NULL.  (reference: fed_aggregator.py:593-613, :455) */
int ceb_sketch_apply_zero(const int32_t* idx, const float* val, int64_t k, const int32_t* packed,
                          int64_t n_tokens, const int32_t* blk, int r, int64_t c, float* Vv, float* Ve,
                          float* w, float lr, const float* lr_vec, uint8_t* changed, void* stream);
/* Same with the scalar learning rate read from device memory at run time (CUDA-graph replays). */
int ceb_sketch_apply_zero_dev(const int32_t* idx, const float* val, int64_t k, const int32_t* packed, int64_t n_tokens,
                              const int32_t* blk, int r, int64_t c, float* Vv, float* Ve, float* w, const float* lr_dev,
                              uint8_t* changed, void* stream);

/* ---- dense modes (reference: fed_aggregator.py:483-566) ---------------------- */
/* Vv = rho*Vv + g*inv_n (+ N(0,noise_std^2)); w -= lr*Vv.  uncompressed / local_topk / fedavg. */
int ceb_dense_server_update(float* Vv, const float* g, float inv_n, float rho, int64_t n, float* w,
                            float lr, const float* lr_vec, float noise_std, unsigned long long seed,
                            unsigned long long noise_offset, uint8_t* changed, void* stream);
/* true_topk server step in one call (momentum, virtual error, exact top-k, zeroing, apply). */
int ceb_true_topk_step(float* Vv, float* Ve, const float* g, float inv_n, float rho, int64_t n, int64_t k,
                       float* w, float lr, const float* lr_vec, uint8_t* changed, void* scratch,
                       int64_t scratch_bytes, int32_t* out_idx, float* out_val, float* client_vel,
                       const int64_t* vel_rows, int n_vel_rows, void* stream);
int ceb_dense_apply_zero(const int32_t* idx, const float* val, int64_t k, float* Vv, float* Ve, float* w,
                         float lr, const float* lr_vec, uint8_t* changed, float* client_vel,
                         const int64_t* vel_rows, int n_vel_rows, int64_t d, void* stream);

/* ---- per-client state (reference: fed_worker.py:190-216) --------------------- */
/* transmit = [err +=] [vel = rho*vel +] scale*g; k == 0: acc += transmit; k > 0: acc[idx] +=
 * topk_k(transmit) and err[idx] = vel[idx] = 0.  vel/err may be NULL. */
int ceb_client_step(float* g, float scale, float* vel, float rho, float* err, int64_t n, int64_t k,
                    float* acc, void* scratch, int64_t scratch_bytes, int32_t* out_idx, float* out_val,
                    void* stream);
int ceb_scatter_add_zero(const int32_t* idx, const float* val, int64_t k, float* acc, float* err,
                         float* vel, void* stream);

/* ---- clipping, DP noise, fedavg, accounting ---------------------------------- */
/* x *= min(1, clip/(||x||+eps)) when eps > 0 (torch clip_grad_norm_, fed_worker.py:290-292);
 * eps == 0: x *= clip/||x|| only when ||x|| >= clip (utils.py:305-313).  sumsq_out: 1 float. */
int ceb_clip_by_norm(float* x, int64_t n, float clip, float eps, float* sumsq_out, void* stream);
/* x += N(0, std^2), Philox4x32-10 (reference: fed_worker.py:306-309). */
int ceb_add_gaussian_noise(float* x, int64_t n, float std, unsigned long long seed,
                           unsigned long long offset, void* stream);
/* acc += weight * (w0 - wloc)   (reference: fed_worker.py:104-108) */
int ceb_fedavg_delta(float* acc, const float* w0, const float* wloc, float weight, int64_t n, void* stream);
/* mask |= (w != prev); prev = w; *count = popcount(mask).  prev/mask may be NULL.
 * (reference: fed_aggregator.py:244-250) */
int ceb_count_changed(const float* w, float* prev, uint8_t* mask, int64_t n, unsigned long long* count,
                      void* stream);
/* Download-byte tracker (no counterpart in the reference, which diffs full weight copies on the host):
 * record: coordinates idx[0..k) move from bucket last[idx] to bucket `round` (last updated, hist
 * decremented / incremented); counts: out[j] = sum of hist[r] for r >= start[j]. */
int ceb_tracker_record(const int* idx, int64_t k, int64_t* last, int64_t* hist, int64_t round, void* stream);
int ceb_tracker_record_dev(const int* idx, int64_t k, int64_t* last, int64_t* hist, const int64_t* round_dev,
                           void* stream);
int ceb_tracker_counts(const int64_t* hist, int n, const int64_t* start, int m, float* out, void* stream);

/* ---- tensor-core layers: tcgen05.mma + TMEM accumulators + TMA-staged operands --
 * (reference: the cuBLAS / cuDNN calls behind nn.Linear / nn.Conv2d, models/resnet9.py:36-37,
 *  models/fixup_resnet18.py:144-154, GPT-2 linears).  bf16 operands, fp32 accumulation.
 *  out_mode 0: bf16 output, 1: fp32 output, 2: fp32 output += result (atomic). */
/* C[M,N] = act(A[M,K] * B[N,K]^T + bias); A, B row-major with K contiguous; K % 8 == N % 8 == 0. */
int ceb_gemm_bf16_tn(const void* A, const void* B, void* C, int M, int N, int K, const float* bias,
                     int relu, int out_mode, void* stream);
/* C[M,N] = act(A[M,K] * B[K,N] + bias): B row-major [K][N] ("MN-major" operand, read in place). */
int ceb_gemm_bf16_nn(const void* A, const void* B, void* C, int M, int N, int K, const float* bias,
                     int relu, int out_mode, void* stream);
/* y[N,H,W,Cout] = act(conv3x3 stride 1 pad 1 (x[N,H,W,Cin], w[Cout,3,3,Cin]) + bias), NHWC / KRSC.
 * Implicit GEMM: the im2col tile of tap (r,s) is a shifted 4-D TMA box, padding = TMA zero fill.
 * Cin % 64 == 0, Cout % 8 == 0, W divides 128.
 * relu_mask (nullable, bf16, same shape as y): y is zeroed where relu_mask <= 0 -- the ReLU
 * backward fused into a data-gradient convolution (dgrad = this kernel on dz with the weights
 * re-laid-out by ceb_conv_w_dgrad_layout). */
int ceb_conv3x3_bf16_nhwc(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout,
                          const float* bias, int relu, int out_mode, const void* relu_mask, void* stream);
/* Same with a fused residual add (bf16 output; Cout % 32 == 0; y, resid, y2, relu_mask 32-byte aligned):
 * y2 == NULL: y = act(conv) + resid;  y2 != NULL: y = act(conv), y2 = y + resid. */
int ceb_conv3x3_bf16_nhwc_res(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout,
                              const float* bias, int relu, const void* relu_mask, const void* resid, void* y2,
                              void* stream);
/* dst[ci][2-r][2-s][co] (bf16) = w[co][r][s][ci] (fp32): weights of the data-gradient convolution. */
int ceb_conv_w_dgrad_layout(const float* w_krsc, void* dst_bf16, int Cout, int Cin, void* stream);
int ceb_cast_f32_bf16(const float* src, void* dst_bf16, int64_t n, void* stream);
/* ceb_cast_f32_bf16 + ceb_conv_w_dgrad_layout for up to 16 conv layers in ONE launch: layer i has
 * fp32 [cout][3][3][cin] weights src[i]; cast_bf16[i] (may be NULL / the array may be NULL) receives
 * the bf16 copy, rot_bf16[i] the [cin][3][3][cout] data-gradient layout.  The pointer arrays are HOST
 * arrays of device pointers (copied into the kernel's parameter block). */
int ceb_conv_weights_prepare(int n_layers, const float* const* src, void* const* cast_bf16, void* const* rot_bf16,
                             const int* cout, const int* cin, void* stream);
/* dst[c][r] (bf16, row stride ldo >= R, ldo even) = src[r][c] (fp32 [R][C]): the K-major bf16 copy of
 * a weight stored the other way round (GPT-2 Conv1D [in][out]; the tied LM head in its dgrad). */
int ceb_cast_transpose_f32_bf16(const float* src, void* dst_bf16, int R, int C, int64_t ldo, void* stream);
/* ceb_gemm_bf16_tn with explicit row strides: lda, ldb multiples of 8 and >= K, ldc >= N and a
 * multiple of 8 (bf16 C) or 4 (fp32 C); M, N, K arbitrary (TMA zero-fills the tails, stores are
 * bounded).  All three bases 16-byte aligned. */
int ceb_gemm_bf16_tn_ld(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N,
                        int K, const float* bias, int relu, int out_mode, void* stream);
/* The same GEMM with a tanh-GELU in its epilogue (bf16 [M][ldc] outputs; N % 32 == 0, ldc % 16 == 0, 32-byte
 * aligned).  act = 2: C = A B^T + bias (pre-activation) and C2 = GELU(C); act = 3: C = (A B^T) * GELU'(aux) with
 * aux a saved pre-activation [M][ldc] (the transformer MLP forward / backward without element-wise passes). */
int ceb_gemm_bf16_tn_gelu(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, void* C2, const void* aux,
                          int64_t ldc, int M, int N, int K, const float* bias, int act, void* stream);
/* Same, computing only rows [0, *m_limit_dev) (whole 128-row tiles, never past M); m_limit_dev is a
 * DEVICE int read by the kernel, so a captured launch follows a data-dependent row count. */
int ceb_gemm_bf16_tn_rows(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N,
                          int K, const float* bias, int relu, int out_mode, const int* m_limit_dev, void* stream);
/* bf16 NHWC max pooling (k = 2 or 4) and its backward; relu_mask != 0 also applies the ReLU
 * backward of y (reference: resnet9.py:48-56,95). */
int ceb_maxpool_nhwc_bf16(const void* y, void* out, int N, int H, int W, int C, int k, void* stream);
int ceb_maxpool_bwd_nhwc_bf16(const void* y, const void* dp, void* dz, int N, int H, int W, int C, int k,
                              int relu_mask, void* stream);
int ceb_relu_bwd_bf16(const void* y, const void* dy, void* dz, int64_t n, void* stream);
/* 3-input-channel stem (reference: the `prep` layer, models/resnet9.py:85): K = 27 is too thin for an
 * implicit-GEMM tile, so it is lowered explicitly.  A[N*H*W][64] (bf16) = im2col of x[N,H,W,3]
 * (fp32), column (r*3+s)*3+ci, zero padded to 64; forward = ceb_gemm_bf16_tn(A, W0pad), weight
 * gradient = ceb_gemm_bf16_mn_acc(dZ, A). */
int ceb_im2col3x3_c3(const float* x_nhwc, void* A_bf16, int N, int H, int W, void* stream);
/* C[M][N] (fp32, row stride ldc) += A[K][M]^T * B[K][N]; A, B bf16 row-major ("both operands
 * MN-major": linear-layer weight gradients dW = dY^T X).  M % 64 == N % 64 == 0; split-K. */
int ceb_gemm_bf16_mn_acc(const void* A, const void* B, float* C, int M, int N, int K, int ldc, void* stream);
/* Same with explicit operand row strides (multiples of 8, lda >= M, ldb >= N); M arbitrary, N % 64 == 0. */
int ceb_gemm_bf16_mn_acc_ld(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N, int K,
                            int ldc, void* stream);
/* Same, reducing over rows [0, *k_limit_dev) only (whole 64-row steps; rows up to the end of the last
 * step must be finite, normally zero). */
int ceb_gemm_bf16_mn_acc_rows(const void* A, int64_t lda, const void* B, int64_t ldb, float* C, int M, int N, int K,
                              int ldc, const int* k_limit_dev, void* stream);
/* dW[Cout][3][3][Cin] (fp32) += sum_p dz[p][co] * x[p + (r-1,s-1)][ci]; dz, x bf16 NHWC.  Both GEMM
 * operands are MN-major (pixels are the reduction dimension); split-K with vector atomics.
 * Cin % 64 == 0, Cout % 64 == 0, W divides 64. */
int ceb_conv3x3_wgrad_bf16_nhwc(const void* dz, const void* x, float* dW, int Nimg, int H, int W, int Cin,
                                int Cout, void* stream);

/* ---- transformer-side ops (reference: pytorch_transformers' GPT-2 behind gpt2_train.py:76-99) ----
 * Softmax cross-entropy over bf16 logits rows (row stride ld elements, V classes, labels ==
 * ignore_index skipped): forward writes loss[row] and lse[row]; backward writes
 * dlogits = grad_rows[row] * (softmax - onehot) (may alias logits). */
int ceb_softmax_ce_fwd(const void* logits_bf16, int64_t ld, int64_t rows, int V, const int64_t* labels,
                       int64_t ignore_index, float* loss, float* lse, void* stream);
int ceb_softmax_ce_bwd(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                       const int64_t* labels, int64_t ignore_index, const float* lse, const float* grad_rows,
                       void* stream);
/* Compacted rows (valid labels first): rows past the 128-row tile holding row *row_limit_dev - 1 are
 * neither read nor written. */
int ceb_softmax_ce_bwd_rows(const void* logits_bf16, void* dlogits_bf16, int64_t ld, int64_t rows, int V,
                            const int64_t* labels, int64_t ignore_index, const float* lse, const float* grad_rows,
                            const int* row_limit_dev, void* stream);
/* LayerNorm over the last dimension: x fp32 [rows][E] -> y bf16; backward gives dx fp32 and
 * accumulates dgamma / dbeta. */
int ceb_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y_bf16, float* mean,
                      float* rstd, int64_t rows, int E, float eps, void* stream);
int ceb_layernorm_bwd(const void* dy_bf16, const float* x, const float* gamma, const float* mean, const float* rstd,
                      float* dx, float* dgamma, float* dbeta, int64_t rows, int E, void* stream);
/* Residual add fused with the LayerNorm that follows it: resid_out (fp32) = x (fp32) + branch (bf16),
 * y = LayerNorm(resid_out); the backward adds the gradient arriving through the residual stream (dres,
 * may be NULL) and also emits the total in bf16 for the branch (dx_bf16, may be NULL).
 * bwd: E in {128, 256, 512, 768, 1024, 1280}, 16-byte aligned tensors. */
int ceb_add_layernorm_fwd(const float* x, const void* branch_bf16, const float* gamma, const float* beta,
                          float* resid_out, void* y_bf16, float* mean, float* rstd, int64_t rows, int E, float eps,
                          void* stream);
int ceb_add_layernorm_bwd(const void* dy_bf16, const float* resid, const float* gamma, const float* mean,
                          const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma, float* dbeta,
                          int64_t rows, int E, void* stream);
/* tanh-GELU on bf16: out = gelu(x) (dy NULL) or dy * gelu'(x). */
int ceb_gelu_bf16(const void* x, const void* dy, void* out, int64_t n, void* stream);
/* Segment-wise BatchNorm (training mode) over NHWC bf16 activations x [N][L][C], C % 8 == 0: sample n
 * belongs to segment seg[n] (int64) and is normalised with the statistics of its own segment only
 * (per-client statistics inside a batched forward; no counterpart in the reference, which runs clients
 * one by one).  samples[s] = number of samples of segment s (fp32); sums = [2][S][C] fp32 scratch.
 * fwd writes y and mean / var (biased) / rstd [S][C]; bwd writes dx and leaves the per-segment sums of dy
 * and dy * xhat in sums[0] / sums[1] (dbeta / dgamma are their sums over segments). */
int ceb_segment_bn_fwd(const void* x_bf16, const int64_t* seg, const float* samples, int64_t N, int L, int C, int S,
                       const float* gamma, const float* beta, float eps, float* sums, float* mean, float* var,
                       float* rstd, void* y_bf16, void* stream);
int ceb_segment_bn_bwd(const void* dy_bf16, const void* x_bf16, const int64_t* seg, const float* samples, int64_t N,
                       int L, int C, int S, const float* gamma, const float* mean, const float* rstd, float* sums,
                       void* dx_bf16, void* stream);
/* The same with fp32 activations (tf32 / fp32 precision modes). */
int ceb_segment_bn_f32_fwd(const float* x, const int64_t* seg, const float* samples, int64_t N, int L, int C, int S,
                           const float* gamma, const float* beta, float eps, float* sums, float* mean, float* var,
                           float* rstd, float* y, void* stream);
int ceb_segment_bn_f32_bwd(const float* dy, const float* x, const int64_t* seg, const float* samples, int64_t N, int L,
                           int C, int S, const float* gamma, const float* mean, const float* rstd, float* sums, float* dx,
                           void* stream);
/* Classifier head + cross-entropy in one kernel: logits = scale * feat (bf16 [N][F]) @ W^T (fp32 [C][F]),
 * C <= 32, F % 64 == 0, F <= 2048.  g == NULL: forward, writes loss[N] and correct[N] (1.0 where the
 * arg-max equals the label).  g != NULL: backward of sum_s g[s] * loss[s]: writes dfeat (bf16 [N][F]) and
 * accumulates dW (fp32 [C][F], +=). */
int ceb_linear_ce_head(const void* feat_bf16, const float* W, float scale, const int64_t* labels, int N, int F, int C,
                       const float* g, float* loss, float* correct, void* dfeat_bf16, float* dW, void* stream);
/* out[c] (fp32) += sum over rows of x[r][c]; x bf16 [rows][ld], cols % 8 == 0, ld % 8 == 0, 16-byte
 * aligned base (bias gradients without an fp32 copy of the activation gradient). */
int ceb_colsum_bf16(const void* x_bf16, int64_t ld, int64_t rows, int cols, float* out, void* stream);

/* ---- peer-memory collectives over NVLink, fused with the server update ---------
 * (reference: the single NCCL reduce + host-side weight update/download,
 *  fed_worker.py:41,138, fed_aggregator.py:327-332,455).  src/dst/flags are HOST arrays of
 *  `world` device pointers, entry q being rank q's buffer mapped into this process
 *  (ceb_peer_open).  n % 4 == 0.  flags: 64 zeroed uint32 per rank; ticket: one zeroed
 *  uint32 (local).  epoch: 1, 2, 3, ... identical on every rank for the same call.
 *  Every rank must issue the same call; the kernel spins on its peers.
 *    mode 0  dst = sum_q src[q]                              (all-reduce)
 *    mode 1  Vv = rho*Vv + sum/N; Ve += Vv (error_mode 2) or Ve = Vv (1); dst = Ve (or Vv)
 *    mode 2  Vv = rho*Vv + sum/N; dst (= every replica's weights) -= lr * Vv
 *    mode 3  true_topk state: Vv = rho*Vv + sum/N; Ve += Vv on the shard, nothing gathered (pass dst = src)
 *  Barrier waits are bounded: after two minutes (globaltimer) without a peer the waiter writes a code into
 *  ticket[2] and goes on.
 *  Vv_shard / Ve_shard hold this rank's ceb_peer_shard_elems(n, world) entries. */
int64_t ceb_peer_shard_elems(int64_t n, int world);
int ceb_peer_fused_reduce(int mode, const void* const* src, const void* const* dst,
                          const void* const* flags, void* ticket, int rank, int world, unsigned int epoch,
                          int64_t n, float* Vv_shard, float* Ve_shard, float inv_n, float rho,
                          int error_mode, float lr, const float* lr_vec, void* stream);
/* The same collective through the NVSwitch (NVLS): mc_src / mc_dst are the MULTICAST addresses of the buffers whose
 * per-rank mappings are src[] / dst[] (memory bound to one multicast object spanning all `world` GPUs:
 * cuMulticastCreate / cuMulticastBindMem, or torch.distributed._symmetric_memory).  The reduce-scatter is one
 * multimem.ld_reduce.add.v4.f32 per 16 bytes (summed inside the switch), the all-gather one multimem.st; flags /
 * ticket / epoch and the modes are those of ceb_peer_fused_reduce (mode 3 needs no mc_dst). */
int ceb_peer_fused_reduce_nvls(int mode, const void* const* src, const void* const* dst,
                               const void* const* flags, void* ticket, int rank, int world, unsigned int epoch,
                               int64_t n, float* Vv_shard, float* Ve_shard, float inv_n, float rho,
                               int error_mode, float lr, const float* lr_vec, void* mc_src, void* mc_dst,
                               void* stream);
/* Symmetric buffers (the only allocating entry points: CUDA IPC needs whole cudaMalloc
 * allocations).  handle: ceb_peer_handle_bytes() (= 64) bytes, host memory. */
int ceb_peer_handle_bytes(void);
int ceb_peer_alloc(int64_t bytes, void** ptr_out, void* handle_out);
int ceb_peer_open(const void* handle, void** ptr_out);
int ceb_peer_release(void* ptr, int is_owner);

/* ---- TF32 tensor-core path (--precision tf32): fp32 operands read by tcgen05.mma.kind::tf32, fp32
 *      accumulation and fp32 outputs -- the arithmetic the reference's fp32 convolutions get from cuDNN
 *      (torch.backends.cudnn.allow_tf32 defaults to True; reference: fed_worker.py:20, no autocast) ------ */
/* C[M,N] (fp32) = act(A[M,K] * B[N,K]^T + bias); lda, ldb multiples of 4; out_mode 1: store, 2: atomic add. */
int ceb_gemm_tf32_tn_ld(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int64_t ldc, int M, int N,
                        int K, const float* bias, int relu, int out_mode, void* stream);
/* 3x3 / s1 / p1 convolution over fp32 NHWC / KRSC tensors, Cin % 32 == 0.  relu_mask (nullable): zero where
 * mask <= 0; resid (nullable): y = act(conv) + resid, or with y2 != NULL: y = act(conv), y2 = y + resid. */
int ceb_conv3x3_tf32_nhwc(const float* x, const float* w, float* y, int Nimg, int H, int W, int Cin, int Cout,
                          const float* bias, int relu, const float* relu_mask, const float* resid, float* y2,
                          void* stream);
int ceb_conv3x3_wgrad_tf32_nhwc(const float* dz, const float* x, float* dW, int Nimg, int H, int W, int Cin, int Cout,
                                void* stream);
/* C[M][N] (fp32) += A[K][M]^T * B[K][N], fp32 operands (MN-major, 32-byte-atom swizzle); N % 32 == 0. */
int ceb_gemm_tf32_mn_acc_ld(const float* A, int64_t lda, const float* B, int64_t ldb, float* C, int M, int N, int K,
                            int ldc, void* stream);
/* fp32 counterparts of the NHWC helpers: pooling windows k in {2,3,4} tile the map from the top-left corner
 * (floor mode), C % 4 == 0. */
int ceb_conv_weights_prepare_f32(int n_layers, const float* const* src, void* const* copy_f32, void* const* rot_f32,
                                 const int* cout, const int* cin, void* stream);
int ceb_maxpool_nhwc_f32(const float* y, float* out, int N, int H, int W, int C, int k, void* stream);
int ceb_maxpool_bwd_nhwc_f32(const float* y, const float* dp, float* dz, int N, int H, int W, int C, int k,
                             int relu_mask, void* stream);
int ceb_relu_bwd_f32(const float* y, const float* dy, float* dz, int64_t n, void* stream);
int ceb_linear_ce_head_f32(const float* feat, const float* W, float scale, const int64_t* labels, int N, int F, int C,
                           const float* g, float* loss, float* correct, float* dfeat, float* dW, void* stream);

/* ---- general convolutions (reference: models/fixup_resnet18.py:144-154 3x3 s2 / 1x1 s2; models/resnet9.py on
 *      the 28 x 28 FEMNIST input of cv_train.py:353-354) ------------------------------------------------- */
/* ks in {1,3}, stride in {1,2}, pad in {0,1}; any map whose OUTPUT width is <= 128 (whole output rows per tile:
 * widths 28 / 14 / 7 / 3 work).  f32 = 0: bf16 tensors, 1: fp32 tensors on the TF32 path. */
int ceb_conv2d_nhwc(const void* x, const void* w, void* y, int Nimg, int H, int W, int Cin, int Cout, int ks,
                    int stride, int pad, int relu, int f32, const void* relu_mask, const void* resid, void* y2,
                    void* stream);
/* 3x3 / pad 1 weight gradient with stride 1 or 2 over an INPUT map H x W; dz is [N, Ho, Wo, Cout]. */
int ceb_conv3x3_wgrad_nhwc(const void* dz, const void* x, float* dW, int Nimg, int H, int W, int Cin, int Cout,
                           int stride, int f32, void* stream);
/* x[N,H,W,Cin] fp32, Cin in {1,3} -> im2col rows of 9*Cin taps zero-padded to one 128-byte row
 * (out_f32 = 0: [N*H*W][64] bf16, 1: [N*H*W][32] fp32). */
int ceb_im2col3x3(const float* x_nhwc, void* A, int N, int H, int W, int Cin, int out_f32, void* stream);
/* sub = x[:, ::2, ::2, :];  up[:, ::2, ::2, :] = g, zero elsewhere.  row_bytes = C * sizeof(element) % 16 == 0. */
int ceb_subsample2_nhwc(const void* x, void* out, int N, int H, int W, int row_bytes, void* stream);
int ceb_upsample2_zero_nhwc(const void* g, void* up, int N, int H, int W, int Ho, int Wo, int row_bytes, void* stream);

/* ---- transformer kernels (reference: the GPT-2 block of pytorch_transformers that gpt2_train.py:88-99,262-273
 *      trains: matmul-softmax-dropout-matmul attention, nn.Embedding x3 + dropout, dropout + residual add) ---- */
/* out[B*T][E] (bf16) = dropout_p(softmax(Q K^T / 8 + causal)) V over the fused projection qkv[B*T][3E] (bf16),
 * head dim 64, T <= 1024; lse[B*H][T] (fp32) is kept for the backward.  The dropout mask is a pure function of
 * (*seed_dev, salt, batch, head, query, key): nothing is stored, the backward regenerates it. */
int ceb_attention_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int E, float p_drop,
                      const uint64_t* seed_dev, uint32_t salt, void* stream);
/* dqkv[B*T][3E] (bf16) from dout[B*T][E]; dq_scratch: fp32 [B*T][E], needed only when T > 128. */
int ceb_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                      float* dq_scratch, int B, int T, int H, int E, float p_drop, const uint64_t* seed_dev,
                      uint32_t salt, void* stream);
/* x[row][E] (fp32) = dropout_p(wte[ids[row]] + wpe[row % T] + (token_types ? wte[token_types[row]] : 0)). */
int ceb_embed_fwd(const int64_t* ids, const int64_t* token_types, const float* wte, const float* wpe, float* x,
                  int64_t rows, int T, int E, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream);
/* dwte / dwpe += scatter of dx (through the regenerated mask): 16-byte vector reductions. */
int ceb_embed_bwd(const int64_t* ids, const int64_t* token_types, const float* dx, float* dwte, float* dwpe,
                  int64_t rows, int T, int E, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream);
/* resid_out = x + dropout_p(branch); y = LayerNorm(resid_out) -- and the backward, in which only the branch's
 * gradient passes through the mask. */
int ceb_add_layernorm_drop_fwd(const float* x, const void* branch_bf16, const float* gamma, const float* beta,
                               float* resid_out, void* y_bf16, float* mean, float* rstd, int64_t rows, int E,
                               float eps, float p_drop, const uint64_t* seed_dev, uint32_t salt, void* stream);
int ceb_add_layernorm_drop_bwd(const void* dy_bf16, const float* resid, const float* gamma, const float* mean,
                               const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                               float* dbeta, int64_t rows, int E, float p_drop, const uint64_t* seed_dev,
                               uint32_t salt, void* stream);

/* ---- sharded server tail (reference: fed_aggregator.py:511-542 true_topk, :584-611 sketch -- there on one
 *      GPU over the full vector) ------------------------------------------------------------------------- */
/* Median estimates of coordinates [lo, hi) only (est[i - lo]); leaves the range's digit-1 histogram in scratch. */
int ceb_sketch_unsketch_est_range(const float* table, const int32_t* packed, int64_t n_tokens, const int32_t* blk,
                                  int num_blocks, int r, int64_t c, int64_t d, int64_t lo, int64_t hi, float* est,
                                  void* scratch, int64_t scratch_bytes, void* stream);
/* Exact top-k of a vector spread over `world` ranks (this rank: v[0..n_local), global index of v[0] =
 * idx_offset): per-pass histograms merged through peer memory, winners written into EVERY rank's lists.
 * exch / flags / out_idx / out_val: host arrays of `world` peer-mapped pointers (ceb_dsel_exch_bytes() bytes,
 * ceb_dsel_flag_words() uint32, k int32, k float each); err, epoch_dev: local device uint32 words (zero at
 * start; the epoch is advanced on the device, so a call captured in a CUDA graph stays valid on replay). */
int64_t ceb_dsel_exch_bytes(void);
int ceb_dsel_flag_words(void);
int ceb_dsel_run(const float* v, int64_t n_local, int64_t idx_offset, int64_t k, void* scratch, int64_t scratch_bytes,
                 const void* const* exch, const void* const* flags, const void* const* out_idx,
                 const void* const* out_val, void* err, int rank, int world, void* epoch_dev, int pass1_done,
                 void* stream);
/* Sharded true_topk apply: Vv / Ve are the shard [lo, hi) (indexed from lo); every rank updates its weights. */
int ceb_dense_apply_zero_shard(const int32_t* idx, const float* val, int64_t k, float* Vv_shard, float* Ve_shard,
                               int64_t lo, int64_t hi, float* w, float lr, const float* lr_vec, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* COMMEFFICIENT_B200_H */
