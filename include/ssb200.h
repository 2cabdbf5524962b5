/* libssb200 -- C ABI of the B200-native Silent-Speech clip-classification hot path.
 *
 * The reference (davdwan21/Silent-Speech) is pure Python/PyTorch and has no FFI; the
 * boundary it exposes for this path is the Python call contract of
 *   BiGRUClassifier.forward(X, lengths, R)      train_model_official.py:279-310, live_infer_official.py:124-138
 *   extract_feature(face, w, h, idxs, prev_xy)  live_infer_official.py:146-169, record_landmarks_official.py:62-100
 *   the train step                              train_model_official.py:433-439
 * This library is what a ctypes binding behind those calls loads (INTEGRATION.md shows it).
 *
 * Conventions
 *  - Every pointer is a raw DEVICE pointer unless the name ends in _host; sizes are explicit.
 *  - `stream` is a cudaStream_t; calls only enqueue work on it and return (asynchronous).
 *  - Return value: 0 = ok, non-zero = error; ssb_last_error() gives the thread-local message.
 *  - The caller owns every buffer (inputs, outputs, saved-for-backward, workspaces); the
 *    library allocates nothing persistent.
 *  - Built for sm_100a only.  There is no CPU fallback: without a B200 the calls fail.
 *  - All floating point is fp32; ROI pixels are uint8; lengths/labels int64; landmark ids int32.
 */
#ifndef SSB200_H_
#define SSB200_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

int ssb_version(void);
const char* ssb_last_error(void);
int ssb_device_info(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor);

/* K1: per-frame landmark features for a batch of clips.
 * Replaces extract_feature + mouth_width_px + the capture loop's distance gate / velocity reset
 * (live_infer_official.py:141-169, 275-296).
 * landmarks [B,T,n_lm,lm_stride] (x,y first), ids [K<=128], anchor_ids = {left corner 61,
 * right corner 291, upper inner 13, lower inner 14} as landmark ids, present [B,T] or NULL.
 * X [B,T,2K+4]; center [B,T,2], mouth_w [B,T], valid [B,T] (required when compact), lengths [B] may be NULL.
 * compact != 0 packs the used frames to the front of each clip (what the reference's bufX holds)
 * and writes their count to lengths; otherwise skipped frames give zero rows in place.
 * xy_px [B,T,K,2] or NULL: the gathered pixel positions (x*w, y*h) of the K landmarks, i.e. `xy` of
 * live_infer_official.py:147 -- the pure data-movement part of the feature, exported so that the gather can be checked
 * bit for bit (rows follow X: compacted or in place, zeros for skipped frames).
 * Velocity history: an undetected frame (present = 0) leaves it untouched, a detected frame outside the band resets it
 * (live_infer_official.py:272-296). */
int ssb_features_fwd(const float* landmarks, int64_t B, int64_t T, int64_t n_lm, int64_t lm_stride,
                     const int32_t* ids, int32_t K, const int32_t anchor_ids_host[4],
                     const uint8_t* present, float frame_w, float frame_h,
                     float gate_min_px, float gate_max_px, int32_t compact,
                     float* X, float* center, float* mouth_w, uint8_t* valid, int64_t* lengths,
                     float* xy_px, void* stream);

/* K2: fused mouth-ROI CNN forward (normalise, per-frame standardise, TinyROICNN).
 * Replaces train_model_official.py:286-294 (standardize=1) / live_infer_official.py:126-127 (0).
 * R [B,T,48,96] u8; weights_host = 8 device pointers {net.0.w,b, net.3.w,b, net.6.w,b, fc.w,b};
 * out[(b*T+t)*out_stride + 0..31] receives the embedding (zeros for t >= lengths[b]). */
int ssb_roi_cnn_fwd(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                    int32_t roi_h, int32_t roi_w, const float* const* weights_host,
                    int32_t standardize, float* out, int64_t out_stride, void* stream);

/* fp32 strided GEMM: C[i,j] (+)= sum_k A[i*sai+k*sak] * B[k*sbk+j*sbj] (+ bias[j]).
 * The ATen matmuls inside nn.GRU / nn.Linear of the reference (train_model_official.py:261-277)
 * and their backward.  lengths/T (optional) skip 64-row tiles made only of padded frames. */
int ssb_sgemm(const float* A, int64_t sai, int64_t sak, const float* B, int64_t sbk, int64_t sbj,
              const float* bias, float* C, int64_t ldc, int64_t M, int64_t N, int64_t K,
              int32_t accumulate, int32_t split_k, const int64_t* lengths, int64_t T, void* stream);

/* K3b: one bidirectional GRU layer's time recurrence on packed-sequence semantics
 * (nn.GRU, train_model_official.py:261-267, 301-305).  gi [B,T,gi_ld]: input projections incl.
 * b_ih, forward direction at column 0, reverse at 576.  out [B,T,384] (zeros past lengths).
 * gates (NULL for inference) [B,T,2,4,192] saves r, z, n, W_hn h + b_hn for the backward pass. */
int ssb_gru_layer_fwd(const float* gi, int64_t gi_ld, const int64_t* lengths, int64_t B, int64_t T,
                      int32_t hidden, const float* whh_f, const float* bhh_f,
                      const float* whh_r, const float* bhh_r, float* out, float* gates, void* stream);

/* K4: AttnPool + LayerNorm + Linear/ReLU/Dropout/Linear (+ softmax, top-k, label-smoothed CE).
 * Replaces train_model_official.py:240-248, 271-277, 405 and live_infer_official.py:223-226.
 * weights_host = 8 device pointers {pool.score.w,b, head.0.w,b, head.1.w,b, head.4.w,b};
 * saved_host = NULL or 6 device pointers {attn[B,T], pooled[B,384], xhat[B,384], rstd[B],
 * ln[B,384], hid[B,128]}.  probs, topk_*, y/loss may be NULL. loss[b] is the per-clip term. */
int ssb_head_fwd(const float* seq, const int64_t* lengths, int64_t B, int64_t T, int32_t dim,
                 const float* const* weights_host, int32_t num_classes, float* const* saved_host,
                 float* logits, float* probs, int32_t* topk_ids, float* topk_probs, int32_t k_top,
                 const int64_t* y, float label_smoothing, float* loss,
                 float dropout_p, uint64_t seed, void* stream);

/* K4-bwd part 1: gradient at the logits (fused CE when dlogits_in == NULL:
 * grad_scale * (softmax - smoothed one-hot)) and back through head.4 / Dropout / ReLU. */
int ssb_head_bwd_logits(const float* logits, const float* dlogits_in, const int64_t* y,
                        float label_smoothing, float grad_scale, const float* w2, const float* hid,
                        float dropout_p, int64_t B, int32_t num_classes,
                        float* dlogits, float* dhid, void* stream);

/* K4-bwd part 2: back through LayerNorm and AttnPool.  Overwrites dseq [B,T,384]; accumulates
 * into d_ln_g[384], d_ln_b[384], d_ws[384], d_bs[1]. */
int ssb_head_bwd_pool(const float* dln, const float* xhat, const float* rstd, const float* attn,
                      const float* seq, const int64_t* lengths, int64_t B, int64_t T,
                      const float* ln_g, const float* ws, float* dseq,
                      float* d_ln_g, float* d_ln_b, float* d_ws, float* d_bs, void* stream);

/* out[j] += sum_i A[i*lda + j]  (bias gradients). */
int ssb_colsum(const float* A, int64_t lda, int64_t M, int64_t N, float* out, void* stream);

/* K2-bwd: weight gradients of the fused ROI CNN (backward of train_model_official.py:209-229 as run by
 * loss.backward(), train:437).  saved = NULL: the forward is recomputed per frame in shared memory; saved = the records
 * of ssb_roi_cnn_fwd_tc3 for the same R / lengths / weights: the recompute is skipped.  grads_host = 8 device pointers
 * in the order of weights_host; they are accumulated into (+=). */
int ssb_roi_cnn_bwd(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                    int32_t roi_h, int32_t roi_w, const float* const* weights_host, int32_t standardize,
                    const float* d_out, int64_t d_out_stride, float* const* grads_host, const void* saved,
                    void* stream);

/* K3-bwd: back-propagation through time of one bidirectional GRU layer (ATen's GRU backward under
 * train:437).  dout/out [B,T,384]; gates from ssb_gru_layer_fwd; whhT_* = W_hh^T [192,576].
 * Writes dgi [B,T,1152] (gradient at the input projections), dghn [B,T,2,192] (n-gate part of the
 * hidden-projection gradient) and hprev [B,T,2,192]; dW_ih, dW_hh, biases and d(input) follow as GEMMs. */
int ssb_gru_layer_bwd(const float* dout, const float* out, const float* gates,
                      const float* whhT_f, const float* whhT_r, const int64_t* lengths,
                      int64_t B, int64_t T, int32_t hidden,
                      float* dgi, float* dghn, float* hprev, void* stream);

/* out[cols][rows] = in[rows][cols]^T. */
int ssb_transpose(const float* in, float* out, int64_t rows, int64_t cols, void* stream);

/* K5: joint-norm clip + Adam over a flat bucket (train_model_official.py:403, 438-439).
 * ssb_grad_sqnorm adds sum(g^2) to *sqnorm (double, zeroed by the caller once per step). */
int ssb_grad_sqnorm(const float* g, int64_t n, double* sqnorm, void* stream);
int ssb_clip_adam_step(float* p, float* g, float* m, float* v, int64_t n, const double* sqnorm,
                       float max_norm, float lr, float beta1, float beta2, float eps, int64_t step,
                       int32_t write_back_grad, float* norm_out, void* stream);

/* Counter-based dropout: y = x * keep(seed, i) / (1-p)  (nn.GRU inter-layer dropout, train:266). */
int ssb_dropout(const float* x, float* y, int64_t n, float p, uint64_t seed, void* stream);

/* K3a on the tensor cores: C[M,N] = [A1 | A2] . W[N, K1+K2]^T + bias with tcgen05.mma (fp16 operands
 * converted on the fly from fp32, fp32 accumulation in TMEM).  passes = 1: fp16; 3: split-fp16 hi/lo x3 (fp32 grade).
 * The matmuls inside nn.GRU (train_model_official.py:261-267); A2 may be NULL. */
int ssb_gemm_tc(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                const float* W, int64_t ldw, const float* bias, float* C, int64_t ldc,
                int64_t M, int64_t N, int32_t passes, int32_t accumulate,
                const int64_t* lengths, int64_t T, void* stream);

/* Backward GEMMs of nn.GRU on the tensor cores (autograd of train_model_official.py:261-267 under
 * loss.backward(), train:437): C[i,j] (+)= sum_k A[i*sai+k*sak] * B[k*sbk+j*sbj], the same strided contract as
 * ssb_sgemm (no bias).  Operands contiguous along their OUTPUT index (the weight gradient dW = dgi^T Z contracts
 * over frames) are transposed on the way into shared memory; split-fp16 x3, fp32 accumulation in TMEM.  a_amax
 * (nullable, device) = max|A| from ssb_absmax: A is scaled by a power of two into fp16's range (gate gradients are
 * far below it) and C un-scaled, without a host synchronisation.
 * A k-contiguous operand needs its other stride and K multiples of 4 floats and a 16-byte aligned base.
 * split_k > 1 requires accumulate = 1 (fp32 atomics).  lengths/T (optional): skip 128-row tiles of padded frames. */
int ssb_gemm_tcx(const float* A, int64_t sai, int64_t sak, const float* B, int64_t sbk, int64_t sbj,
                 float* C, int64_t ldc, int64_t M, int64_t N, int64_t K,
                 int32_t accumulate, int32_t split_k, const int64_t* lengths, int64_t T,
                 const float* a_amax, void* stream);
/* out[0] = max |x[i]| over n floats (device scalar; feeds ssb_gemm_tcx's a_amax). */
int ssb_absmax(const float* x, int64_t n, float* out, void* stream);

/* K3b on the tensor cores (cluster of 4 CTAs, W_hh resident in shared memory as fp16, tcgen05.mma per
 * step, h slices exchanged by bulk DSMEM copies).  Same semantics/layouts as ssb_gru_layer_fwd, except that
 * gi must already include b_ih + (b_hr, b_hz, 0) -- build that bias with ssb_gru_fold_bias and pass it to the
 * projection GEMM.  order (nullable) = clip index per batch slot (e.g. sorted by length). passes = 1 or 3.
 * gi_layout 0: gi columns [dir][gate][192] (nn.GRU's row order of weight_ih); gi_layout 1 (passes = 1, gates = NULL):
 * GRU-blocked columns [dir][rank][gate][48] written through ssb_proj_prepack / ssb_gru_fold_bias with gru_blocked = 1,
 * staged per step with one 576-byte bulk copy per clip; gi_layout 2: the same column order stored as fp16
 * (ssb_proj_gemm out_half = 1, gi_ld in fp16 elements), 288-byte copies into two staging buffers, two steps ahead. */
int ssb_gru_fold_bias(const float* bih_f, const float* bhh_f, const float* bih_r, const float* bhh_r,
                      float* out1152, int32_t gru_blocked, void* stream);
int ssb_gru_layer_fwd_tc(const void* gi, int64_t gi_ld, const int64_t* lengths, const int32_t* order,
                         int64_t B, int64_t T, int32_t hidden,
                         const float* whh_f, const float* bhh_f, const float* whh_r, const float* bhh_r,
                         float* out, float* gates, int32_t passes, int32_t gi_layout, void* stream);

/* K2 on the tensor cores: same contract as ssb_roi_cnn_fwd (train_model_official.py:286-294, 209-229) plus `passes`:
 *   1 = single fp16 pass (stated tolerance 1e-3 on the logits);
 *   2 = weights split hi + lo, activations rounded to fp16 once (x.W_hi + x.W_lo; conv1's raw-pixel input is exact):
 *       the inference form of the fp32-parity mode -- embedding within 3e-5, logits unchanged at the 1e-6 level;
 *   3 = both operands split (fp32-parity grade everywhere; the training forward, the only mode that fills `saved`).
 * All three convolutions are tcgen05 GEMMs with the 2x2 max-pools folded into N; frames are staged by one bulk async copy
 * each, handed out by an atomic counter; four 128-thread frame slots per CTA (passes 1, 2) or two 256-thread slots
 * (passes 3) with one MMA issuer each; hi | lo weights stacked along N.   */
int ssb_roi_cnn_fwd_tc3(const uint8_t* R, const int64_t* lengths, int64_t B, int64_t T,
                        int32_t roi_h, int32_t roi_w, const float* const* weights_host, int32_t standardize,
                        float* out, int64_t out_stride, int32_t passes, void* saved, void* stream);
/* Training: bytes per frame of the record ssb_roi_cnn_fwd_tc3(passes = 3, saved != NULL) writes for every valid frame
 * (saved = B*T records): pooled conv1 / conv2 activations (fp32), both max-pool argmax maps, the conv3 ReLU mask and the
 * channel means -- what ssb_roi_cnn_bwd would otherwise recompute. */
int64_t ssb_roi_cnn_saved_bytes(void);

/* K3a v2 (fp16 single-pass mode): the input projections of BOTH directions as one persistent GEMM,
 * C[M, N] = [A1 | A2] . [W_a ; W_b]^T + bias (nn.GRU's x W_ih^T + b_ih, train_model_official.py:261-267, forward and
 * reverse direction stacked: N = 2 x 576).  ssb_proj_prepack converts the weights once into fp16 blocks that the GEMM
 * streams with bulk async copies (a derived cache of the state-dict tensors; rebuild when they change);
 * `packed` is caller-owned, ssb_proj_packed_bytes(N, K1, K2) bytes.  Rows of padded frames are not written.
 * gru_blocked = 1 permutes the output columns from [dir][gate][192] to [dir][rank 0..3][gate][48], the order in which
 * one CTA of the recurrence cluster reads them (ssb_gru_layer_fwd_tc gi_layout = 1); the bias must be in the same order
 * (ssb_gru_fold_bias gru_blocked = 1). */
int64_t ssb_proj_packed_bytes(int64_t N, int64_t K1, int64_t K2);
int ssb_proj_prepack(const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int64_t K1, int64_t K2,
                     int32_t gru_blocked, void* packed, void* stream);
/* out_half = 1: C is fp16 (ldc counts fp16 elements; N, ldc multiples of 8) -- gi for ssb_gru_layer_fwd_tc gi_layout = 2. */
int ssb_proj_gemm(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                  const void* packed, const float* bias, void* C, int64_t ldc, int64_t M, int64_t N,
                  const int64_t* lengths, int64_t T, int32_t out_half, void* stream);
/* K3a v2, fp32-parity grade (the default precision of BiGRUClassifier, the training mode): the same persistent GEMM with
 * both operands split x = hi + lo (fp16 each) and D += W_hi.A_hi + W_lo.A_hi + W_hi.A_lo.  ssb_proj_prepack_split writes
 * every 16 KB weight block followed by the block of its fp16 residuals (ssb_proj_packed_bytes_split bytes); the kernel
 * keeps a 64-row activation tile (hi and lo) resident, uses the weights as the M = 128 operand and stores the transposed
 * accumulator as full 128-byte row segments.  C is fp32. */
int64_t ssb_proj_packed_bytes_split(int64_t N, int64_t K1, int64_t K2);
int ssb_proj_prepack_split(const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int64_t K1, int64_t K2,
                           int32_t gru_blocked, void* packed, void* stream);
int ssb_proj_gemm_split(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                        const void* packed, const float* bias, float* C, int64_t ldc, int64_t M, int64_t N,
                        const int64_t* lengths, int64_t T, void* stream);
/* Inference form of the split projection (same call sites, train_model_official.py:261-267 under torch.no_grad):
 * the packed weights of ssb_proj_prepack_split, activations rounded to fp16 once, D += W_hi.A + W_lo.A on 128-row tiles
 * (two N = 128 MMAs per k-step instead of three N = 64 ones per 64 rows).  Same arguments as ssb_proj_gemm_split. */
int ssb_proj_gemm_wsplit(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                         const void* packed, const float* bias, float* C, int64_t ldc, int64_t M, int64_t N,
                         const int64_t* lengths, int64_t T, void* stream);

/* K3a v2 without any allocation: ssb_proj_prepack + ssb_proj_gemm in one call, the packed-weight workspace being a
 * file-scope static device array of the library instead of caller-owned memory.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  The call performs no cudaMalloc / cudaFree / allocating library call.
 * Limit: the packed weights must fit ssb_proj_static_max_bytes() = 884,736 bytes, i.e. N = (W_b ? 2 : 1) * n_each <= 1152
 * and round_up(K1, 8) + round_up(K2, 8) <= 384 (rounded up to 64); larger shapes are rejected with SSB_ERR_ARG and a
 * message.  Every GPU that loads the library holds the array until exit.  All calls on a device share it and it keeps
 * the last call's contents (every byte the GEMM reads is rewritten by the call's own pack kernel first, so nothing needs
 * clearing); calls that overlap on the GPU race on it -- serialise them on one stream.
 * W_a / W_b fp32 [n_each, K1 + K2] with row stride ldw (W_b nullable); every other argument as for ssb_proj_gemm. */
int64_t ssb_proj_static_max_bytes(void);
int ssb_proj_gemm_static(const float* A1, int64_t lda1, int64_t K1, const float* A2, int64_t lda2, int64_t K2,
                         const float* W_a, const float* W_b, int64_t ldw, int64_t n_each, int32_t gru_blocked,
                         const float* bias, void* C, int64_t ldc, int64_t M,
                         const int64_t* lengths, int64_t T, int32_t out_half, void* stream);

/* K6: batched mouth-ROI crop + BGR->gray + resize to 96x48, bit-exact with OpenCV 4.13.
 * Replaces crop_roi_gray (live_infer_official.py:171-187; flavor 0: INTER_AREA, box in double arithmetic) and
 * crop_roi (record_landmarks_official.py:102-118; flavor 1: INTER_LINEAR, box in float32 arithmetic).
 * frames u8 [n_frames,Hf,Wf,3] BGR; frame_index i32 [N] or NULL (crop n reads frame n); centers f32 [N,2];
 * mouth_w f64 [N] (4th return value of extract_feature); out u8 [N,48,96], zeros where the reference returns
 * None; ok u8 [N] or NULL. */
int ssb_roi_crop(const uint8_t* frames, const int32_t* frame_index, int64_t Hf, int64_t Wf,
                 const float* centers, const double* mouth_w, int64_t N, int32_t flavor,
                 uint8_t* out, uint8_t* ok, void* stream);

/* Sliding-window stream state (windowing policy of inactive/live_feed.py:155-213 driving the official model).
 * ring f32 [S,W,D]: per-stream ring of rows [point features | ROI-CNN embedding].  ssb_ring_push writes row i
 * (x [n,dx], e [n,de] or NULL) at ring[sid[i]][pos[i]]; ssb_ring_gather builds the time-ordered, zero-padded
 * windows Xw [m,T,dx], Ew [m,T,de] (nullable) from rows start[i].. (mod W) and writes lengths[i] = min(count[i],T).
 * Head / count bookkeeping is the caller's (host integers). */
int ssb_ring_push(float* ring, int64_t W, int64_t D, const int32_t* sid, const int32_t* pos,
                  const float* x, int64_t dx, const float* e, int64_t de, int64_t n, void* stream);
int ssb_ring_gather(const float* ring, int64_t W, int64_t D, const int32_t* sid, const int32_t* start,
                    const int32_t* count, int64_t m, float* Xw, int64_t dx, float* Ew, int64_t de,
                    int64_t T, int64_t* lengths, void* stream);

/* K7: GRU recurrence for small hidden sizes (H = 32/64/96/128, W_hh resident in one SM) with the time-mean pooling fused:
 * the earlier model generation GRUClassifier (inactive/train_reduced.py:129-145, weights inactive/word_model_5.pt).
 * gi f32 [B,T,gi_ld] = x W_ih^T + b_ih, direction d in columns [d*3H,(d+1)*3H), gates (r,z,n); lengths NULL = every clip
 * runs all T steps (the reference's un-packed call); out f32 [B,T,ndir*H] or NULL; pooled f32 [B,ndir*H] or NULL. */
int ssb_gru_small_fwd(const float* gi, int64_t gi_ld, const float* whh_f, const float* bhh_f,
                      const float* whh_r, const float* bhh_r, const int64_t* lengths,
                      int64_t B, int64_t T, int64_t H, int32_t ndir, float* out, float* pooled, void* stream);

/* Mouth-open EMA + hysteresis of important_landmarks.py:130-144 for S independent streams x F frames, in double.
 * lm f32 [S,F,n_lm,3]; valid u8 [S,F] or NULL (frame without a face: state untouched); ema_state f64 [S] / open_state
 * u8 [S] in-out; ema_out f64 [S,F] / open_out u8 [S,F] nullable. */
int ssb_mouth_open_ema(const float* lm, const uint8_t* valid, int64_t S, int64_t F, int64_t n_lm,
                       double alpha, double open_thr, double close_thr, double* ema_state, uint8_t* open_state,
                       double* ema_out, uint8_t* open_out, void* stream);

/* Device-resident clip store (".npz ingestion"): dst[i,:] = src[index[i],:], rows of row_bytes.
 * Replaces the per-item np.load + clip_pad_trim / roi_pad_trim + torch.stack of NPZWordDataset.__getitem__ and
 * collate_fn (train_model_official.py:93-117, 134-204) for clips ingested once into HBM.  index i64 [n] on the device,
 * every value in [0, rows of src). */
int ssb_gather_rows(const void* src, int64_t row_bytes, const int64_t* index, int64_t n, void* dst, void* stream);

/* Training augmentation of the point features on the device: the policy of NPZWordDataset.__getitem__
 * (train_model_official.py:143-152): N(0, noise_std) noise with probability p_noise; if T > 12, with probability p_drop
 * 1..drop_max frames from [1, T-2] are removed from X; the ROI clip is only cut to the new length (the reference does
 * not thin it).  Counter-based RNG in (seed, clip); decisions i32 [B,4] = {noise, k, drop_0, drop_1} (nullable).
 * Xo must not alias X; R (nullable) is modified in place. */
int ssb_augment_clips(const float* X, const int64_t* lengths, int64_t B, int64_t T, int64_t D,
                      uint8_t* R, int64_t roi_frame_bytes, uint64_t seed, float p_noise, float noise_std,
                      float p_drop, int32_t drop_max, float* Xo, int64_t* lengths_out, int32_t* decisions,
                      void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSB200_H_ */
