/*
 * lipvq.h -- C ABI of the B200-native LipVQ-VAE action-tokenizer kernels (liblipvq.so).
 *
 * This is the drop-in boundary underneath the reference's PyTorch module API.  The reference
 * (Alishah-Naushad/Robot-Manipulation-VQ-VAE) has no FFI of its own: its "operator interface"
 * for this path is the nn.Module ctor/forward/state_dict surface.  Every entry
 * point below replaces a group of torch ops of the reference and cites the file:line it stands
 * in for (paths relative to the reference checkout).  The Python shim that keeps the module API
 * (robot-manipulation-vq-vae_b200/lipvq/) binds these symbols with ctypes; INTEGRATION.md shows
 * the stub a reference maintainer would add.
 *
 * Conventions
 *   - plain pointers + sizes; no torch / C++ types.  All pointers are DEVICE pointers unless
 *     the name ends in _host.  Row-major, contiguous, fp32 unless stated.  Indices are int64
 *     (what torch.argmin returns).
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*), performs no
 *     allocation and no host synchronisation, and is CUDA-graph capturable.  Scratch memory is
 *     supplied by the caller (see *_workspace_bytes).
 *   - return value: LIPVQ_OK or an error code; lipvq_last_error() gives the message of the
 *     last failing call on the calling thread.
 *   - sm_100a only.  There is no CPU path in this library.
 */
#ifndef LIPVQ_H_
#define LIPVQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LIPVQ_ABI_VERSION 9
#define LIPVQ_API __attribute__((visibility("default")))

enum {
  LIPVQ_OK = 0,
  LIPVQ_ERR_INVALID = 1, /* bad argument (shape, alignment, null pointer, workspace too small) */
  LIPVQ_ERR_CUDA = 2,    /* a CUDA runtime/driver call or kernel launch failed */
  LIPVQ_ERR_UNSUPPORTED = 3
};

/* nearest-code metric */
enum {
  LIPVQ_METRIC_L2 = 0,     /* argmin_k sqrt(sum_d (z-e)^2): LFQQuantizer, backbone_lfqvae_v5.py:39-46 */
  LIPVQ_METRIC_L2SQ = 1,   /* argmin_k sum_d (z-e)^2 (no sqrt): VQVAE.quantize, backbone.py:57-63 */
  LIPVQ_METRIC_DOT = 2     /* argmax_k <z,e> on caller-normalised rows: heirarchical_vqvae.py:441-449 */
};

/* search algorithm */
enum {
  LIPVQ_ALGO_AUTO = 0,
  LIPVQ_ALGO_SIMT_F32 = 1, /* shared-memory staged fp32 FFMA tiles, warp-shuffle argmin */
  LIPVQ_ALGO_TCGEN05 = 2   /* fused tcgen05/TMEM search: fp16 pass + 3-term fp16 pass, argmin + certification in the epilogue */
};

/* activations understood by the fused linear kernels */
enum {
  LIPVQ_ACT_NONE = 0,
  LIPVQ_ACT_RELU = 1,
  LIPVQ_ACT_GELU = 2,    /* exact erf GELU (nn.GELU() default) */
  LIPVQ_ACT_SIGMOID = 3,
  LIPVQ_ACT_GIVEN = 4    /* lipvq_linear_bwd_x_f32 only: X already holds act'(pre-activation), as written by
                            lipvq_linear_fwd_dual_f32 */
};

/* EMA codebook normalisation variants */
enum {
  LIPVQ_EMA_SIMPLE = 0,   /* codebook = sum / (size + eps)                      hq_vae.py:277-279 */
  LIPVQ_EMA_LAPLACE = 1,  /* codebook = sum / ((size+eps)/(sum(size)+K*eps))    hq_vae.py:115-118 */
  LIPVQ_EMA_NORMALIZE = 2 /* codebook = l2normalize(sum / (size + eps))         heirarchical_vqvae.py:533-541 */
};

LIPVQ_API int lipvq_abi_version(void);
LIPVQ_API const char* lipvq_last_error(void);
/* number of kernels this library has launched from the calling process since load (all threads) */
LIPVQ_API uint64_t lipvq_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * K3  fused distance + argmin (nearest code).  Replaces the [N,K,D] broadcast, torch.norm and
 * torch.argmin of LFQQuantizer.forward (backbone_lfqvae_v5.py:39-46; dup hq_vae.py:92-97,
 * 247-255), the squared form of VQVAE.quantize (backbone.py:57-63) and the matmul+argmax of
 * HierarchicalVQVAE._compute_similarities/quantize_* (heirarchical_vqvae.py:421-449).
 *
 * Never materialises [N,K].  Scores ||e-mu||^2 - 2 (z-mu).(e-mu) are produced by the chosen
 * contraction with a per-row running top-2; a row is decided only when its two best scores
 * are further apart than a rigorous error bound of that contraction, otherwise it moves to the
 * next precision level: fp16 tcgen05 MMAs (fp32 accumulation in TMEM) over every row, a
 * three-term fp16 split over the rows left, and finally an exact fp32 direct-difference scan
 * (sqrt applied for METRIC_L2, first-min tie rule).  The fp32 FFMA path has two levels (FFMA
 * tiles, exact scan).  Indices therefore equal the reference's except where the reference's
 * own fp32 rounding decides (relative distance gap < 1e-6).  Codebooks that stay resident in shared memory (every
 * tokenizer codebook of the reference) take ONE launch: the fp16 pass settles the rows it cannot certify itself -- a
 * candidate filter on the CUDA cores from the resident fp16 tiles, then the exact fp32 evaluation of the candidates.
 *
 *   z [N,D], codebook [K,D] fp32; idx_out [N] int64; best_out [N] fp32 or NULL (exact winning
 *   distance / squared distance / similarity, as the metric defines).
 *   stats_out: NULL or 4 x int32 on device: {rows decided by the exact scan, rows the first
 *   (fastest) level could not certify, and -- when the single-launch pass settles its own rows (codebooks resident in
 *   shared memory) -- codes evaluated exactly for those rows, rows for which every code was evaluated; else 0, 0}.
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API size_t lipvq_nn_search_workspace_bytes(int64_t N, int D, int K, int metric, int algo);
LIPVQ_API int lipvq_nn_search_f32(const float* z, int64_t N, int D, const float* codebook, int K, int metric, int algo,
                        int64_t* idx_out, float* best_out, int32_t* stats_out, void* workspace,
                        size_t workspace_bytes, void* stream);
/* which algorithm AUTO resolves to for a shape (host-side, no launch) */
LIPVQ_API int lipvq_nn_search_pick_algo(int64_t N, int D, int K, int metric);

/* The same search split in two for callers whose codebook changes less often than they search it (bulk tokenisation:
 * LFQQuantizer.forward under torch.no_grad with fixed weights, backbone_lfqvae_v5.py:39-46; a training step: the
 * codebook changes once per step, hq_vae.py:277-279, and is prepared right after that update, off the search's
 * critical path).  prepare: everything that depends on the codebook only -- centre, centred fp16 code tiles in the
 * tensor core's shared-memory layout, the fp16 split of the code norms, the maxima the certification margin uses --
 * is written into the workspace (one small launch).  prepared: ONE launch per search; it bulk-copies that image,
 * runs the fp16 tcgen05 pass over all rows and settles the rows it cannot certify itself (exact fp32 scan, same
 * arithmetic and tie rule as lipvq_nn_search_f32), so results are identical to lipvq_nn_search_f32's.
 *   center_rows [n_center_rows, D] or NULL: rows whose mean becomes the centre (a batch of latents of the kind that
 *   will be searched; <= 32 evenly spaced rows are read).  NULL: the codebook's own rows.  Any centre is valid.
 *   The workspace is lipvq_nn_search_workspace_bytes(N, D, K, metric, LIPVQ_ALGO_TCGEN05) bytes for the largest N
 *   searched; the prepared state sits at an offset that does not depend on N.  lipvq_nn_search_f32 on the same
 *   workspace leaves the prepared state intact.
 *   CONTRACT: `codebook` must hold the values it held when prepare ran (it is re-read only for near-ties); searching
 *   a workspace that was never prepared for (K, D, metric) is detected on the device and costs an exact scan of every
 *   row, not a wrong result.  supported(): 1 when the shape takes this path (codebook resident in shared memory and
 *   N * K large enough for the tensor path), else 0 -- use lipvq_nn_search_f32. */
LIPVQ_API int lipvq_nn_search_prepared_supported(int64_t N, int D, int K, int metric);
LIPVQ_API int lipvq_nn_search_prepare_f32(const float* codebook, int K, int D, int metric, const float* center_rows,
                        int64_t n_center_rows, void* workspace, size_t workspace_bytes, void* stream);
LIPVQ_API int lipvq_nn_search_prepared_f32(const float* z, int64_t N, int D, const float* codebook, int K, int metric,
                        int64_t* idx_out, int32_t* stats_out, void* workspace, size_t workspace_bytes, void* stream);

/* row-wise F.normalize(x, dim=-1, eps=1e-12) (heirarchical_vqvae.py:441-445). y may alias x. */
LIPVQ_API int lipvq_l2_normalize_rows_f32(const float* x, int64_t rows, int D, float* y, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K4  gather + straight-through + commitment/codebook loss.  Replaces codebook[indices],
 * z + (q - z).detach(), the two mse_loss calls and z_q.clone().detach()
 * (backbone_lfqvae_v5.py:47,74,80-81; backbone.py:66-74; heirarchical_vqvae.py:450-459).
 *
 *   z_q_out   [N,D]: codebook[idx] (bit-exact copy of the code rows)
 *   z_st_out  [N,D] or NULL: z + (z_q - z) evaluated in fp32 exactly as the reference's STE
 *   sqerr_sum (double*, 1 element, ACCUMULATED INTO): sum over N*D of (z_q - z)^2
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_gather_ste_loss_fwd_f32(const float* z, const int64_t* idx, const float* codebook, int64_t N, int D, int K,
                                  float* z_q_out, float* z_st_out, double* sqerr_sum, void* stream);

/* backward of the two MSE terms + gather:
 *   dz_out[n,:]  = w_commit * 2/(N*D) * (z - z_q) * gscale            (+ dzq_up[n,:] if ste != 0)
 *   dcodebook[idx[n],:] += w_codebook * 2/(N*D) * (z_q - z) * gscale  (+ dzq_up[n,:] if ste == 0)
 * gscale is read from device (*gscale_dev, the upstream d loss) or 1 if NULL.  dzq_up (gradient
 * arriving at z_q / z_st from the decoder) may be NULL.  If sigmoid_out != NULL the result is
 * multiplied by y(1-y) with y = sigmoid_out (= z), i.e. dz_out is the gradient at the Lipschitz
 * layer's pre-activation (fuses torch.sigmoid's backward, backbone_lfqvae_v5.py:24). */
LIPVQ_API int lipvq_gather_ste_loss_bwd_f32(const float* z, const float* z_q, const int64_t* idx, const float* dzq_up,
                                  const float* gscale_dev, float w_commit, float w_codebook, int ste,
                                  const float* sigmoid_out, int64_t N, int D, int K, float* dz_out,
                                  float* dcodebook, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K5  EMA codebook statistics and update.  Replaces F.one_hot, .sum(0), one_hot^T @ z and the
 * mul_/add_ EMA + normalise chain (hq_vae.py:103-118, 262-279; heirarchical_vqvae.py:498-541).
 * counts [K] fp32 and sums [K,D] fp32 are ACCUMULATED INTO (zero them first; in data-parallel
 * runs they are all-reduced between the two calls).
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_ema_stats_f32(const float* z, const int64_t* idx, int64_t N, int D, int K, float* counts, float* sums,
                        void* stream);
/* Deterministic variants of the two per-code sums above (scatter_det.cu).  lipvq_code_sort_i64 puts the tokens in code
 * order with a STABLE counting sort (integer work only); the *_sorted calls then add every code's rows in ascending token
 * order, one writer per output row: bit-identical results run to run, no floating-point atomics.  The legacy calls above
 * accumulate with atomics (sums correct to rounding, order varies).  Semantics are otherwise the same (ACCUMULATE into
 * counts / sums / dcodebook; out-of-range codes are clamped to [0, K)).
 * workspace: lipvq_code_sort_workspace_bytes(N, D, K) bytes of device memory, 16-byte aligned, owned by the caller; it
 * must be passed unchanged from the sort to the sums, and re-sorted whenever idx changes.  Shapes: 1 <= N < 2^31,
 * D % 4 == 0, 4 <= D <= 128, 1 <= K <= 8192 (workspace_bytes returns 0 otherwise).  Replaces the same reference lines as
 * lipvq_ema_stats_f32 / lipvq_gather_ste_loss_bwd_f32 (hq_vae.py:262-275; algo/heirarchical_vqvae.py:518-530). */
LIPVQ_API size_t lipvq_code_sort_workspace_bytes(int64_t N, int D, int K);
LIPVQ_API int lipvq_code_sort_i64(const int64_t* idx, int64_t N, int D, int K, void* workspace, size_t workspace_bytes,
                                  void* stream);
LIPVQ_API int lipvq_ema_stats_sorted_f32(const float* z, int64_t N, int D, int K, void* workspace, size_t workspace_bytes,
                                         float* counts, float* sums, void* stream);
LIPVQ_API int lipvq_gather_ste_loss_bwd_sorted_f32(const float* z, const float* z_q, const float* dzq_up,
                                                   const float* gscale_dev, float w_commit, float w_codebook, int ste,
                                                   const float* sigmoid_out, int64_t N, int D, int K, void* workspace,
                                                   size_t workspace_bytes, float* dz_out, float* dcodebook, void* stream);
/* Allocation-free variant of code_sort + ema_stats_sorted.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  One call sorts idx and ACCUMULATES counts[k] += #tokens of code k, sums[k,:] += their z rows in
 * ascending token order (bit-identical to the two-call sequence above).  Its scratch is a file-scope __device__ array of
 * lipvq_ema_stats_static_max_bytes() = 8 MiB inside the library (held by every GPU that loads the library until exit);
 * the call performs no cudaMalloc / cudaFree / cudaMallocAsync, only kernel launches.  Shapes as for lipvq_code_sort_i64
 * AND lipvq_code_sort_workspace_bytes(N, D, K) <= 8 MiB (e.g. N = 65536, D = 64, K = 512 needs 2.9 MB), otherwise
 * LIPVQ_ERR_INVALID with the sizes in lipvq_last_error().  Every call on a device shares the one array, which keeps the
 * last call's contents (nothing in it needs to start at zero): calls that overlap on the GPU -- two streams, or two
 * threads on one device -- race; serialise them. */
LIPVQ_API size_t lipvq_ema_stats_static_max_bytes(void);
LIPVQ_API int lipvq_ema_stats_static_f32(const float* z, const int64_t* idx, int64_t N, int D, int K, float* counts,
                                         float* sums, void* stream);
/* size <- decay*size + (1-decay)*counts ; ema_sum <- decay*ema_sum + (1-decay)*sums ;
 * codebook <- variant(ema_sum, size).  one_minus_decay is passed separately so the host can form
 * it in double like the reference does (1 - self.decay).  scratch: >= 8 bytes device. */
LIPVQ_API int lipvq_ema_apply_f32(const float* counts, const float* sums, int K, int D, float decay, float one_minus_decay,
                        float eps, int variant, float* ema_size, float* ema_sum, float* codebook, void* scratch,
                        void* stream);
/* usage bookkeeping of LFQQuantizerEMA_KMeans (hq_vae.py:123-130):
 * usage_counts += counts; usage_ma = .99*usage_ma + .01*(counts>0); entropy_ma (1 elem) EMA of
 * -(p log(p+1e-8)).sum(), p = counts/(counts.sum()+1e-8). */
LIPVQ_API int lipvq_usage_stats_f32(const float* counts, int K, float* usage_counts, float* usage_ma, float* entropy_ma,
                          void* stream);
/* codebook-health summary (heirarchical_vqvae.py:687-706): out[0]=#(size>0), out[1]=#(size<dead_thr)
 * as int32 on device -- one D2H copy instead of ~8 .item() syncs. */
LIPVQ_API int lipvq_codebook_usage_f32(const float* cluster_size, int K, float dead_thr, int32_t* out2, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K1/K2/K7  fused linear layers (encoder / Lipschitz / decoder MLPs).
 * Replaces nn.Linear+nn.GELU chains (backbone_lfqvae_v5.py:54-68, 71, 75-76), nn.Linear+ReLU
 * (backbone.py:17-32), normalization()+LipschitzMLP.forward (backbone_lfqvae_v5.py:6-24).
 *
 * A layer may apply an activation to its INPUT while loading it (act_in) and/or to its output in the epilogue
 * (act_out).  The training step uses the producing side: lipvq_linear_fwd_dual_f32 stores act(pre) AND act'(pre) from
 * one evaluation, every consumer then reads the stored values (act_in = NONE / LIPVQ_ACT_GIVEN).  The consumer side
 * (act_in = GELU/RELU on a stored pre-activation) remains for stand-alone module calls.
 *
 *   fwd:   Y[n,o] = act_out( sum_i act_in(X[n,i]) * W[o,i] + bias[o] )           X:[N,in] W:[out,in]
 *   bwd_x: dX[n,i] = ( sum_o dY[n,o] * W[o,i] ) * act_in'(X[n,i])                 (gradient at X, the stored pre-activation)
 *   bwd_w: dW[o,i] += sum_n dY[n,o] * act_in(X[n,i]) ;  db[o] += sum_n dY[n,o]    (ACCUMULATED INTO; db may be NULL)
 *
 * Two implementations behind each entry point, chosen per shape (LIPVQ_LINEAR_ALGO=simt forces the second):
 * tcgen05 MMAs on three-way bf16 splits of both operands (six chains, fp32 accumulation in TMEM; as
 * accurate as fp32 FFMA) for N >= 4096 tokens and dims <= 128, fp32 FFMA tiles otherwise.
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_linear_fwd_f32(const float* X, int64_t N, int in_dim, int act_in, const float* W, const float* bias,
                         int out_dim, int act_out, float* Y, void* stream);
/* Training-time forward: Y = act_out(pre) AND dYdpre = act_out'(pre), pre = act_in(X) W^T + b, from one evaluation of
 * the activation in the producing layer's epilogue.  Consumers then read Y with act_in = NONE and the backward pass
 * multiplies by dYdpre (act_in = LIPVQ_ACT_GIVEN) instead of re-evaluating erf per consumer.  Same arithmetic as
 * nn.Linear -> nn.GELU and its autograd (backbone_lfqvae_v5.py:54-68). */
LIPVQ_API int lipvq_linear_fwd_dual_f32(const float* X, int64_t N, int in_dim, int act_in, const float* W, const float* bias,
                              int out_dim, int act_out, float* Y, float* dYdpre, void* stream);

/* Three linear layers in ONE launch, for inference (no stored activations):
 *   Y = act3(act2(act1(X W1^T + b1) W2^T + b2) W3^T + b3),  X [N,d0], Wl [dl, d(l-1)] row-major, bl [dl] or NULL.
 * The tokenizer's encoder under torch.no_grad -- nn.Linear + GELU, nn.Linear + GELU, the Lipschitz layer's clamped weight
 * + sigmoid (backbone_lfqvae_v5.py:54-62, 70-76: self.encoder, self.to_latent) -- and, with other widths, its decoder
 * (:63-68).  Same arithmetic as three lipvq_linear_fwd_f32 calls on the tensor path (three-way bf16 splits, six MMA
 * chains, fp32 accumulation, fp32 bias + activation); hidden activations stay in shared memory / TMEM.
 * supported(): 1 when the widths fit (weights of all layers resident in shared memory, N >= 4096), else 0 -- chain three
 * lipvq_linear_fwd_f32 calls. */
/* lipvq_mlp3_fwd_train_f32: the same chain as the forward of a TRAINING step -- the two hidden layers also store
 * their activation H_l = act_l(pre_l) [N, d_l] and its derivative G_l = act_l'(pre_l) (what two
 * lipvq_linear_fwd_dual_f32 calls store; consumed by lipvq_linear_bwd_x_f32 with LIPVQ_ACT_GIVEN and by
 * lipvq_linear_bwd_w_f32).  Replaces the encoder / decoder forward of LLFQVAE_V4.forward with autograd's saved
 * activations (backbone_lfqvae_v5.py:70-84). */
LIPVQ_API int lipvq_mlp3_fwd_train_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1,
                        const float* W2, const float* b2, int d2, int act2, const float* W3, const float* b3, int d3,
                        int act3, float* Y, float* H1, float* G1, float* H2, float* G2, void* stream);
/* lipvq_mlp3_fwd_bf16_f32: the bf16 mode of the same chain (north_star's reduced-precision gate: results within 1e-2
 * of the fp32 reference).  Inputs, outputs, biases and accumulation stay fp32; activations and weights enter the tensor
 * core as plain bf16 (one MMA chain instead of six).  No counterpart in the reference, which runs the tokenizer in fp32
 * (algo/icl.py:788, 976: self.nets.float()); BASELINE config 5 names bf16 for the policy step around it. */
LIPVQ_API int lipvq_mlp3_fwd_supported(int64_t N, int d0, int d1, int d2, int d3);
LIPVQ_API int lipvq_mlp3_fwd_bf16_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1,
                        const float* W2, const float* b2, int d2, int act2, const float* W3, const float* b3, int d3,
                        int act3, float* Y, void* stream);
LIPVQ_API int lipvq_mlp3_fwd_f32(const float* X, int64_t N, int d0, const float* W1, const float* b1, int d1, int act1,
                        const float* W2, const float* b2, int d2, int act2, const float* W3, const float* b3, int d3,
                        int act3, float* Y, void* stream);
LIPVQ_API int lipvq_linear_bwd_x_f32(const float* dY, int64_t N, int out_dim, const float* W, int in_dim, const float* X,
                           int act_in, float* dX, void* stream);
LIPVQ_API int lipvq_linear_bwd_w_f32(const float* dY, int64_t N, int out_dim, const float* X, int in_dim, int act_in, float* dW,
                           float* db, void* stream);

/* dX = dY * act'(.) evaluated from the layer OUTPUT y (sigmoid: y(1-y); relu: [y>0]; none: 1).  For modules used
 * stand-alone (LipschitzMLP as q_encoder, hq_vae.py:375) whose upstream gradient arrives at the activation output. */
LIPVQ_API int lipvq_act_bwd_from_output_f32(const float* dY, const float* Y, int act, int64_t numel, float* dX, void* stream);

/* L-inf Lipschitz weight clamp (normalization(), backbone_lfqvae_v5.py:6-12):
 * scale[o] = min(1, softplus(ci[o]) / sum_i |W[o,i]|);  W_norm = W * scale.  */
LIPVQ_API int lipvq_lipschitz_scale_fwd_f32(const float* W, const float* ci, int out_dim, int in_dim, float* scale,
                                  float* W_norm, void* stream);
/* given G = dL/dW_norm: dW += G*scale + [scale<1] * (-(sum_i G W) * softplus(ci)/l1^2) * sign(W);
 * dci += [scale<1] * (sum_i G W)/l1 * sigmoid(ci).  (autograd of the expression above) */
LIPVQ_API int lipvq_lipschitz_scale_bwd_f32(const float* G, const float* W, const float* ci, int out_dim, int in_dim, float* dW,
                                  float* dci, void* stream);

/* ---------------------------------------------------------------------------------------------
 * K8  MS-TCN layers of the hierarchical tokenizer.  Replaces DilatedTemporalConvLayer.forward
 * (heirarchical_vqvae.py:39-57: Conv1d(E,E,3,dilation=d,padding=d) -> ReLU -> Dropout -> + x), the
 * [B,T,E]<->[B,E,T] transposes of ActionEncoder/ActionDecoder (:147-158, 202-210) and nn.LayerNorm(E) (:137,159).
 *
 * Activations are token-major [N = B*T, E]; time index of row n is n % T.  Taps that can only read padding
 * (dilation >= T) are skipped.  E % 4 == 0 and Ei == Eo (the reference never builds the 1x1 residual conv).
 *   pack:   W [Eo][Ei][3] (nn.Conv1d layout) -> Wf [Eo][3][Ei] and/or Wb [3][Eo][Ei]
 *   fwd:    OUT[n,o] = keep ? relu(pre)/(1-p) : 0  + X[n,o];  mask[n,o] = keep (uint8);  keep = pre>0 && not dropped
 *           (dropout: counter-based hash of (seed, element index); p = 0 in eval)
 *   bwd_x:  dX = dOUT + conv^T(dOUT * mask /(1-p))
 *   bwd_w:  dWf [Eo][3][Ei] += ... ; db[o] += sum_n dOUT*mask/(1-p)   (ACCUMULATED INTO)
 *   unpack: dW [Eo][Ei][3] += dWf
 * E % 128 == 0 with at least 32 tokens: the three entry points run on the tensor cores (csrc/conv_tc.cu: three bf16
 * terms per operand, fp32 accumulation; values within 1e-5 of the fp32 FFMA tiles, same keep-mask and dropout stream;
 * needs N * E < 2^32 and an 8-byte-aligned mask, else -- or with LIPVQ_CONV_ALGO=simt -- the FFMA tiles run;
 * LIPVQ_CONV_PRECISION=bf16: plain bf16 operands and one MMA chain, values within 1e-2).
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_conv_pack_weights_f32(const float* W, int Eo, int Ei, float* Wf, float* Wb, void* stream);
LIPVQ_API int lipvq_conv_unpack_grad_f32(const float* dWf, int Eo, int Ei, float* dW, void* stream);
LIPVQ_API int lipvq_dilated_conv_fwd_f32(const float* X, int64_t N, int T, int Ei, const float* Wf, const float* bias, int Eo,
                               int dilation, float drop_p, uint32_t seed, float* OUT, uint8_t* mask, void* stream);
LIPVQ_API int lipvq_dilated_conv_bwd_x_f32(const float* dOUT, const uint8_t* mask, int64_t N, int T, int Eo, const float* Wb,
                                 int Ei, int dilation, float drop_p, float* dX, void* stream);
LIPVQ_API int lipvq_dilated_conv_bwd_w_f32(const float* dOUT, const uint8_t* mask, const float* X, int64_t N, int T, int Eo,
                                 int Ei, int dilation, float drop_p, float* dWf, float* db, void* stream);
/* nn.LayerNorm over the last dim (biased variance, eps inside the sqrt); mean/rstd [rows] are kept for backward.
 * bwd: dw/db ACCUMULATED INTO (may be NULL). */
LIPVQ_API int lipvq_layernorm_fwd_f32(const float* X, int64_t rows, int E, const float* w, const float* b, float eps, float* Y,
                            float* mean, float* rstd, void* stream);
LIPVQ_API int lipvq_layernorm_bwd_f32(const float* dY, const float* X, const float* mean, const float* rstd, const float* w,
                            int64_t rows, int E, float* dX, float* dw, float* db, void* stream);
/* backward of F.normalize(x, dim=-1): dx = (dy - y (y.dy)) / max(||x||, 1e-12), y = x/max(||x||,1e-12)
 * (cosine quantisation and the normalised commitment losses, heirarchical_vqvae.py:441-459, 652-667) */
LIPVQ_API int lipvq_l2_normalize_rows_bwd_f32(const float* x, const float* dy, int64_t rows, int D, float* dx, void* stream);

/* y = a*x + b (one fma per element; y may alias x).  `distances = 1.0 - sims` of heirarchical_vqvae.py:461,494. */
LIPVQ_API int lipvq_affine_f32(const float* x, float a, float b, int64_t numel, float* y, void* stream);

/* mean-squared error (F.mse_loss, reduction='mean'), fused with its gradient:
 * sqerr_sum (double, accumulated into) += sum (y-t)^2 ; if dY != NULL: dY = 2/(numel) * (y - t) * gscale,
 * and multiplied by act'(y) for act_out in {NONE, RELU} when the layer applied it in its epilogue. */
LIPVQ_API int lipvq_mse_fwd_bwd_f32(const float* Y, const float* T, int64_t numel, const float* gscale_dev, int act_out,
                          double* sqerr_sum, float* dY, void* stream);

/* T1 loss tail (backbone_lfqvae_v5.py:79-83): losses_out[0..3] = {loss, recon, commit, codebook} with
 * recon = sq[0]/n_recon, commit = codebook = sq[1]/n_lat, loss = recon + w_commit*commit + w_cb*codebook (fp32). */
LIPVQ_API int lipvq_t1_loss_finalize(const double* sq2, double n_recon, double n_lat, float w_commit, float w_cb,
                           float* losses_out4, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Optimiser: torch.optim.AdamW single step over one flat fp32 buffer (robomimic/algo/icl.py:1437-1441:
 * AdamW(lr=1e-3, weight_decay=1e-4) on the tokenizer).  grad_scale multiplies g first (1/world_size
 * after a sum all-reduce).  step is 1-based.
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_adamw_step_f32(float* p, const float* g, float* m, float* v, int64_t n, int step, const int32_t* step_dev,
                         float lr, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                         void* stream);
/* *ctr += delta (device-resident step counter so a captured CUDA graph can be replayed step after step);
 * when step_dev != NULL above, the step number is read from it instead of `step`. */
LIPVQ_API int lipvq_counter_add_i32(int32_t* ctr, int delta, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Data-parallel reduction slab: one-shot sum all-reduce over NVLink peer memory (csrc/allreduce.cu).
 * Every rank holds a symmetric buffer [slab (n floats) | flags] mapped into all peers
 * (torch.distributed._symmetric_memory); peer_ptrs_dev is a device array of the W base addresses as seen from this
 * process, in rank order.  out (local, 16-byte aligned) receives sum over ranks of slab[0..n); all ranks get the same
 * bits.  cta_epoch: lipvq_allreduce_oneshot_ctas() zero-initialised uint32 on the device, private to this rank; the
 * flags region (lipvq_allreduce_oneshot_flag_bytes(world) bytes at flag_offset_bytes) must be zero before the first
 * call on any rank.  Every rank must issue the call the same number of times (it is a collective).
 * Replaces the dist.all_reduce a data-parallel port of the reference would add around
 * hq_vae.py:264-275 (EMA statistics) and optimizer.step() (gradients); NCCL stays available as the fallback.
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_allreduce_oneshot_ctas(void);
LIPVQ_API size_t lipvq_allreduce_oneshot_flag_bytes(int world);
LIPVQ_API int lipvq_allreduce_oneshot_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n,
                                int64_t flag_offset_bytes, float* out, uint32_t* cta_epoch, void* stream);
/* The same reduction with its second flag exchange taken off the critical path: ..._deferred_f32 only SIGNALS "I have
 * read your slab" on its way out, and ..._clear_f32 -- which replaces the memset that clears the slab at the start of
 * the next step -- waits for the peers' signals of the last reduction (about a step old by then) before it zeroes
 * slab[0..n).  The pair must be used together: after a deferred reduction the slab may only be written again behind
 * lipvq_allreduce_oneshot_clear_f32.  Same arguments as above (clear takes no `out`). */
LIPVQ_API int lipvq_allreduce_oneshot_deferred_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n,
                        int64_t flag_offset_bytes, float* out, uint32_t* cta_epoch, void* stream);
LIPVQ_API int lipvq_allreduce_oneshot_clear_f32(const uint64_t* peer_ptrs_dev, int rank, int world, int64_t n,
                        int64_t flag_offset_bytes, const uint32_t* cta_epoch, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Callers and data formats either side of the tokenizer.
 *
 * Action-chunk input side.  Replaces, for demonstrations resident in HBM, the per-sample Python of
 * SequenceDataset.get_item / get_sequence_from_demo (robomimic/utils/dataset.py:499-632): window of
 * n_frame_stack-1 frames back and seq_length frames forward around index_in_demo, rows outside the demo
 * repeated from its first / last frame ("pad_same"), optional (x - offset) / scale normalisation
 * (obs_utils.normalize_dict, obs_utils.py:464-501).
 *   data [L_total, A]: all demos back to back; demo_row0[d] = first row of demo d, demo_len[d] its length,
 *   demo_seq0[d] = first sequence index of demo d in the dataset's index map (dataset.py:229-260);
 *   sample_demo[b], sample_index[b]: the sampled sequences; out [B, T, A], T = n_frame_stack-1+seq_length;
 *   pad_mask [B, T] uint8 (1 = real frame) or NULL; offset/scale [A] or both NULL.
 * --------------------------------------------------------------------------------------------- */
LIPVQ_API int lipvq_window_gather_f32(const float* data, const int64_t* demo_row0, const int32_t* demo_len,
                                      const int64_t* demo_seq0, const int32_t* sample_demo, const int64_t* sample_index,
                                      int64_t B, int A, int n_frame_stack, int seq_length, int pad_frame_stack,
                                      const float* offset, const float* scale, float* out, uint8_t* pad_mask, void* stream);

/* AdaptiveBinActionEmbedding (robomimic/models/bin_action/backbone.py:5-89): the A/B baseline tokenizer behind the same
 * boundary (obs_nets.py:1214-1217, 1346).  fwd: optionally running_min/max <- min/max with the batch (:36-39), bin
 * boundaries torch.linspace(min, max, num_bins+1) (:41-52), idx = clamp(bucketize(x) - 1, 0, num_bins-1) (:54-67),
 * out[n, a*E:(a+1)*E] = tables[a][idx] (:80-86); idx_out [N, A] int64 or NULL.  bwd: dtables[a][idx[n,a]] += dout[n, a, :].
 * tables [A][num_bins][E]. */
LIPVQ_API int lipvq_bin_embed_fwd_f32(const float* actions, int64_t N, int A, int num_bins, int E, int update_stats,
                                      float* running_min, float* running_max, const float* tables, float* out,
                                      int64_t* idx_out, void* stream);
LIPVQ_API int lipvq_bin_embed_bwd_f32(const float* dout, const int64_t* idx, int64_t N, int A, int num_bins, int E,
                                      float* dtables, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LIPVQ_H_ */
