/* gfedntm_b200 — C ABI of the sm_100a kernel library (libgfedntm_kernels.so).
 *
 * Every entry point takes raw device pointers plus a cudaStream_t (passed as void*) and returns
 * 0 on success, non-zero on error (a message is printed to stderr).  Nothing here allocates
 * device memory unless stated; callers (gfedntm_b200.ops) own all buffers.
 *
 * "Ref" comments cite the reference op each kernel replaces (paths relative to the reference
 * repository root).
 */
#ifndef GFEDNTM_KERNELS_H_
#define GFEDNTM_KERNELS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------
 * Library info
 * ---------------------------------------------------------------------------------------- */
/* Returns the ABI version of the loaded library. */
int gf_abi_version(void);
/* Number of kernel launches issued through this library since load (all entry points). */
long long gf_launch_count(void);
/* sizeof() of argument struct number `which` (order of declaration in this header), -1 if unknown. */
int gf_sizeof(int which);

/* ------------------------------------------------------------------------------------------
 * Decoder sweeps (tcgen05 / TMEM / TMA).
 * Ref: src/models/base/pytorchavitm/avitm_network/decoder_network.py:121-126 (theta.beta ->
 *      BatchNorm1d(V) -> softmax over V) and avitm.py:225 (BoW NLL), plus their autograd.
 * ---------------------------------------------------------------------------------------- */
enum GfSweepMode {
  GF_ROW_LOGITS = 0, /* out[b,v] = theta_bf[b,:].beta_bf[:,v] (fp32)                           */
  GF_ROW_LSE = 1,    /* lse_part[b,chunk] = online (max2,sum) of 2^(logit*colc.x + colc.y)     */
  GF_ROW_GRAD = 2,   /* dtheta[b,:] += sum_v dlogit[b,v] * beta[:,v]      (dense part)         */
  GF_COL_STATS = 3,  /* colc[v] = {rstd*log2e, bias, 0, 0}; running stats update               */
  GF_COL_GRAD = 4,   /* colc[v].zw = BN-backward constants; dbeta[:,v] = sum_b dlogit*theta_c  */
  GF_COL_LOGITS = 5, /* same result as GF_ROW_LOGITS through the word-major operand family     */
  /* Folded modes (no side band): beta_bf is the SCALED shadow written by gf_scale_beta, theta_bf
     carries the constant-one rows K, K+1 and the -e_b rows K+2, K+3 (see gf_theta_center).      */
  GF_ROW_LSE2 = 6,   /* lse_part[b,chunk] = online (max2,sum) of 2^D1                           */
  GF_COL_GRAD2 = 7,  /* one pass: colc[v].zw, dbeta[:,v], beta3 (gamma-scaled beta for gf_gram)  */
  GF_ROW_GRAD2 = 8,  /* dtheta[b,:K] += ln2 * sum_v 2^D1[b,v] * beta'[:,v]                      */
  GF_COL_STATS2 = 9  /* GF_COL_STATS from beta_v^T C beta_v (C hi/lo tiles at ext_col0 of theta_bf) */
};

typedef struct GfSweepArgs {
  int mode;              /* GfSweepMode */
  int B, V, K, Kp;       /* documents, vocabulary, topics, topics padded to a multiple of 16  */
  int n_chunks;          /* ROW modes: vocabulary split per document tile (>=1)                */
  const void* theta_bf;  /* theta_c TRANSPOSED: [Kp, ld_theta] bf16 (documents contiguous,
                            ld_theta >= B, multiple of 8), rows >= K zero                       */
  int ld_theta;
  const void* beta_bf;   /* [Kp, ld_beta] bf16, rows >= K zero                                  */
  int ld_beta;
  const void* sb_stream; /* float4 side band of the streamed axis, padded to a multiple of 128:
                            ROW modes: colc (per word); COL modes: {lse*log2e, S, 0, 0} per doc.
                            Padding entries must be finite zeros.                               */
  float* out;            /* LOGITS modes: [B,V] fp32                                            */
  void* lse_part;        /* ROW_LSE: float2 [B, n_chunks]                                       */
  const void* doc_const; /* ROW_GRAD: float2 [B] {lse*log2e, S*ln2}                             */
  float* dtheta;         /* ROW_GRAD: [B, Kp] fp32, accumulated with atomics                    */
  void* colc;            /* float4 [ceil(V/128)*128]                                            */
  const float* theta_mean; /* COL_STATS: [K] batch mean of theta (for running_mean), may be 0  */
  float* running_mean;   /* [V] or null                                                         */
  float* running_var;    /* [V] or null                                                         */
  float bn_eps, bn_momentum;
  int use_running;       /* COL_STATS: 1 = eval mode (normalise with running statistics)        */
  const float* asp;      /* COL_GRAD: [V] sparse part of sum_b dy                               */
  const float* csp;      /* COL_GRAD: [V] sparse part of sum_b dy*y                             */
  float* dbeta;          /* COL_GRAD: [K, ld_dbeta] fp32                                        */
  int ld_dbeta;
  const float* corr;     /* COL_GRAD: [V, Kp] sparse correction (subtracted), may be null       */
  const int* colptr;     /* COL_GRAD: [V+1] CSC pointer of the batch (required with corr)       */
  /* --- folded modes --- */
  int theta_cols;        /* readable columns of theta_bf (0 => B): padded documents + pseudo docs */
  int beta_cols;         /* readable columns of beta_bf (0 => V): vocabulary padded to 128        */
  int ext_col0;          /* COL_STATS2 / COL_GRAD2: first column of the C = theta_c^T theta_c block (hi, then lo)
                            in theta_bf; ROW_GRAD2: first column of the M hi / lo blocks in beta_bf   */
  const void* betaT_bf;  /* COL_GRAD2: [V, Kp] bf16 UNSCALED beta^T                               */
  void* beta3;           /* COL_GRAD2 out: [Kp, ld_beta3] bf16, rows k<K gamma_v*beta[k,v], row K alpha_v */
  int ld_beta3;
  int ld_corr;           /* COL_GRAD2: corr is TRANSPOSED, [Kp, ld_corr] (gf_csc_corr_t)                */
  const void* fed;       /* ROW_GRAD2: DEVICE copy of a GfFedArgs (n_rank_ctx 1, step_ctr set) or null.  The
                            last fed_ctas CTAs of the grid then wait for every peer's
                            GF_FED_PHASE_BETA_READY flag and run the beta part of the aggregation (TMA
                            pulls of the peers' bf16 gradient tiles + sharded Adam + pushes of the bf16
                            operand shadows) while the other CTAs do the sweep; a sweep CTA that runs out
                            of sweep work joins them (word tiles are handed out by an atomic counter).
                            Finish the step with gf_fed_aggregate(parts = 1).                            */
  int fed_ctas;
  const void* fed_tmaps; /* ROW_GRAD2 with fed: HOST pointer to the GF_MAX_RANKS tensor maps written by
                            gf_fed_make_tmaps (one per rank's bf16 beta gradient; passed to the kernel as
                            a parameter).  They cover plain (non-multicast) loads of peer memory.        */
  void* dbeta_bf;        /* COL_GRAD2: when non-null the beta gradient is written as bf16 [K, ld_dbeta_bf]
                            here INSTEAD of fp32 `dbeta` (the federated exchange format)                 */
  int ld_dbeta_bf;
} GfSweepArgs;

int gf_dec_sweep(const GfSweepArgs* args, void* stream);
/* Vocabulary chunk count for the ROW modes that fills the persistent grid in whole waves.       */
int gf_dec_row_chunks(int mode, int B, int V, int Kp);
/* Same for an explicit number of sweep CTAs (GF_ROW_GRAD2 with fed_ctas CTAs set aside).            */
int gf_dec_row_chunks_grid(int B, int V, int grid);
/* Number of SMs (= persistent CTAs of a sweep) on the current device.                            */
int gf_num_sms(void);

/* ------------------------------------------------------------------------------------------
 * Encoder / latent layer (CUDA-core kernels on [B,H] / [B,K] data; the BoW input is CSR).
 * Ref: avitm_network/inference_network.py:76-85, decoder_network.py:102-118, avitm.py:207-220;
 *      ctm_network/inference_network.py:79-94,160-193.
 * ---------------------------------------------------------------------------------------- */
#define GF_MAX_HIDDEN 8

enum GfActivation {
  GF_ACT_SOFTPLUS = 0, GF_ACT_RELU, GF_ACT_SIGMOID, GF_ACT_TANH,
  GF_ACT_LEAKYRELU, GF_ACT_RRELU, GF_ACT_ELU, GF_ACT_SELU
};

typedef struct GfEncArgs {
  int B, V, K, L;            /* L = number of hidden layers, H[l] their widths (<= 256)       */
  int H[GF_MAX_HIDDEN];
  int act, training;
  /* sparse first-layer input (may be null for ZeroShotTM): CSR of the minibatch             */
  const int* indptr; const int* indices; const float* counts;
  const float* W1T;          /* [V, H0]  input_layer.weight[:, :V] transposed                 */
  const float* b1;           /* [H0]                                                          */
  /* dense first-layer input (CTM): xdense [B,C], WdT [C,H0]                                  */
  const float* xdense; int C; const float* WdT;
  const float* Wh[GF_MAX_HIDDEN]; const float* bh[GF_MAX_HIDDEN]; /* hidden l: [H[l],H[l-1]]  */
  const float* Wmu; const float* bmu; const float* Wsig; const float* bsig; /* [K, H_last]    */
  float* pre[GF_MAX_HIDDEN]; /* out: pre-activations [B, H[l]]                                */
  float enc_drop_p;
  const float* mask_enc_in;  /* optional injected dropout multipliers [B, H_last]             */
  float* mask_enc;           /* out: multipliers used                                         */
  float* mu_pre; float* ls_pre; /* out: [B,K] head outputs before BatchNorm                   */
  double* bn_sums;           /* out (+=): [4K] sum mu, sum mu^2, sum ls, sum ls^2 (zeroed by caller) */
  unsigned long long seed;
  const long long* step_ctr; /* optional device counters (gf_step_begin): seed advances by ctr[0]    */
  int w_smem;                /* set by gf_enc_fwd (callers leave 0): head weights staged in shared memory */
  const float* pre0_ext;     /* optional [B, H0]: first-layer contribution computed elsewhere (the dense contextual
                                input on the tensor cores, gf_ctm_rc); added to the pre-activation, xdense unused */
} GfEncArgs;
/* Device-side step counters, so that a whole training step can be replayed from a CUDA graph:
 * ctr[0] counts forward passes (advances the dropout / reparameterisation seeds), ctr[1] counts
 * optimiser steps (Adam bias correction, federated epoch flag).                                  */
int gf_step_begin(long long* ctr, int inc_rng, int inc_opt, void* stream);
/* gf_step_begin (ctr may be null) and the zero fill of up to GF_ZERO_MAX buffers (16-byte aligned, sizes % 4 == 0;
 * `ptrs` / `bytes` are HOST arrays) in one launch. */
#define GF_ZERO_MAX 6
int gf_step_begin_zero(long long* ctr, int inc_rng, int inc_opt, void* const* ptrs, const long long* bytes, int n,
                       void* stream);
int gf_enc_fwd(const GfEncArgs* args, void* stream);

/* Supervised CTM (label_size > 0; ref: ctm_network/inference_network.py:79-94,178-183, decoding_network.py:84-85,
 * 156-159, ctm.py:292-296).  gf_small_linear_fwd: out[b,:] (+)= x[b,:] W (the one-hot labels' rows of the first
 * layer); gf_small_linear_wgrad: dW = x^T g; gf_label_head: logits = theta Wc^T + bc, loss += mean cross-entropy
 * against argmax(labels), and with dtheta != null: dtheta[b,:K] += d/d theta, dWc / dbc += (atomics, zeroed by the
 * caller); gradients carry gscale / B.                                                                        */
int gf_small_linear_fwd(const float* x, const float* W, float* out, int B, int L, int H, int accumulate, void* stream);
int gf_small_linear_wgrad(const float* x, const float* g, float* dW, int B, int L, int H, void* stream);
int gf_label_head(const float* theta, const float* Wc, const float* bc, const float* labels, int B, int K, int L,
                  float gscale, double* loss, float* dtheta, int ld_dtheta, float* dWc, float* dbc, void* stream);
/* First-layer contribution of the sparse BoW input: out[b,:] (+)= sum over the document's non-zeros of
 * count * W1T[word,:]  (several warps per document, rows kept in flight; feed the result to gf_enc_fwd through
 * GfEncArgs.pre0_ext with indptr = null).  Ref: inference_network.py:78.                                    */
int gf_bow_gather(const int* indptr, const int* indices, const float* counts, const float* W1T, int H0, int B,
                  float* out, int accumulate, void* stream);
/* Monte-Carlo document-topic inference in one launch: mean over n_samples reparameterised soft-max draws from the
 * eval-mode encoder outputs (gf_enc_fwd with training = 0), then entries < thr zeroed and rows L1-renormalised
 * (thr <= 0: plain mean).  Ref: avitm.py:470-523 + federated_model.py:170-173.                              */
int gf_theta_mc(const float* mu_pre, const float* ls_pre, const float* mu_rm, const float* mu_rv, const float* ls_rm,
                const float* ls_rv, float bn_eps, int B, int K, int n_samples, unsigned long long seed, float thr,
                float* out, void* stream);

typedef struct GfLatentArgs {
  int B, K, training, sample;   /* sample: draw eps ~ N(0,1) (else eps = 0 unless eps_in)     */
  const float* mu_pre; const float* ls_pre;
  const double* bn_sums;        /* [4K] from gf_enc_fwd (training)                            */
  float* mu_rm; float* mu_rv; float* ls_rm; float* ls_rv; /* running stats [K]                */
  float bn_eps, bn_momentum;
  const float* prior_mean; const float* prior_var;
  const float* eps_in; const float* mask_theta_in; /* optional injected noise / multipliers   */
  float theta_drop_p;
  float* mu; float* logvar; float* eps; float* theta_s; float* mask_theta; float* theta; /* [B,K] */
  float* theta_colsum;          /* (+=) [K], zeroed by caller; may be null                    */
  float* kl;                    /* [B]                                                        */
  double* loss;                 /* (+=) scalar, may be null                                   */
  float kl_weight;
  unsigned long long seed;
  const long long* step_ctr;
} GfLatentArgs;
int gf_latent_fwd(const GfLatentArgs* args, void* stream);

/* theta_c = theta - colsum/B (center=1) or theta (center=0) as bf16 in two layouts:
 * out_bf16 [B, ld] (row per document, zero padded to ld columns; sparse kernels) and
 * outT_bf16 [Kp, ldT] (row per topic, documents contiguous; tensor-core sweeps).               */
int gf_theta_center(const float* theta, const float* colsum, void* out_bf16, void* outT_bf16,
                    float* mean_out, int B, int K, int ld, int ldT, int center, void* stream);
/* Same, plus the folded-mode rows of outT for all documents up to Bpad (multiple of 128, <= ldT):
   rows K, K+1 = 1; rows K+2, K+3 = 0 for b < B (filled in by gf_dec_sparse_doc) and row K+2 =
   -30000 for padded documents (their probabilities vanish).  Needs ld >= K + 4.                 */
int gf_theta_center_fold(const float* theta, const float* colsum, void* out_bf16, void* outT_bf16,
                         float* mean_out, int B, int K, int ld, int ldT, int Bpad, int center, void* stream);
/* Scaled operand shadow for the folded sweeps: out[k,v] = bf16(beta_bf[k,v] * colc[v].x) for k < K,
   rows K, K+1 = hi/lo bf16 split of colc[v].y, rows K+2, K+3 = 1; words v in [V, cols) get row K =
   -30000 and zeros elsewhere.  colc is the float4 array written by GF_COL_STATS.                */
int gf_scale_beta(const void* beta_bf, int ld_beta, const void* colc, void* out, int ld_out, int K, int Kp,
                  int V, int cols, void* stream);
/* C = theta_c^T theta_c (fp32 accumulate) written as a bf16 hi / lo pair into outT[j, ext_col0 + k]
   and outT[j, lo_col0 + k], j,k < K.  c_scratch: [K*K] fp32, zero on entry, left zero on return.  */
int gf_theta_gram(const void* theta_bf, int ld, int B, int K, void* outT_bf16, int ldT, int ext_col0,
                  int lo_col0, float* c_scratch, void* stream);
/* Split-V tcgen05 GEMM: partial[g] [Kp,Kp] = sum over the CTA's words of x[:,v] y[:,v]^T
   (x = beta3 of GF_COL_GRAD2, y = unscaled beta shadow), each partial stored TRANSPOSED.
   partial must hold gf_gram_parts()*Kp*Kp floats. */
int gf_gram_parts(void);
int gf_gram(const void* x_bf, int ldx, const void* y_bf, int ldy, int V, int Kp, float* partial, void* stream);
/* M = sum of the partials (rows j <= K; row K is a = beta.alpha) written as a bf16 hi / lo pair into
   out[j, col_hi + k] and out[j, col_lo + k]: the pseudo-word blocks of beta' read by GF_ROW_GRAD2,
   whose MMA1 then yields a_k + (theta_c M)[b,k] (theta_c^T row K is the constant one).           */
int gf_gram_finish(const float* partial, int n_parts, int K, int Kp, void* out_bf16, int ld_out, int col_hi,
                   int col_lo, void* stream);

typedef struct GfLatentBwdArgs {
  int B, K;
  const float* dtheta; int ld_dtheta;  /* gradient w.r.t. theta (post dropout)                */
  const float* mask_theta; const float* theta_s; const float* mu; const float* logvar; const float* eps;
  const float* prior_mean; const float* prior_var;
  float kl_weight, gscale;
  float* dmu; float* dlv;              /* out: gradient w.r.t. BatchNorm outputs [B,K]        */
  float* bn_bsums;                     /* (+=) [4K] sum d, sum d*xhat for mu then logvar      */
  float* d_prior_mean; float* d_prior_var; /* (+=) [K], may be null                           */
} GfLatentBwdArgs;
int gf_latent_bwd(const GfLatentBwdArgs* args, void* stream);

typedef struct GfEncBwdArgs {
  int B, V, K, L;
  int H[GF_MAX_HIDDEN];
  int act;
  const float* dmu; const float* dlv; const float* mu; const float* logvar;
  const double* bn_sums; const float* bn_bsums; float bn_eps;
  const float* pre[GF_MAX_HIDDEN]; const float* mask_enc;
  const float* Wh[GF_MAX_HIDDEN]; const float* Wmu; const float* Wsig;
  float* dWh[GF_MAX_HIDDEN]; float* dbh[GF_MAX_HIDDEN]; /* (+=)                                */
  float* dWmu; float* dbmu; float* dWsig; float* dbsig; float* db1; /* (+=)                   */
  float* dpre0;              /* out: [B,H0] gradient w.r.t. first pre-activation              */
  const float* xdense; int C; float* dWdT; /* CTM dense input: (+=) [C,H0]                     */
  unsigned long long seed;
  const long long* step_ctr;
} GfEncBwdArgs;
int gf_enc_bwd(const GfEncBwdArgs* args, void* stream);

/* ------------------------------------------------------------------------------------------
 * Sparse side: minibatch transpose, decoder non-zeros, scatter kernels, NeuralLDA decoder.
 * ---------------------------------------------------------------------------------------- */
/* CSR -> CSC of a minibatch.  colptr [V+1], csc_row/csc_src [nnz]; cursor_scratch is V ints of
 * scratch owned by the caller.  Ref: client.py:461-462 builds a dense [n_docs,V] array instead. */
int gf_csr_to_csc(const int* indptr, const int* indices, int B, int V, int nnz, int* colptr,
                  int* cursor_scratch, int* csc_row, int* csc_src, void* stream);
/* Same, with the number of non-zeros read on the device (indptr[B]); `nnz_capacity` only sizes the
 * launch.  Lets a captured CUDA graph be replayed on minibatches with different nnz.            */
int gf_csr_to_csc_dyn(const int* indptr, const int* indices, int B, int V, int nnz_capacity, int* colptr,
                      int* cursor_scratch, int* csc_row, int* csc_src, void* stream);

/* Same transpose, for callers that cannot allocate device memory (a host whose own pool already owns
 * the whole device, or a latency-critical loop where cudaMalloc/cudaFree would stall every stream).
 * No counterpart in the reference: it exists only because this library is callable through its C ABI
 * from such hosts.  The V write cursors live in a fixed array inside the library image
 * (GF_CSC_STATIC_MAX_V ints = 4 MiB, held by every GPU that loads the library until process exit).
 *   - V > GF_CSC_STATIC_MAX_V is refused with an error (use gf_csr_to_csc).
 *   - The array is shared by all calls on a GPU and keeps the previous call's contents; each call
 *     zeroes the cursors it needs on `stream` first.  Calls that can overlap on the GPU race.
 *   - The first call on a device resolves the array's device address; make it during set-up, before
 *     starting any stream capture on that device.
 * Inputs and outputs are passed exactly as for gf_csr_to_csc; nothing is allocated or freed.        */
#define GF_CSC_STATIC_MAX_V (1 << 20)
int gf_csr_to_csc_static(const int* indptr, const int* indices, int B, int V, int nnz, int* colptr,
                         int* csc_row, int* csc_src, void* stream);

typedef struct GfDecSparseArgs {
  int B, Kp, n_chunks;
  const int* indptr; const int* indices; const float* counts;
  const void* theta_bf; int ld_theta;   /* bf16 [B, ld_theta]                                  */
  const void* betaT_bf;                 /* bf16 [V, Kp]                                        */
  const void* colc;                     /* float4 [Vpad]                                       */
  const void* lse_part;                 /* float2 [B, n_chunks]                                */
  float gscale;                         /* upstream gradient scale (federated client weight)   */
  float* rl;                            /* out [B] reconstruction loss                         */
  double* loss;                         /* (+=) scalar, may be null                            */
  float* asp; float* csp;               /* (+=) [V], zeroed by caller; null => forward only    */
  float* tnnz;                          /* out [nnz]                                           */
  float* dtheta;                        /* out [B,Kp] sparse part (plain store); may be null   */
  void* doc_const;                      /* out float2 [B]                                      */
  void* sb_doc;                         /* out float4 [Bpad]                                   */
  void* thetaT_bf; int ldT; int K;      /* folded modes: rows K+2, K+3 of theta_c^T receive the
                                           hi/lo bf16 split of -(lse2_b - log2 S_b); may be null */
} GfDecSparseArgs;
int gf_dec_sparse_doc(const GfDecSparseArgs* args, void* stream);

int gf_csc_corr(const int* colptr, const int* csc_row, const int* csc_src, const float* tnnz,
                const void* theta_bf, int ld_theta, float* corr, int V, int Kp, void* stream);
/* Same values, stored transposed: corrT[k, v] ([Kp, ldc], only words with entries are written).     */
int gf_csc_corr_t(const int* colptr, const int* csc_row, const int* csc_src, const float* tnnz,
                  const void* theta_bf, int ld_theta, float* corrT, int ldc, int V, int Kp, void* stream);
/* out[v,:N] (+)= sum_{entries of column v} val[src] * dense[row,:N]                          */
int gf_csc_spmm(const int* colptr, const int* csc_row, const int* csc_src, const float* val,
                const float* dense, int ldd, float* out, int ldo, int V, int N, int accumulate,
                void* stream);

/* NeuralLDA decoder.  Ref: decoder_network.py:127-132.                                        */
int gf_lda_beta_fwd(const float* beta, int ldb, int K, int V, float eps, float momentum, int training,
                    float* running_mean, float* running_var, float* xhat, float* col_rstd, float* bp,
                    void* stream);
int gf_lda_doc(const int* indptr, const int* indices, const float* counts, const float* theta,
               const float* bp, int B, int K, int V, float gscale, float* rl, double* loss, float* dtheta,
               int ld_dtheta, float* gnnz, void* stream);
int gf_lda_beta_bwd(const float* bp, const float* dbpT, int ldt, const float* xhat, const float* col_rstd,
                    float* dxhat, float* dbeta, int ldg, int K, int V, int training, void* stream);

/* Row top-k (k <= 32) of a [rows, V] matrix.  Ref: avitm.py:574.                               */
int gf_row_topk(const float* mat, int ld, int rows, int V, int topk, int* out_idx, float* out_val,
                void* stream);

/* ------------------------------------------------------------------------------------------
 * Optimiser.  Ref: torch.optim.Adam in avitm.py:141-143 / federated_avitm.py:78.
 * ---------------------------------------------------------------------------------------- */
/* Solvers of the reference's optimiser factory (avitm.py:141-153), torch defaults.  Two state buffers per
 * parameter (m, v): adam (exp_avg, exp_avg_sq), sgd (momentum buffer, -), adagrad (-, sum), adadelta (acc_delta,
 * square_avg), rmsprop (momentum buffer, square_avg).  b1 = momentum, b2 = 0.99 (adam / rmsprop alpha) or rho.   */
enum GfSolver { GF_ADAM = 0, GF_SGD = 1, GF_ADAGRAD = 2, GF_ADADELTA = 3, GF_RMSPROP = 4 };
/* `step_ctr` (optional, device): when non-null the step number is read from step_ctr[1] on the
 * device instead of the `step` argument.                                                        */
int gf_adam_flat(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                 float eps, float wd, int step, float gscale, const long long* step_ctr, int solver, void* stream);
/* Adam on beta [K,ldp] plus the bf16 operand shadows beta_bf [Kp,ldb] and betaT_bf [V,Kp];
 * do_update = 0 only refreshes the shadows.                                                  */
int gf_adam_beta(float* p, const float* g, float* m, float* v, int K, int V, int ldp, int ldg,
                 void* beta_bf, int ldb, void* betaT_bf, int Kp, float lr, float b1, float b2, float eps,
                 float wd, int step, float gscale, int do_update, const long long* step_ctr, int solver, void* stream);
/* The same update for a caller that runs it on a second stream NEXT TO GF_ROW_GRAD2: no shared memory, <= 64
 * registers x 128 threads, `blocks_per_sm` (default 2) persistent blocks per SM, streaming cache hints. */
int gf_adam_beta_bg(float* p, const float* g, float* m, float* v, int K, int V, int ldp, int ldg,
                    void* beta_bf, int ldb, void* betaT_bf, int Kp, float lr, float b1, float b2, float eps,
                    float wd, int step, float gscale, const long long* step_ctr, int solver, int blocks_per_sm,
                    void* stream);

/* ------------------------------------------------------------------------------------------
 * Skinny tcgen05 GEMMs of the CombinedTM encoder (csrc/ctm_gemm.cu): the adapt_bert layer collapsed into the
 * first encoder layer, M^T = A^T W, and its exact gradients dW = A dM^T + ab (x) dsum, dA = W dM, d ab = W dsum.
 * Ref: ctm_network/inference_network.py:160-193 and its autograd.  Operands are bf16 hi + lo shadows
 * (gf_ctm_split); every product is three MMAs (hi.hi + lo.hi + hi.lo) accumulated in fp32.
 * ---------------------------------------------------------------------------------------- */
/* hi = bf16(x), lo = bf16(x - hi) of src [rows, cols] (row pitch ld_src floats) into [rows_out, ldo] bf16;
 * column `cols` <- extra_col[row] (may be null), row `rows` <- tail_row[0..cols) (may be null; rows_out = rows+1
 * then), everything else zero.                                                                             */
int gf_ctm_split(const float* src, long long ld_src, const float* extra_col, const float* tail_row, int rows, int cols,
                 void* out_hi, void* out_lo, int ldo, void* stream);
/* Adam on a row-major [rows, cols] block (+ an optional per-row scalar p_col, folded in as column `cols`) fused with
 * the refresh of its bf16 hi / lo shadows [rows, ldo] (ldo % 8 == 0, padding columns zeroed); do_update = 0 only
 * refreshes the shadows.  step_ctr as in gf_adam_flat.  Ref: optimizer.step(), federated_ctm.py:76-78.       */
int gf_adam_split(float* p, const float* g, float* m, float* v, long long ld, int rows, int cols, float* p_col,
                  const float* g_col, float* m_col, float* v_col, void* out_hi, void* out_lo, int ldo, float lr, float b1,
                  float b2, float eps, float wd, int step, float gscale, int do_update, const long long* step_ctr,
                  int solver, void* stream);
/* out [rows, ld] <- 0, except row bias_row <- bias[0..cols)  (accumulation target of gf_ctm_vc).            */
int gf_ctm_init_out(float* out, int rows, int cols, int ld, int bias_row, const float* bias, void* stream);
/* out[m, n] += sum_r X[r, m] Y[r, n]   X hi/lo [R, ldx] (m_total valid columns), Y hi/lo [R, 64 nb]         */
int gf_ctm_vc(const void* x_hi, const void* x_lo, int ldx, int m_total, const void* y_hi, const void* y_lo, int nb,
              int R, float* out, int ldo, int n_valid, void* stream);
/* out[r, n] = sum_k X[r, k] Z[k, n]    X hi/lo [R, ldx] (k_total valid columns), Z hi/lo [z_rows, 64 nb]    */
int gf_ctm_rc(const void* x_hi, const void* x_lo, int ldx, int k_total, const void* z_hi, const void* z_lo, int nb,
              int z_rows, int R, float* out, int ldo, int n_valid, void* stream);
/* out[r, c] = sum_k X[r, k] Z[c, k] (c < n_valid), extra[r] = column n_valid (may be null)
 * X hi/lo [R, 64 nb], Z hi/lo [z_rows, 64 nb]                                                               */
int gf_ctm_rr(const void* x_hi, const void* x_lo, const void* z_hi, const void* z_lo, int nb, int z_rows, int R,
              float* out, int ldo, int n_valid, float* extra, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused federated aggregation over NVLink peer memory (no NCCL on this path).
 * Ref: src/federation/server.py:442-521 (gather / weighted mean / broadcast over gRPC) and the
 *      clients' optimizer.step() + load_state_dict (federated_avitm.py:78, federated_model.py:117-131).
 * ---------------------------------------------------------------------------------------- */
#define GF_MAX_RANKS 8

typedef struct GfFabricRank {
  int rank;
  float* G[GF_MAX_RANKS];         /* flat gradient buffer of every rank (peer-mapped)           */
  float* P[GF_MAX_RANKS];         /* flat parameter buffer of every rank                        */
  float* S[GF_MAX_RANKS];         /* flat BatchNorm running statistics of every rank            */
  void* beta_bf[GF_MAX_RANKS];    /* bf16 [Kp, ldb] shadow of every rank (may be null)          */
  void* betaT_bf[GF_MAX_RANKS];   /* bf16 [V, Kp] shadow of every rank (may be null)            */
  uint32_t* flags[GF_MAX_RANKS];  /* flag array of every rank: uint32 [GF_FED_FLAG_SLOTS][GF_MAX_RANKS]
                                     (GF_FED_FLAG_BYTES), zeroed once: slot 0 gradients ready, 1 pushes
                                     done, 2 beta gradient ready                                      */
  float* M; float* V2;            /* this rank's Adam moments (same indexing as P)              */
  unsigned int* done_counter;     /* this rank, zeroed once                                     */
  /* Optional NVLS multicast aliases of the same buffers (one multimem.st reaches every rank, the
     NVSwitch replicates it); null => one P2P store per rank.                                    */
  float* P_mc; float* S_mc; void* beta_bf_mc; void* betaT_bf_mc;
  /* Overlapped beta aggregation (GF_ROW_GRAD2 with `fed`): every rank's beta gradient as bf16
     [K, ld_gb] (written by GF_COL_GRAD2 via dbeta_bf) and this rank's word-tile counter.        */
  void* Gb[GF_MAX_RANKS];
  int ld_gb;
  unsigned int* agg_counter;      /* this rank, reset by gf_fed_signal(GF_FED_PHASE_BETA_READY)  */
  /* One-shot small all-reduce (gf_fed_allreduce_small): every rank's staging area, float [world][ar_cap] */
  float* AR[GF_MAX_RANKS];
  long long ar_cap;
} GfFabricRank;

/* A contiguous range of the flat parameter space [0, flat_n) with its sharing rule (grads_to_share of
 * the reference, federated_model.py:98-131): shared ranges are aggregated, the others stay client-local
 * (mode 0: local Adam on the client's own unweighted gradient; mode 1: untouched).               */
#define GF_FED_MAX_RANGES 48
typedef struct GfFedRange {
  long long off, len;             /* floats, both multiples of 4                                */
  int shared;                     /* 0 client-local, 1 aggregated, 2 replicated: the gradient in G already is the
                                     global one on every rank (mode 0: local Adam with unit gradient scale; mode 1:
                                     aggregated like 1); 3 = like 2 but the caller runs the optimiser on it
                                     itself (gf_adam_split keeps the operand shadows in step)                    */
  int pad_;
} GfFedRange;

typedef struct GfFedArgs {
  int world;                      /* number of federated clients (ranks)                        */
  int mode;                       /* 0 = weighted gradients + global Adam, 1 = weighted parameters */
  int epoch;                      /* strictly increasing per call (flag value)                  */
  int n_rank_ctx;                 /* 1 in production; = world for the single-GPU emulation      */
  long long flat_n;               /* elements before beta in the flat buffers (multiple of 4)   */
  long long beta_off;             /* offset of beta [K,V] in the flat buffers                   */
  long long s_n;                  /* BatchNorm statistics elements (multiple of 4)              */
  int K, V, Kp, ldb;              /* K = 0 skips the beta region                                */
  float w[GF_MAX_RANKS];          /* client weights n_c / sum n                                 */
  float lr, b1, b2, eps, wd;
  int step;                       /* Adam step count (>= 1) for the bias correction             */
  const long long* step_ctr;      /* optional device counters: epoch = step = step_ctr[1]       */
  int parts;                      /* 0 or 3 = everything; bit 0 = flat region + BatchNorm statistics, bit 1 =
                                     beta (which GF_ROW_GRAD2 can run on spare CTAs, see GfSweepArgs.fed) */
  /* Subset sharing.  n_ranges = 0: all of [0, flat_n), every BatchNorm statistic and beta are shared.  */
  int n_ranges;
  GfFedRange ranges[GF_FED_MAX_RANGES];
  int s_mask;                     /* bit i: block i of the statistics buffer is shared; blocks are
                                     {mu mean, mu var, sigma mean, sigma var}[K] then {beta mean, beta var}[V] */
  int beta_shared;                /* 0: beta stays local (the caller runs its own optimiser on it)        */
  float local_gscale;             /* mode 0, unshared ranges: gradient scale of the local Adam (1 / w[rank],
                                     undoing the client weight folded into the backward pass)             */
  int solver;                     /* GfSolver of the global / local update (b1 = momentum, b2 = rho / alpha)  */
  GfFabricRank ranks[GF_MAX_RANKS];
} GfFedArgs;

/* One call per rank and step, on every rank of the federation, with the same `epoch`.          */
int gf_fed_aggregate(const GfFedArgs* args, void* stream);
/* Raise flag slot `phase` (2 <= phase < GF_FED_FLAG_SLOTS) of this rank on every peer with the current
 * epoch.  GF_FED_PHASE_BETA_READY also resets this rank's word-tile counter (ranks[0].agg_counter).   */
int gf_fed_signal(const GfFedArgs* args, int phase, void* stream);
#define GF_FED_PHASE_BETA_READY 2
#define GF_FED_PHASE_SMALL_AR 3
#define GF_FED_FLAG_SLOTS 8
#define GF_FED_FLAG_BYTES (GF_FED_FLAG_SLOTS * GF_MAX_RANKS * 4)
/* Writes GF_FED_TMAPS CUtensorMap (128 bytes each, 64-byte aligned HOST buffer): GF_MAX_RANKS maps of every
 * rank's bf16 beta gradient ranks[0].Gb[c] as [K, ld_gb] (entries of absent ranks are zeroed), then three maps of
 * this rank's fp32 beta / first / second moment [K, V]; boxes of 16 topics x 128 words.  The maps are used for
 * plain TMA loads; they are never used with cluster multicast.                                        */
#define GF_FED_TMAPS (GF_MAX_RANKS + 3)
int gf_fed_make_tmaps(const GfFedArgs* args, void* out_tmaps_host);
/* One-shot all-reduce (sum) of a small fp32 buffer x[0..n) (n <= ranks[0].ar_cap, x 16-byte aligned, n % 4 == 0)
 * over peer memory: every rank stores its x into its slot of every rank's staging area, raises flag slot
 * GF_FED_PHASE_SMALL_AR, waits for all ranks and sums the slots in rank order (bit-identical result everywhere).
 * Used for the gradient of CombinedTM's collapsed weight (150 KB instead of the 38 M-element adapt_bert
 * gradient).  Every rank calls it once per step, before gf_fed_aggregate of the same step.                  */
int gf_fed_allreduce_small(const GfFedArgs* args, float* x, long long n, void* stream);
/* Overlapped aggregation keeps the fp32 master copy of beta on the owner of each word-tile range only;
 * this copies every range this rank does not own from its owner's parameter buffer (P2P loads) so that
 * state_dict() / checkpoints see the complete tensor.  Collective in spirit: every rank must be between
 * steps (host barrier before and after).                                                             */
int gf_fed_gather_beta(const GfFedArgs* args, void* stream);

/* ------------------------------------------------------------------------------------------
 * Host-side data loader helpers (no device code).
 * Ref: torch DataLoader over a dense [n_docs, V] array, federated_model.py:82-86.
 * ---------------------------------------------------------------------------------------- */
/* Gather CSR rows `rows[0..n_rows)` of a corpus into staging buffers (pinned memory welcome).  */
int gf_host_gather_csr(const int* indptr, const int* indices, const float* data, const long long* rows,
                       int n_rows, int* out_indptr, int* out_indices, float* out_data,
                       long long out_capacity, int n_threads);
int gf_host_gather_rows(const float* src, long long row_len, const long long* rows, int n_rows, float* out);

#ifdef __cplusplus
}
#endif
#endif /* GFEDNTM_KERNELS_H_ */
