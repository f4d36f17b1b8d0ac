/*
 * collie_b200 -- C ABI of the B200-native matrix-factorization hot path.
 *
 * The reference (ShopRunner/collie v1.3.1, pure Python + stock torch ops) has no FFI seam for this
 * path; every entry point below states which reference call site(s) it stands in for. Paths are
 * relative to the reference root. The host-side mirror of the reference API
 * (collie_b200/model.py, trainer.py, metrics.py, interactions.py) binds these through ctypes; see
 * INTEGRATION.md for the stub a reference maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless named host_*;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - return value 0 on success, negative cmf_status on failure; cmf_last_error() returns a
 *     thread-local message for the last failure on the calling thread;
 *   - the library never allocates device memory: scratch comes from the caller, sized by the
 *     *_workspace_bytes() queries, and initialised once with cmf_workspace_init();
 *   - all tables are fp32 row-major [rows, D] contiguous, exactly the reference's
 *     nn.Embedding.weight layout (collie/model/matrix_factorization.py:106-117), so state_dicts
 *     round-trip; bias tables are [rows, 1].
 */
#ifndef COLLIE_B200_H_
#define COLLIE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CMF_API __attribute__((visibility("default")))
#else
#define CMF_API
#endif

typedef enum {
  CMF_OK = 0,
  CMF_ERR_INVALID = -1,   /* bad argument (message says which) */
  CMF_ERR_CUDA = -2,      /* a CUDA runtime call or launch failed */
  CMF_ERR_WORKSPACE = -3, /* workspace missing or too small */
  CMF_ERR_UNSUPPORTED = -4
} cmf_status;

/* loss selected by collie/model/base/base_pipeline.py:277-340 */
typedef enum {
  CMF_LOSS_BPR = 0,            /* collie/loss/bpr.py:75-89     (one negative)   */
  CMF_LOSS_HINGE = 1,          /* collie/loss/hinge.py:71-85   (one negative)   */
  CMF_LOSS_ADAPTIVE_BPR = 2,   /* collie/loss/bpr.py:153-167                    */
  CMF_LOSS_ADAPTIVE_HINGE = 3, /* collie/loss/hinge.py:157-171                  */
  CMF_LOSS_WARP = 4,           /* collie/loss/warp.py:73-154                    */
  CMF_LOSS_MSE = 5,            /* torch.nn.MSELoss, base_pipeline.py:287-289    */
  CMF_LOSS_MAE = 6             /* torch.nn.L1Loss,  base_pipeline.py:290-292    */
} cmf_loss;

/* optimizer built by collie/model/base/base_pipeline.py:406-437 */
typedef enum {
  CMF_OPT_SPARSE_ADAM = 0, /* torch.optim.SparseAdam: every gathered row moves, even with g == 0 */
  CMF_OPT_SGD = 1,         /* torch.optim.SGD (no momentum)                                      */
  CMF_OPT_ADAM = 2,        /* torch.optim.Adam on dense grads: EVERY row of the table moves      */
  CMF_OPT_ADAGRAD = 3      /* torch.optim.Adagrad (lr_decay 0, eps 1e-10)                        */
} cmf_opt_rule;

typedef struct {
  int32_t rule;        /* cmf_opt_rule */
  int32_t reserved;
  double lr;           /* hyper-parameters are doubles, as the Python floats torch receives */
  double beta1, beta2; /* Adam family */
  double eps;
  double weight_decay; /* dense rules only (the reference forces 0 when sparse, base_pipeline.py:230-240) */
  int64_t step;        /* 1-based step count of this parameter group (torch keeps it per tensor) */
} cmf_opt_t;

#define CMF_MAX_METADATA 8

/*
 * One training (or validation) step of MatrixFactorizationModel on one batch:
 *   negatives (injected or drawn on device) -> forward -> loss -> backward -> optimizer updates.
 * Stands in for one iteration of CollieMinimalTrainer._train_loop_single_epoch
 * (collie/model/base/trainer.py:458-470) = BasePipeline.calculate_loss (base_pipeline.py:582-654)
 * + MatrixFactorizationModel.forward x2 (matrix_factorization.py:120-159) + collie/loss/*.py
 * + autograd backward + optimizer.step() for both optimizers, and -- when negs_in is NULL -- for
 * Interactions._negative_sample (collie/interactions/datasets.py:390-445).
 */
typedef struct {
  /* parameter tables and optimizer state (state pointers may be NULL when the rule has none) */
  float *U, *I, *bu, *bi;
  float *U_m, *U_v, *I_m, *I_v;
  float *bu_m, *bu_v, *bi_m, *bi_v;
  int64_t num_users, num_items;
  int32_t D;

  /* batch: B interactions. implicit: (users, items) + negatives; explicit: + ratings */
  const int32_t *users;   /* [B] */
  const int32_t *items;   /* [B] positive (implicit) or rated (explicit) item ids */
  const float *ratings;   /* [B], explicit losses only */
  const int64_t *negs_in; /* [B, N] injected negatives, or NULL to draw them on device */
  const int64_t *gidx;    /* [B] dataset position of each interaction (RNG counter), or NULL */
  int64_t gidx_base;      /* used as gidx_base + b when gidx == NULL */
  int64_t B;
  int32_t N;              /* negatives per interaction (0 for explicit) */
  int32_t users_sorted;   /* nonzero: `users` is already non-decreasing, skip the internal sort */

  /* on-device sampler (counter-based Philox4x32-10) */
  uint64_t seed;
  uint32_t epoch;
  const int64_t *csr_indptr; /* [num_users + 1] positives CSR for exact sampling, or NULL */
  const int32_t *csr_items;  /* sorted item ids per user row */
  int32_t max_tries;         /* max_number_of_samples_to_consider, datasets.py:291 */

  /* loss */
  int32_t loss; /* cmf_loss */
  int32_t has_y_range;
  float y_lo, y_hi;              /* matrix_factorization.py:152-157 */
  int32_t n_meta;                /* partial-credit metadata keys, metadata_utils.py:60-71 */
  const int32_t *meta[CMF_MAX_METADATA]; /* each [num_items] categorical labels */
  float meta_w[CMF_MAX_METADATA];

  /* optimizers: embeddings and biases are separate groups (base_pipeline.py:450-467) */
  cmf_opt_t emb_opt, bias_opt;
  int32_t train; /* 0: loss only (validation step, trainer.py:487-504) */

  /* outputs */
  float *loss_out;       /* [1] */
  int64_t *negs_out;     /* [B, N] negatives actually used, or NULL */
  int32_t *sel_out;      /* [B] selected negative slot (adaptive: argmax; warp: first violation or
                            N when none; bpr/hinge: 0), or NULL */
  float *pos_scores_out; /* [B] or NULL */
  float *neg_scores_out; /* [B, N] or NULL */

  void *workspace;
  size_t workspace_bytes;

  /* multi-GPU (one process per GPU; 0 / 0 on a single GPU):
   * B_global: interactions in the global batch; the loss is a mean over it, so gradients and the
   * returned loss are scaled by 1 / B_global and the ranks' losses add up to the global loss.
   * skip_item_update: leave the item-row gradients (and item "touched" flags) in the workspace
   * for the caller to reduce across ranks (cmf_train_workspace_items) instead of applying them. */
  int64_t B_global;
  int32_t skip_item_update;
  /* number of real items when the item table is padded to a multiple of the rank count: negatives
   * are drawn from [0, num_items_logical) and WARP weights use it (0 = num_items) */
  int32_t num_items_logical;
  /* optional cudaEvent_t recorded on the stream once the item-side accumulators of this step are
   * complete (after k_bwd_update / the fused kernel, before the row-update kernel): with
   * skip_item_update a caller can start the exchange on another stream while the remaining
   * user-row updates run */
  void *item_grads_ready_event;
  /* dropout on both embeddings of every scored pair, training mode only
   * (collie/model/matrix_factorization.py:118, :147): each element of the user row and of the
   * item row of every (interaction, positive-or-negative) pair is zeroed with probability
   * dropout_p and scaled by 1 / (1 - dropout_p) otherwise, independently. The masks come from
   * the counter-based stream of the negatives (other key), keyed by (seed, emb_opt.step, epoch,
   * dataset position, pair, side, element), so a step can be replayed. 0 = off. */
  float dropout_p;
  /* non-zero: bit-repeatable step. The item-row gradients are summed in batch order from a
   * radix-sorted (item, contribution) list instead of vector RED in arrival order, and every user
   * is summed by one warp (chunks cut at user boundaries), so two runs of the same steps give the
   * same bits whatever the schedule, chunk length or GPU. Slower (one sort of 2 B pairs and a
   * serial sum per item row per step). Needs embedding_dim % 4 == 0 and dropout_p == 0. With
   * skip_item_update the rank's accumulators are repeatable; the sum over the ranks is so only
   * with the plain (non-multicast) exchange. The reference's CPU path is deterministic, torch's
   * CUDA index_add is not (no counterpart of this switch in collie). */
  int32_t deterministic;
} cmf_step_t;

/* scratch for cmf_train_step; depends only on the table shapes and the largest batch */
CMF_API size_t cmf_train_workspace_bytes(int64_t max_B, int32_t max_N, int64_t num_users, int64_t num_items,
                                 int32_t D);
/* Views of the item-side accumulators inside a training workspace (layout depends only on the
 * table shapes): GI [num_items, D] fp32 gradient rows, gbi [num_items] fp32 bias gradients,
 * item_flag [num_items] int32 (non-zero = the row was gathered this step). With
 * skip_item_update the caller sums them over ranks, applies cmf_optimizer_rows_masked to the rows
 * it owns and zeroes all three again before the next step. */
CMF_API int cmf_train_workspace_items(void *workspace, int64_t num_users, int64_t num_items, int32_t D,
                              float **GI, float **gbi, int32_t **item_flag);

/* zero the persistent part of the workspace; call once after allocation (and after a failure) */
CMF_API int cmf_workspace_init(void *workspace, size_t workspace_bytes, void *stream);

CMF_API int cmf_train_step(const cmf_step_t *args, void *stream);

/*
 * All steps of one epoch in one call: `first` describes step 0 (B = the batch size; users / items /
 * ratings / gidx / negs_in point at arrays of n_total entries laid out batch after batch, e.g. by
 * cmf_epoch_plan*); step s takes entries [s B, min((s + 1) B, n_total)), the optimizers' step
 * counters advance by one per step, losses[s] receives the loss of step s (device memory,
 * ceil(n_total / B) floats). Replaces the for-loop of
 * CollieMinimalTrainer._train_loop_single_epoch, collie/model/base/trainer.py:451-478.
 */
CMF_API int cmf_train_epoch(const cmf_step_t *first, int64_t n_total, float *losses, void *stream);
/* same call, restricted to the implicit / explicit loss families (base_pipeline.py:603 / :635) */
CMF_API int cmf_train_step_implicit(const cmf_step_t *args, void *stream);
CMF_API int cmf_train_step_explicit(const cmf_step_t *args, void *stream);

/*
 * preds[m] = <U[users[m]], I[items[m]]> + bu[..] + bi[..]  (+ sigmoid range if y_range != NULL).
 * Replaces MatrixFactorizationModel.forward, collie/model/matrix_factorization.py:120-159.
 * idx_bytes = 4 (int32 ids) or 8 (int64 ids, what the reference passes after .long()).
 */
CMF_API int cmf_forward_scores(const float *U, const float *I, const float *bu, const float *bi, int32_t D,
                       const void *users, const void *items, int32_t idx_bytes, int64_t M,
                       const float *host_y_range /* NULL or host float[2] */, float *out,
                       void *stream);

/*
 * Backward of cmf_forward_scores: accumulates (atomically) into dense, caller-zeroed gradient
 * tables gU [num_users,D], gI [num_items,D], gbu, gbi given dL/dpreds. `preds` (the forward output)
 * is needed only when y_range is set. Replaces autograd through matrix_factorization.py:143-157
 * (embedding_dense_backward x4), triggered at collie/model/base/trainer.py:463.
 */
CMF_API int cmf_backward_scores(const float *U, const float *I, int32_t D, const void *users,
                        const void *items, int32_t idx_bytes, int64_t M, const float *dpreds,
                        const float *preds, const float *host_y_range, float *gU, float *gI,
                        float *gbu, float *gbi, void *stream);

/*
 * out[r] = cosine similarity of row r with row id of the table E [rows, D] (each row divided by
 * its L2 norm, then dotted). Replaces BasePipeline._calculate_embedding_similarity,
 * collie/model/base/base_pipeline.py:849-859 (item_item_similarity :783-815, user_user_similarity
 * :817-847).
 */
CMF_API int cmf_row_similarity(const float *E, int64_t rows, int32_t D, int64_t id, float *out,
                       void *stream);

/*
 * Negative item ids [B, N] for interactions at dataset positions gidx (or gidx_base + b).
 * Replaces Interactions._negative_sample, collie/interactions/datasets.py:390-445: approximate mode
 * when csr_indptr == NULL (:433-443), else the exact rejection loop (:399-426).
 */
CMF_API int cmf_sample_negatives(uint64_t seed, uint32_t epoch, const int64_t *gidx, int64_t gidx_base,
                         const int32_t *users, int64_t B, int32_t N, int64_t num_items,
                         const int64_t *csr_indptr, const int32_t *csr_items, int32_t max_tries,
                         int64_t *negs_out, void *stream);

/*
 * Epoch plan of a data loader, on the device: one pass over the n dataset positions as a random
 * permutation (shuffle != 0; keyed by (seed, epoch)) cut into batches of `batch`, every batch
 * ordered by user when sort_within_batch != 0 (ties keep the permutation's order). Outputs, all
 * of length n: the users / items (/ ratings) in plan order and out_gidx = the dataset position of
 * every entry (gpos[position] when gpos != NULL: positions in a global, unsharded dataset), which
 * keys the negative sampler. n < 2^31.
 * Replaces the host-side shuffle + batch slicing of the reference's loaders:
 * collie/interactions/samplers.py:51-60 (index shuffle) under dataloaders.py:70-173.
 */
CMF_API size_t cmf_epoch_plan_workspace_bytes(int64_t n);
CMF_API int cmf_epoch_plan(const int32_t *users, const int32_t *items, const float *ratings,
                   const int64_t *gpos, int64_t n, int64_t batch, int64_t num_users,
                   int32_t shuffle, int32_t sort_within_batch, uint64_t seed, uint32_t epoch,
                   int32_t *out_users, int32_t *out_items, float *out_ratings, int64_t *out_gidx,
                   void *workspace, size_t workspace_bytes, void *stream);

/*
 * The same plan (shuffled, every batch ordered by user) when the dataset is handed over ordered by
 * user: rec[p] = {user, item, gidx low word, gidx high word} (16 bytes) of the p-th entry in user
 * order, ratings[p] alongside (or NULL). A keyed bijection of [0, n) sends entry p to batch
 * sigma(p) / batch and one stable radix partition by batch number leaves every batch ordered by
 * user (equal users in dataset order): no sort by user per epoch. Same workspace size.
 */
CMF_API int cmf_epoch_plan_presorted(const void *rec, const float *ratings, int64_t n, int64_t batch,
                             uint64_t seed, uint32_t epoch, int32_t *out_users,
                             int32_t *out_items, float *out_ratings, int64_t *out_gidx,
                             void *workspace, size_t workspace_bytes, void *stream);

/*
 * Row-wise optimizer on a table given a dense gradient table G [rows, D]:
 *   rows == NULL: every row (torch.optim.Adam / SGD / Adagrad on dense grads);
 *   rows != NULL: only the n_rows listed unique rows (torch.optim.SparseAdam on the coalesced
 *   gradient; SGD / Adagrad on sparse grads). Used rows of G are reset to zero when zero_grad != 0.
 * Replaces optimizer.step() reached from collie/model/base/trainer.py:465-470 for the optimizers
 * built at base_pipeline.py:418-433.
 */
CMF_API int cmf_optimizer_rows(float *P, float *M, float *V, float *G, int64_t num_rows, int32_t D,
                       const int32_t *rows, int64_t n_rows, const cmf_opt_t *opt, int32_t zero_grad,
                       void *stream);

/*
 * cmf_optimizer_rows restricted to the rows whose touched[row] != 0 (all other rows are left
 * alone unless the rule is dense); biases ride along: bP/bM/bV/bG are [num_rows] (bG may be NULL
 * to skip them). This is the owner-side update of a row-sharded table after the ranks'
 * gradients have been reduced: SparseAdam moves every row some rank gathered, even with a zero
 * gradient (torch.optim.SparseAdam on the coalesced sparse gradient, base_pipeline.py:430-433).
 */
CMF_API int cmf_optimizer_rows_masked(float *P, float *M, float *V, const float *G, float *bP, float *bM,
                              float *bV, const float *bG, int64_t num_rows, int32_t D,
                              const int32_t *touched, const cmf_opt_t *opt,
                              const cmf_opt_t *bias_opt, void *stream);

/*
 * Item-side exchange of a data-parallel step (users partitioned over the ranks, item table
 * row-sharded in contiguous blocks of rows_per_rank), done by the owner of each row in ONE kernel
 * over peer memory: sum the ranks' gradient rows (GI / gbi, read out of every rank's training
 * workspace, only where that rank's item_flag is set), apply the rule to the owned rows (state
 * M / V / bM / bV are [rows_per_rank, D] / [rows_per_rank], local), store the new rows into every
 * rank's table copy I / bi and clear the accumulators that were consumed. All pointers of rank q
 * must be addresses valid in THIS process (peer mappings), at the same offset of equally sized
 * buffers when the multicast addresses are given. The caller orders it across ranks:
 * all ranks' cmf_train_step(skip_item_update) before, all ranks' next step after.
 * Replaces the gradient all-reduce + replicated optimizer.step of Lightning DDP around
 * collie/model/base/trainer.py:463-470.
 */
#define CMF_MAX_PEERS 8
typedef struct cmf_peers_t {
  int32_t rank, world;
  int32_t D, reserved;
  int64_t rows_per_rank;
  float *I[CMF_MAX_PEERS], *bi[CMF_MAX_PEERS];
  float *GI[CMF_MAX_PEERS], *gbi[CMF_MAX_PEERS];
  int32_t *flag[CMF_MAX_PEERS];
  /* NVSwitch multicast addresses of the same five buffers (one address = that offset in every
   * rank's buffer), or all NULL: with them the sums are multimem.ld_reduce (reduced in the
   * switch) and the stores multimem.st, so the kernel's NVLink traffic does not grow with world */
  float *mc_I, *mc_bi, *mc_GI, *mc_gbi;
  int32_t *mc_flag;
  /* Rendezvous inside the kernel (all NULL / 0: the caller fences, and the owner clears the
   * accumulators it consumed on every rank). pad[q] = rank q's signal pad, cmf_exchange_pad_bytes()
   * bytes of zero-initialised peer-mapped memory; epoch = 1, 2, 3, ... one value per call, the
   * same on every rank. The kernel then (1) waits until every rank has entered it -- so it must
   * be stream-ordered after this rank's cmf_train_step(skip_item_update) --, (2) leaves only when
   * every rank's stores have landed everywhere, and (3) clears [clear_ptr, clear_ptr +
   * clear_bytes) = this rank's own accumulators (item_flag .. gbi of its workspace) itself. Its
   * grid is sized to be resident at once; a peer that never arrives makes it trap after 20 s
   * instead of hanging the GPU. */
  uint32_t *pad[CMF_MAX_PEERS];
  uint32_t epoch, reserved2;
  void *clear_ptr;
  size_t clear_bytes;
} cmf_peers_t;
CMF_API size_t cmf_exchange_pad_bytes(void);
CMF_API int cmf_item_exchange_update(const cmf_peers_t *peers, float *M, float *V, float *bM, float *bV,
                             const cmf_opt_t *opt, const cmf_opt_t *bias_opt, void *stream);

/*
 * Evaluation of one tile of users against the whole item table (or an item shard):
 * scores S = U[user_ids] . I^T + bu + bi are not returned (precision 1 never writes them to
 * memory; precision 0 keeps the tile's scores in the caller's workspace between its scoring and its
 * ranking kernel); per user the call keeps the
 * top-k (score descending, item id ascending), and -- given the users' target items in CSR form --
 * the MAP@k / MRR / AUC ingredients. Replaces get_preds + mapk + mrr + auc,
 * collie/metrics.py:77-107, 144-198, 201-240, 243-282, for one iteration of the loop at :379-384.
 *
 *   tgt_indptr [n_users_tile + 1], tgt_items: sorted target item ids per user of the tile (global
 *   ids); item_offset/num_items_shard describe the item shard this call scores (0 / num_items for
 *   the unsharded case).
 * Outputs (any may be NULL): topk_ids [n,k] int32 (global ids, -1 padded), topk_scores [n,k],
 *   ap [n] double (AP@k per user, fp64 like the reference's weights), n_before_best [n] int32: #items ranked before the user's best-ranked
 *   target in this shard, best_score/best_id [n]: that target's score/id (-inf/-1 when none),
 *   auc_pairs [n] double: sum over targets of (#non-targets with lower sigmoid + .5 ties).
 * precision: 0 = fp32 SIMT scoring (exact fp32; all outputs), 1 = tcgen05 tensor cores select
 *   (bf16 operands, fp32 accumulation, a stated error bound) and fp32 decides (every candidate of
 *   the top-k and every item whose order against the best target is in doubt is re-scored in
 *   the exact path's summation order): embedding_dim <= 128, k <= 16, every output except
 *   auc_pairs (want_auc must be 0); ids, scores, ap and n_before_best equal precision 0 bit for bit.
 */
typedef struct {
  const float *U, *I, *bu, *bi;
  int32_t D;
  const int32_t *user_ids; /* [n_users_tile] */
  int64_t n_users_tile;
  int64_t item_offset, num_items_shard;
  int32_t k;
  const int64_t *tgt_indptr;
  const int32_t *tgt_items;
  int32_t want_mrr, want_auc;
  int32_t precision;
  /* item-sharded second pass (all NULL otherwise): count against the globally best target and
   * the globally gathered target sigmoids instead of this shard's own */
  const float *ext_best_score;
  const int32_t *ext_best_id;
  int32_t *topk_ids;
  float *topk_scores;
  double *ap;
  int32_t *n_before_best;
  float *best_score;
  int32_t *best_id;
  double *auc_pairs;
  void *workspace;
  size_t workspace_bytes;
  /* optional: rank an already materialised score tile [n_users_tile, num_items_shard] instead of
   * scoring (the metric callables' `preds` argument, collie/metrics.py:144-147); U/I/bu/bi and
   * user_ids are then ignored */
  const float *scores_in;
  /* precision = 1 only: non-zero = the item table, the shard and want_mrr are those of the
   * previous call with this workspace, so the bf16 copy of the items in the workspace is reused */
  int32_t reuse_item_prep;
  /* precision = 1 only: bf16 terms of the operands, 0 = automatic (1 without the MRR counter, 3
   * with it), 1 or 3 = as said (3 costs three times the MMA work and settles users whose top
   * scores lie closer together than the one-term bound can tell; results are exact either way;
   * part of what reuse_item_prep assumes unchanged) */
  int32_t tc_terms;
  /* precision = 1 only, optional: int32[2] on the device, ADDED to by every call: [0] users whose
   * candidates the error bound could not separate and that were therefore ranked by the exact
   * path inside the call (results are exact either way), [1] such users beyond what the
   * workspace can take over in one call -- their outputs are NOT written: the caller repeats the
   * tile with precision = 0 when it finds [1] != 0 */
  int32_t *tc_status;
  /* optional, with tc_status (then int32[4]): the users of [1] are listed for the caller instead
   * of being lost -- tc_overflow_rows[tc_status[2]++] = tc_row_base + (row of the tile) -- so that
   * only they are repeated with precision = 0 (tc_overflow_cap entries at most) */
  int32_t *tc_overflow_rows;
  int64_t tc_row_base;
  int32_t tc_overflow_cap;
  int32_t reserved4_;
} cmf_eval_t;

CMF_API size_t cmf_eval_workspace_bytes(int64_t n_users_tile, int64_t num_items_shard, int32_t D, int32_t k,
                                int32_t precision);
CMF_API int cmf_eval_tile(const cmf_eval_t *args, void *stream);
/* Facts about the precision = 1 path for measurement: out[0] / out[1] = K extent the GEMM issues
 * per output with one bf16 term (top-k / MAP@k only) / three terms (with the MRR counter);
 * out[2] / out[3] = the error-bound coefficients c1 / c3. */
CMF_API int cmf_eval_tc_info(int32_t D, double *out);

/*
 * cmf_eval_tile without a caller-provided workspace. (No counterpart in the reference: Collie's
 * evaluation allocates freely through torch.)
 *
 * For hosts that cannot hand the library device memory: a serving or monitoring process that was
 * given parameter tables and output buffers by another runtime but has no allocator of its own on
 * the GPU, or one whose framework has already reserved the whole device so a further cudaMalloc
 * would fail or stall the device. The scratch is a fixed 16 MiB array inside the library image
 * (every GPU that loads the library holds it until the process exits); `workspace` and
 * `workspace_bytes` in `args` are ignored. Inputs and outputs are still the caller's.
 *
 *   - Limit: cmf_eval_workspace_bytes(n_users_tile, num_items_shard, ...) must be <= 16 MiB
 *     (12 bytes per user x item, e.g. 52 users against 26,744 items); larger tiles are refused with
 *     CMF_ERR_WORKSPACE and a message that names both sizes.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress);
 *     make that first call before any stream capture begins.
 *   - The call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *   - All calls share the one array and it keeps whatever the previous call left in it. Nothing in
 *     it has to start at zero (the score tile and the sigmoid buffers are written before they are
 *     read), but two calls whose kernels overlap on the GPU -- different streams, or different
 *     host threads -- race on it: serialise them.
 */
CMF_API int cmf_eval_tile_noalloc(const cmf_eval_t *args, void *stream);

/*
 * Dataset ingest on the device (collie/interactions/datasets.py:81-155, 359-366): of every
 * duplicated (user, item) pair the FIRST position and the LAST value are kept, as the reference's
 * dict-backed construction does -- keep_pos[j] / keep_val[j] (j < *n_out, dataset order; both
 * arrays have room for n) are the position whose (user, item) and the position whose rating
 * entry j of the de-duplicated dataset takes -- and the positives CSR falls out of the same sort:
 * indptr [num_users + 1], csr_items [n] (the first *n_out are valid: distinct items per user,
 * ascending). n_out, indptr, csr_items are device memory; indptr / csr_items may both be NULL.
 */
CMF_API size_t cmf_ingest_workspace_bytes(int64_t n);
CMF_API int cmf_ingest_pairs(const int32_t *users, const int32_t *items, int64_t n, int64_t num_users,
                     int64_t num_items, int32_t *keep_pos, int32_t *keep_val, int64_t *n_out,
                     int64_t *indptr, int32_t *csr_items, void *workspace, size_t workspace_bytes,
                     void *stream);

/*
 * Forward / backward of the metric-learning model's score (collie/model/
 * collaborative_metric_learning.py:100-124): out[r] = || U[users[r]] - I[items[r]] + 1e-6 ||_2
 * (torch.nn.functional.pairwise_distance: both gathers and the reduction in one kernel), and
 * gU[users[r]] += grad_out[r] * (U[u] - I[i] + 1e-6) / dist[r], gI[items[r]] -= the same (dense
 * gradient tables, zero where dist is 0; either may be NULL). ids: int32 or int64.
 */
CMF_API int cmf_pairwise_distance(const float *U, const float *I, int32_t D, const void *users,
                          const void *items, int32_t idx_bytes, int64_t n, float *out,
                          void *stream);
CMF_API int cmf_pairwise_distance_backward(const float *U, const float *I, int32_t D, const void *users,
                                   const void *items, int32_t idx_bytes, int64_t n,
                                   const float *grad_out, const float *dist, float *gU, float *gI,
                                   void *stream);

/*
 * Item-sharded evaluation, merge step (distributed.py evaluate_item_sharded): ids / scores are the
 * all-gathered per-shard top-k lists [n_shards][n_users][k] (ids < 0 = padding; the shards score
 * disjoint item blocks), merged per user into the global top-k in the canonical order (score
 * descending, item id ascending); AP@k (collie/metrics.py:176-196) against the users' sorted
 * target ids (tgt_indptr [n_users + 1] absolute offsets into tgt_items). Outputs may be NULL.
 * The reference scores the whole row on one device (metrics.py:144-198); the merge has no
 * counterpart there. k <= 32.
 */
CMF_API int cmf_merge_topk(const int32_t *ids, const float *scores, int32_t n_shards, int64_t n_users,
                   int32_t k, const int64_t *tgt_indptr, const int32_t *tgt_items,
                   int32_t *out_ids, float *out_scores, double *ap, void *stream);

/*
 * Row gather into a column block of a wider matrix, and its transpose (vector RED): the input of
 * the first dense layer of HybridPretrainedModel, X[r] = [user_meta[u] | U[u] | I[i] | item_meta[i]],
 * is written by one call per source table (out + col0 with row pitch ld floats), instead of the
 * reference's four gathers + torch.cat (collie/model/hybrid_pretrained_matrix_factorization.py
 * :401-464); cmf_scatter_add_rows is the embedding backward of the same (grad_table[ids[r]] +=
 * grad_out[r, col0 : col0 + D]). ids: int32 or int64 (idx_bytes 4 / 8).
 */
CMF_API int cmf_gather_rows(const float *table, int64_t num_rows, int32_t D, const void *ids,
                    int32_t idx_bytes, int64_t n, float *out, int64_t ld, int32_t col0,
                    void *stream);
CMF_API int cmf_scatter_add_rows(float *grad_table, int64_t num_rows, int32_t D, const void *ids,
                         int32_t idx_bytes, int64_t n, const float *grad_out, int64_t ld,
                         int32_t col0, void *stream);

/* Materialised score tile out[n, num_items] (for user-supplied metric callables and get_preds,
 * collie/metrics.py:77-107). */
CMF_API int cmf_score_tile(const float *U, const float *I, const float *bu, const float *bi, int32_t D,
                   const int32_t *user_ids, int64_t n_users_tile, int64_t num_items, float *out,
                   void *stream);

/* Measurement helper (no reference counterpart): reads `bytes` of `buf` `repeats` times with a
 * persistent grid of 16-byte loads; timed by the caller, it gives the read bandwidth of L2 (buffer
 * below the L2 size) or of HBM (buffer far above it) that bench.py quotes next to its kernels. */
CMF_API int cmf_read_probe(const void *buf, size_t bytes, int32_t repeats, float *out, void *stream);
/* Same idea for the training step's access pattern: reads the 512-byte rows idx[0..n_idx) of
 * `table` ([rows, 128] fp32), `unroll` rows in flight per warp, layout 0 = one row per warp
 * instruction, 1 = the 8-lanes-per-row layout of the forward kernel. */
/* ... and for the backward kernel's item-gradient pattern: vector atomic adds (RED) of 512-byte
 * rows into `table` at idx[0..n_idx). */
CMF_API int cmf_red_probe(void *table, const int32_t *idx, int64_t n_idx, int32_t ctas_per_sm,
                  void *stream);
CMF_API int cmf_gather_probe(const void *table, const int32_t *idx, int64_t n_idx, int32_t unroll,
                     int32_t layout, int32_t ctas_per_sm, float *out, void *stream);

/* Tuning knobs (process-wide). key 0: chunk length L of the fused step (0 = automatic); key 1:
 * widest lane group per row in the generic fused kernel (8, 16 or 32); key 2: kernel choice
 * (0 automatic, 1 generic fused kernel, 2 split k_fwd_loss + k_bwd_update); key 3: experiments on
 * the tensor-core evaluation kernel (0 = off; 1 = skip the ranking, 2 = skip the inserts: timing
 * only, results are then meaningless); key 4: bf16 terms of the tensor-core evaluation (0 =
 * automatic: 1 without the MRR counter, 3 with it; 1 or 3 = forced, results stay exact); key 5: candidate-list length of the tensor-core
 * evaluation (0 = automatic, 8 / 16 / 24; must exceed k); key 7: its epilogue warps (0 = automatic, 8 or 16); key 6: programmatic dependent launch of k_bwd_update and
 * k_update_rows (0 = plain stream order, the default; 1 = on: 1 % faster steps when nothing else
 * shares the GPU, slower pipelines when other streams need SMs between the step's kernels). key 8: the item-row gradient contributions are sent by k_fwd_loss_g8 (which holds the user row) instead of k_bwd_update (0 / 1; also CMF_RED_IN_FWD). Results do not depend on keys 0-2, 6 and 8 beyond fp32 summation
 * order; tests use them to cross-check the in-place and the accumulate paths. Returns the previous
 * value, or a negative cmf_status for an unknown key. */
CMF_API int cmf_set_tuning(int32_t key, int32_t value);

/* Phase timing of cmf_train_step for bench.py's roofline line. While enabled, every step records
 * CUDA events on its launch stream between its phases. cmf_profile_read waits for them and returns
 * the accumulated milliseconds of the four intervals, which path ran (*host_split = 1: split
 * kernels, intervals = k_fwd_loss | rest of the overlapped sort | k_bwd_update | k_update_rows;
 * 0: generic kernel, intervals = sort | k_fused_step | - | k_update_rows), the number of steps
 * since the last read and the touched-row list lengths of the LAST step; then clears the record.
 * Not for use inside the timed region of a throughput measurement. */
CMF_API int cmf_profile_enable(int32_t on);
CMF_API int cmf_profile_read(double *host_ms /* [4] */, int32_t *host_split, int64_t *host_steps,
                     int64_t *host_last_item_rows, int64_t *host_last_user_rows);

CMF_API const char *cmf_last_error(void);
CMF_API int cmf_version(void);
/* number of kernels this library has launched since load (process-wide, all streams) */
CMF_API int64_t cmf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* COLLIE_B200_H_ */
