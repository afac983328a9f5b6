/*
 * wxddp.h -- C ABI of libwxddp.so: the B200 (sm_100a) implementation of the one data-parallel hot
 * path of MustaphaMensouri/Distributed-Continuous-Training-with-Airflow-PyTorch-Distributed-DDP-:
 * the training step of jobs/train_lightning_ddp.py plus the standardisation / label encoding of
 * jobs/preprocess.py that feeds it.
 *
 * The reference has no FFI of its own (it is pure Python over torch / Lightning / Spark wheels);
 * each entry point below names the reference lines whose arithmetic it replaces.  Citations are
 * relative to /root/reference/jobs/.
 *
 * Conventions
 *   - extern "C"; every call returns 0 (WX_OK) or a negative WX_E* code; wx_last_error() gives a
 *     thread-local message.  Nothing throws, nothing calls exit().
 *   - All tensors are CALLER-OWNED DEVICE pointers unless a parameter says "host".  They must stay
 *     alive until the work queued on `stream` has completed.
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream).  Launch
 *     functions are asynchronous; none of them synchronises the device or allocates memory.
 *   - wx_ctx is an opaque per-process / per-GPU object that owns only scratch memory (partial-sum
 *     slabs, tickets).  One ctx per GPU; a ctx is not thread-safe; calls that share a ctx must be
 *     ordered on one stream.
 *   - Flat parameter layout (fp32, P = 64*F + 64 + 128 + 2 values), identical to the reference
 *     module's state_dict order (train_lightning_ddp.py:57-62):
 *         W1[64][F] (net.0.weight) | b1[64] (net.0.bias) | W2[2][64] (net.3.weight) | b2[2] (net.3.bias)
 *   - There is no CPU fallback anywhere in this library.
 */
#ifndef WXDDP_H_
#define WXDDP_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WX_OK 0
#define WX_EINVAL (-1)   /* bad argument (message says which) */
#define WX_ECUDA (-2)    /* a CUDA runtime call failed */
#define WX_ENOMEM (-3)   /* scratch too small for the request */
#define WX_EARCH (-4)    /* device is not compute capability 10.x */

#define WX_HIDDEN 64     /* train_lightning_ddp.py:58 */
#define WX_CLASSES 2     /* train_lightning_ddp.py:61 */
#define WX_MAX_FEATURES 32

/* column element types accepted by the preprocessing kernels */
#define WX_F32 0
#define WX_F64 1

/* null handling in wx_normalize_pack / wx_train_step_raw */
#define WX_IMPUTE_NONE 0 /* reference behaviour: null stays null -> NaN feature (preprocess.py:38-41) */
#define WX_IMPUTE_MEAN 1 /* extension: null -> column mean, i.e. 0.0f after standardisation */

/* dropout source in the training step (train_lightning_ddp.py:60) */
#define WX_DROPOUT_OFF 0    /* eval semantics */
#define WX_DROPOUT_MASK 1   /* caller supplies keep bits: uint64 per row, bit j = keep hidden unit j */
#define WX_DROPOUT_PHILOX 2 /* in-kernel Philox4x32-10 keyed by (seed, step, row): two calls per row, eight base-5 digits per
                             * word, keep <=> digit != 0 (p_keep = 0.8); the stream is specified in oracle/philox_ref.py */

typedef struct wx_ctx wx_ctx;

typedef struct wx_model_desc {
    int32_t n_features; /* F, 1..32   (input_dim, train_lightning_ddp.py:52,125) */
    int32_t hidden;     /* must be 64 (:58) */
    int32_t classes;    /* must be 2  (:61) */
    float dropout_p;    /* 0.2        (:60); only 0.2 and 0.0 are supported */
    double lr;          /* 0.01       (:88) -- doubles, as torch holds them (Python floats) */
    double beta1;       /* 0.9   torch.optim.Adam default used by :88 */
    double beta2;       /* 0.999 */
    double eps;         /* 1e-8 */
    int32_t optimizer;  /* WX_OPT_ADAM (:88, the reference's) or WX_OPT_SGD */
} wx_model_desc;

/* update rule applied by wx_adam_step and by the tails of the fused step kernels */
#define WX_OPT_ADAM 0 /* torch.optim.Adam(lr), single-tensor op order (train_lightning_ddp.py:88) */
#define WX_OPT_SGD 1  /* p -= lr * g (torch.optim.SGD without momentum); no counterpart in the reference, which only uses Adam --
                       * the m / v buffers are left untouched, the step counter still advances */

/* ---- library / context --------------------------------------------------------------------- */
int wx_version(void);               /* 10000*major + 100*minor + patch */
const char* wx_build_arch(void);    /* "sm_100a" */
const char* wx_last_error(void);    /* thread-local, never NULL */

/* Creates the per-GPU context on `device` and its scratch (one cudaMalloc, here and only here). */
int wx_create(int device, const wx_model_desc* desc, wx_ctx** out);
int wx_destroy(wx_ctx* ctx);
int64_t wx_n_params(const wx_ctx* ctx);   /* P */
int wx_num_sms(const wx_ctx* ctx);
/* Number of kernels this context has launched since creation (bench.py's gpu_launches claim). */
int64_t wx_launch_count(const wx_ctx* ctx);

/* ---- multi-GPU: gradient exchange inside the step kernel ------------------------------------------------ */
/* Replaces DDP's bucket all-reduce and the sync_dist loss reduce (train_lightning_ddp.py:70,:136) for the fused steps:
 * after wx_peer_connect, wx_tc_train_step / wx_tc_train_step_raw / wx_train_steps_small sum the [P+1] bucket over all
 * ranks INSIDE the kernel (8-byte {value, tag} words stored into every rank's inbox over NVLink peer memory, ordered sum)
 * and scale by 1/(world*batch).  One process per GPU on one node; every rank must launch the same steps.
 *   wx_peer_export : allocates this rank's inbox (cudaMalloc, once) and writes its 64-byte CUDA IPC handle to handle_out (host).
 *   wx_peer_connect: handles = host [world][64] bytes gathered from all ranks (any transport, e.g. torch.distributed).
 * Protocol: export on every rank, gather, connect on every rank, then a BARRIER across the ranks before the first step. */
int wx_peer_export(wx_ctx* ctx, void* handle_out);
int wx_peer_connect(wx_ctx* ctx, int rank, int world, const void* handles);

/* ---- K1: preprocessing (jobs/preprocess.py) ------------------------------------------------ */

/* Label encode: labels_out[i] = 1 iff string i is valid and equals "rain" exactly, else 0.
 * Replaces `when(col("Rain") == "rain", 1).otherwise(0)` (preprocess.py:23-25) and the int64 widening
 * at train_lightning_ddp.py:46.  Strings are in Arrow utf8 layout: int32 offsets[n+1], byte data,
 * optional LSB-first validity bitmap (NULL = all valid).  labels32_out may be NULL. */
int wx_label_encode(wx_ctx* ctx, const int32_t* offsets, const uint8_t* data, const uint8_t* validity,
                    int64_t n, int64_t* labels_out, int32_t* labels32_out, void* stream);

/* Column statistics: for each of the F columns, over non-null rows, in float64:
 *   stats_out[j] = { count, mean, M2 = sum (x-mean)^2, std_val }
 * std_val = sqrt(M2/(count-1)) with the reference's zero guard (`std if std != 0 else 1.0`), NaN when
 * count < 2 (Spark returns null there).  Replaces `df.select(mean(c), stddev(c)).first()` and the guard
 * (preprocess.py:32-36) -- one pass over all columns instead of F Spark jobs.
 * cols: HOST array of F DEVICE pointers to columns of `dtype`; a cell is null if it is NaN or if
 * validity (HOST array of F device bitmaps, or NULL, entries may be NULL) has its bit clear. */
int wx_col_stats(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity,
                 int64_t n, double* stats_out /* device [F][4] */, void* stream);

/* Chan merge of per-shard statistics (multi-GPU: all-gather the [F][4] blocks, then merge in rank
 * order).  parts: device [n_parts][F][4]; out: device [F][4], std_val recomputed.  SURVEY 8(e). */
int wx_stats_merge(wx_ctx* ctx, const double* parts, int n_parts, double* out, void* stream);

/* Standardise + pack: features_out[i][j] = (float)((double)col_j[i] - mean_j) / std_val_j), row-major
 * fp32 [n][F] -- exactly what WeatherDataset.features holds after preprocess.py:38-41 and the cast at
 * train_lightning_ddp.py:45.  Nulls: WX_IMPUTE_NONE -> NaN, WX_IMPUTE_MEAN -> 0.0f.
 * nullmask_out (may be NULL): uint32 per row, bit j set <=> feature j was null. */
int wx_normalize_pack(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity,
                      int64_t n, const double* stats /* device [F][4] */, int impute_mode,
                      float* features_out, uint32_t* nullmask_out, void* stream);

/* Columnar float64 z-score, the `*_norm` columns as preprocess.py:38-41 writes them to parquet:
 * out_cols[j][i] = ((double)col_j[i] - mean_j) / std_val_j with IEEE float64 subtraction and division (bit-identical
 * to the float64 expression Spark evaluates); a clear validity bit gives NaN (SQL null propagates, :38-41).
 * out_cols: HOST array of F device pointers to double[n]. */
int wx_zscore_f64(wx_ctx* ctx, const void* const* cols, int dtype, const uint8_t* const* validity,
                  int64_t n, const double* stats /* device [F][4] */, double* const* out_cols, void* stream);

/* Ascending flat indices (row*F + col) of every null cell recorded in nullmask -- the "imputation
 * indices" of the north star.  idx_out has room for `capacity` entries; *count_out (device) receives
 * the total number of null cells even when it exceeds capacity (then only the first `capacity` are
 * written).  No counterpart in the reference (it never imputes); defined by oracle/preprocess_ref.py. */
int wx_impute_index_list(wx_ctx* ctx, const uint32_t* nullmask, int64_t n, int64_t* idx_out,
                         int64_t capacity, int64_t* count_out, void* stream);

/* ---- K2/K3/K5: the training step (jobs/train_lightning_ddp.py) ------------------------------- */

/* Forward + cross-entropy + backward over one local batch, gradients only (no update).
 * Replaces WeatherClassifier.forward / training_step and autograd's backward (:57-71).
 *   x: fp32 row-major [*, F] (WeatherDataset.features, :45); y: int64 labels (:46).
 *   Rows of the batch: row_idx[0..batch) if row_idx != NULL (the sampler's indices), else
 *   row0, row0+1, ..., row0+batch-1.
 *   keep_mask: [batch] uint64 for WX_DROPOUT_MASK (bit j of entry b keeps unit j of batch row b).
 *   out[0..P) = scale * sum_rows d(loss_row)/d(param); out[P] = scale * sum_rows loss_row.
 * With scale = 1/batch this is the mean loss and its gradient; under DDP pass 1/(W*batch) and SUM
 * across ranks (what torch's reducer does at :136). */
int wx_mlp_grad(wx_ctx* ctx, const float* params, const float* x, const int64_t* y,
                const int64_t* row_idx, int64_t row0, int64_t batch, int dropout_mode,
                const uint64_t* keep_mask, uint64_t seed, uint64_t step, float scale,
                float* out /* device [P+1] */, void* stream);

/* wx_mlp_grad for callers that must not allocate device memory: no context (wx_create allocates), same
 * inputs, outputs and arithmetic as wx_mlp_grad, on the CURRENT device.  This entry point has no counterpart
 * in the reference; it exists for hosts that run under a framework-owned memory budget, inside latency-critical
 * loops that cannot take cudaMalloc's implicit synchronisation, or under a no-allocation-after-start policy.
 * The partial-sum scratch is a fixed 1.5 MiB array inside the library image (each GPU that loads the library
 * holds it until the process exits).  Limits: n_features <= 32 and batch <= 16,777,216 rows per call; larger
 * requests are rejected with WX_EINVAL and a message (split the batch and add the outputs -- they are sums).  The
 * launch never uses more thread blocks than the array holds partial-sum slabs for, so every in-limit request runs.
 * The first call on a device looks the array's device address up and queries the device; make that first call
 * before starting any stream capture.  The call itself performs no cudaMalloc / cudaFree / allocating library
 * call.  All calls share the one array: two calls whose kernels can overlap on the GPU (different streams,
 * threads or processes sharing a context) race on it -- order them on one stream. */
int wx_mlp_grad_static(int n_features, const float* params, const float* x, const int64_t* y,
                       const int64_t* row_idx, int64_t row0, int64_t batch, int dropout_mode,
                       const uint64_t* keep_mask, uint64_t seed, uint64_t step, float scale,
                       float* out /* device [P+1] */, void* stream);

/* Tensor-core (tcgen05, kind::tf32 with hi/lo operand splits -> fp32-level accuracy) versions of wx_mlp_grad and
 * wx_train_step: same outputs and the same 1e-5 contract (train_lightning_ddp.py:57-71,:88); the two F x 64 contractions
 * run on the 5th-generation tensor cores with operands and accumulators in tensor memory (csrc/tc_step_kernel.cuh).
 * Restrictions: n_features <= 31, contiguous rows only (no row_idx), x + row0*F*4 (or every column + row0*4) and
 * y + row0*8 16-byte aligned.  After wx_peer_connect the train-step forms sum the [P+1] bucket over all ranks inside
 * the kernel.  wx_tc_last_status reads a sticky protocol error code (0 = none): a bounded barrier wait that expired, or
 * 20 = a peer did not arrive in the exchange (that step was not applied on this rank). */
int wx_tc_mlp_grad(wx_ctx* ctx, const float* params, const float* x, const int64_t* y, int64_t row0, int64_t batch,
                   int dropout_mode, const uint64_t* keep_mask, uint64_t seed, uint64_t step, float scale,
                   float* out /* device [P+1] */, void* stream);
int wx_tc_train_step(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const float* x, const int64_t* y,
                     int64_t row0, int64_t batch, int dropout_mode, const uint64_t* keep_mask, uint64_t seed,
                     float* loss_out, void* stream);
/* Same step from RAW fp32 columns (standardisation fused in front, preprocess.py:38-41): the tensor-core counterpart
 * of wx_train_step_raw.  Columns must be 16-byte aligned at row0 (row0 % 4 == 0 for 16-byte-aligned column bases). */
int wx_tc_train_step_raw(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const void* const* cols,
                         const double* stats, int impute_mode, const int64_t* y, int64_t row0, int64_t batch,
                         int dropout_mode, const uint64_t* keep_mask, uint64_t seed, float* loss_out, void* stream);
int wx_tc_last_status(wx_ctx* ctx, int* host_out);

/* Adam update of a flat fp32 vector, torch.optim.Adam semantics with the desc's hyper-parameters
 * (:88; torch/optim/adam.py _single_tensor_adam).  step is 1-based.  grads may hold extra trailing
 * values (they are ignored).  Also the stand-alone HBM-bound K3 for long vectors (28 B/param). */
int wx_adam_step(wx_ctx* ctx, float* params, const float* grads, float* m, float* v, int64_t n,
                 int64_t step, void* stream);

/* One optimizer step in ONE launch: wx_mlp_grad + grid reduction + Adam in the kernel's tail.
 * state: device int64[2] = { step counter (number of completed steps), reserved }; the tail
 * increments it, so the call can be replayed from a CUDA graph.  loss_out: device fp32, receives
 * the batch-mean loss (`train_loss`, :70).  Single-GPU form of :66-71 + :88. */
int wx_train_step(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state, const float* x,
                  const int64_t* y, const int64_t* row_idx, int64_t row0, int64_t batch,
                  int dropout_mode, const uint64_t* keep_mask, uint64_t seed, float* loss_out,
                  void* stream);

/* Same step, but reading RAW columnar features and standardising in registers on the way in
 * (preprocess.py:38-41 fused in front of train_lightning_ddp.py:66-71,:88): rows row0..row0+batch of
 * F columns of `dtype`, stats from wx_col_stats.  Nulls (NaN cells) follow impute_mode. */
int wx_train_step_raw(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state,
                      const void* const* cols, int dtype, const double* stats, int impute_mode,
                      const int64_t* y, int64_t row0, int64_t batch, int dropout_mode,
                      const uint64_t* keep_mask, uint64_t seed, float* loss_out, void* stream);

/* Many small optimizer steps in one launch (the reference's regime: batch 4, :122): resident CTAs run
 * steps s = 0..n_steps-1 over rows row_idx[s*batch .. (s+1)*batch) (or row0 + s*batch ...),
 * forward/backward/Adam each, writing loss_out[s].  keep_mask is [n_steps*batch].
 * batch <= 64: one CTA.  64 < batch <= WX_RESIDENT_MAX_BATCH: up to 32 co-resident CTAs (cooperative launch) that keep
 * their own copy of params/m/v in shared memory, exchange per-CTA gradient sums through global memory and cross one grid
 * barrier per step (the latency-bound end of the batch sweep: the per-step launch, parameter reload and last-CTA tail of
 * wx_train_step are gone).  Same arithmetic and dropout stream as wx_train_step.  Measured at F = 20: 5 / 9.5 / 16.6 us per
 * step at batch 8 / 64 / 256 against ~20 us for one launch per step; from ~1024 rows on a launch per step is faster.
 * After wx_peer_connect (batch <= 64 per rank) every step is a DDP step (:136): the [P+1] sums cross NVLink inside the
 * step loop, every rank must launch the same n_steps and batch; if a peer stops arriving the launch ends early with
 * state[0] counting the steps actually applied and status 20 (wx_tc_last_status). */
#define WX_RESIDENT_MAX_BATCH 16384
int wx_train_steps_small(wx_ctx* ctx, float* params, float* m, float* v, int64_t* state,
                         const float* x, const int64_t* y, const int64_t* row_idx, int64_t row0,
                         int64_t batch, int64_t n_steps, int dropout_mode, const uint64_t* keep_mask,
                         uint64_t seed, float* loss_out /* [n_steps] */, void* stream);

/* Validation pass (validation_step, :73-85): eval-mode forward over the given rows.
 * out: device double[2] = { sum of per-row cross-entropy, number of rows with argmax == label }. */
int wx_mlp_eval(wx_ctx* ctx, const float* params, const float* x, const int64_t* y,
                const int64_t* row_idx, int64_t row0, int64_t batch, double* out, void* stream);

/* Inference forward (score.py run(), dags/azure_auto_deploy.py:99-103): probs_out[i][c] = softmax. */
int wx_mlp_predict(wx_ctx* ctx, const float* params, const float* x, int64_t n, float* probs_out,
                   void* stream);

/* Eval-mode forward() itself (train_lightning_ddp.py:64 with dropout off): logits_out[i][c] = the two raw logits,
 * what WeatherClassifier.forward returns to callers that read logits rather than probabilities. */
int wx_mlp_logits(wx_ctx* ctx, const float* params, const float* x, int64_t n, float* logits_out,
                  void* stream);

/* ---- wide tabular MLP on bf16 tcgen05 GEMMs (BASELINE.json configs[3]; SURVEY 8(b) "wx_wide_mlp_fwd/bwd") ----------
 * No counterpart in the reference, whose only model is Linear(F,64)-ReLU-Dropout-Linear(64,2)
 * (jobs/train_lightning_ddp.py:57-62).  Same layer pattern widened: n_hidden x [Linear, ReLU, Dropout(p)] then
 * Linear(hidden, 2), CE loss (:69), Adam (:88).  Parameters are ONE flat fp32 vector in torch module order
 * (W_1 [H,F], b_1 [H], W_2 [H,H], b_2 [H], ..., W_o [2,H], b_o [2]); gradients use the same layout plus one extra
 * element [P] for the loss, so the DDP exchange (:136) is a SUM all-reduce of (slices of) one buffer.
 * Hidden-layer products run in bf16 with fp32 accumulation (activations and dZ are stored as bf16); the head, the loss,
 * bias gradients and Adam are fp32.  All buffers are caller-owned device memory; `ws` is a caller-allocated workspace of
 * wx_wide_workspace_bytes() bytes (1024-byte aligned) that carries the bf16 weight copies between calls.
 * Constraints: hidden % 256 == 0, n_features <= 64, classes == 2, batch % 128 == 0. */
typedef struct wx_wide_desc {
    int32_t n_features; /* F */
    int32_t hidden;     /* H */
    int32_t n_hidden;   /* number of hidden layers; "4 layers, hidden 4096" = 3 hidden + the head */
    int32_t classes;    /* 2 */
    float dropout_p;
    double lr, beta1, beta2, eps;
} wx_wide_desc;

int64_t wx_wide_n_params(const wx_wide_desc* desc);                      /* P, or -1 (see wx_last_error) */
int64_t wx_wide_workspace_bytes(const wx_wide_desc* desc, int64_t batch); /* bytes, or -1 */
/* (Re)build the bf16 copies of the hidden weights inside ws from the fp32 parameters (once after init / load). */
int wx_wide_cast_weights(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, void* stream);
/* Forward + loss + head gradients: x fp32 [batch,F] row-major standardised features (wx_normalize_pack output),
 * y int64 [batch].  grads fp32 [P+1] receives db_o and, at [P], sum(CE) * grad_scale (dW_o: WX_WIDE_HEAD_DW of
 * wx_wide_backward_layer, so that H_L is read once on the critical path -- logits and dZ_L come out of one pass); grad_scale is normally
 * 1/(world*batch).  dropout_mode: WX_DROPOUT_OFF or WX_DROPOUT_PHILOX (keep <=> 16 Philox4x32-10 bits <
 * round((1-p)*65536); counter (row, col/8, layer, step), key = seed). */
int wx_wide_forward(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, const float* x,
                    const int64_t* y, int64_t batch, int dropout_mode, uint64_t seed, uint64_t step, float grad_scale,
                    float* grads, void* stream);
/* Backward of hidden layer `layer` = n_hidden, ..., 1 (in that order, after wx_wide_forward): dW_layer and db_layer into
 * grads; the caller may start reducing that slice across ranks while the next layer's backward runs.  `parts` selects
 * which of the three pieces this call enqueues (WX_WIDE_ALL for everything), so db can go to a side stream: it only
 * reads dZ of this layer, which stays valid until the dZ product of the layer BELOW the next one overwrites it. */
#define WX_WIDE_DW 1 /* weight gradient of the layer */
#define WX_WIDE_DB 2 /* bias gradient (column sums of dZ) */
#define WX_WIDE_DZ 4 /* dZ of the layer below (skipped for layer 1) */
#define WX_WIDE_DW_FOLD 8 /* add the workspace slabs of dW's split tail tiles into the gradient (no-op when dW was not split); must follow DW */
#define WX_WIDE_HEAD_DW 16 /* with layer == n_hidden only: the head's weight gradient dW_o (wx_wide_forward leaves it to this piece) */
#define WX_WIDE_DB_FUSED 32 /* layer < n_hidden only: the bias gradient from the 32-row column sums that the WX_WIDE_DZ piece of
                             * layer + 1 wrote beside dZ_layer (no pass over dZ); use instead of WX_WIDE_DB, after that piece */
#define WX_WIDE_ALL 31
int wx_wide_backward_layer(wx_ctx* ctx, const wx_wide_desc* desc, void* ws, int64_t batch, int layer, int dropout_mode,
                           int parts, float* grads, void* stream);
/* Adam over all P parameters (step is 1-based), rewriting the bf16 weight copies in ws in the same pass. */
int wx_wide_adam(wx_ctx* ctx, const wx_wide_desc* desc, float* params, const float* grads, float* m, float* v,
                 int64_t step, void* ws, void* stream);
/* The same for one layer's tensors (layer 1..n_hidden, n_hidden+1 = head), e.g. on a side stream once that slice is final. */
int wx_wide_adam_layer(wx_ctx* ctx, const wx_wide_desc* desc, float* params, const float* grads, float* m, float* v,
                       int64_t step, void* ws, int layer, void* stream);
/* Keep `n` SMs (even, 0..64; default 0) out of the wide GEMM grids, for a collective that runs concurrently on another stream. */
int wx_wide_reserve_sms(wx_ctx* ctx, int n);
/* Validation pass of the wide net (the :73-85 pattern): eval-mode forward, out[0] = sum of per-row CE, out[1] = number of
 * rows whose argmax equals the label (fp32 device scalars; first index wins a tie, as torch.argmax). */
int wx_wide_eval(wx_ctx* ctx, const wx_wide_desc* desc, const float* params, void* ws, const float* x, const int64_t* y,
                 int64_t batch, float* out, void* stream);
/* The GEMM underneath, exposed for tests: out fp32 [M,N] = op(A) op(B) (+ bias[N]); A is [M,K] (a_kmajor) or [K,M],
 * B is [N,K] (b_kmajor) or [K,N], row-major bf16 with pitches lda/ldb in elements.  N % 16 == 0, K % 64 == 0.
 * Supported layout pairs: (1,1), (1,0), (0,0) -- the three the wide MLP uses. */
int wx_wide_gemm(wx_ctx* ctx, int a_kmajor, int b_kmajor, const void* A, int64_t lda, const void* B, int64_t ldb,
                 int64_t M, int64_t N, int64_t K, const float* bias, float* out, void* stream);


#ifdef __cplusplus
}
#endif
#endif /* WXDDP_H_ */
